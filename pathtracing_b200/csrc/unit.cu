// unit.cu — deterministic probes of the DEVICE build of the shading code (shade*.cuh as render.cu compiles it:
// sm_100a, -use_fast_math), one thread per record.  The reference has per-function entry points only implicitly
// (Material::scatter / PDF / calc_attenuation, Light::sample / PDF / L, HenyeyGreenstein::Sample, the interaction that
// Shape::Intersect builds); tests/test_shading_gpu.py feeds the same seeded inputs to these kernels and to the
// reference's own functions (oracle/ref_driver.cpp ref_unit_*) and compares field by field with stated tolerances.
// Record layouts are those of ref_unit_* (and of the host emulation tests/emul/emul.cpp, which checks the IEEE host
// build of the same headers in the CPU tier).  Diagnostic entry points: synchronous, host buffers, no state kept.
#include <cuda_runtime.h>
#include <math_constants.h>

#include "scene_host.hpp"
#include "shade_mat.cuh"
#include "shade_light.cuh"
#include "trace.cuh"
#include "alpha.cuh"

namespace ptb {

template <int WIDTH>
__global__ void __launch_bounds__(128) k_unit_shade(DevScene D, ShadeScene S, const float4* __restrict__ rays, uint32_t n, const float* __restrict__ in,
                                                    float* __restrict__ out, int stride){
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if(i >= n)return;
    float* o = out + (size_t)i * stride;
    for(int k = 0; k < stride; k++)o[k] = 0.0f;
    HitRec h;
    if(!trace_scene<WIDTH, false, false>(D, rays, i, h, nullptr, NoAlphaWrap()))return;
    const float4 ra = rays[2 * (size_t)i], rb = rays[2 * (size_t)i + 1];
    const V3 ro = v3(ra.x, ra.y, ra.z), rd = v3(rb.x, rb.y, rb.z);
    SurfInt si;
    build_interaction(S, ro, rd, rb.w, h.t, h.u, h.v, h.prim, h.inst, si);
    o[0] = 1.0f;
    const V3 vs[4] = { si.p, si.n, si.ns, si.tangent };
    for(int k = 0; k < 4; k++){ o[1 + 3 * k] = vs[k].x; o[2 + 3 * k] = vs[k].y; o[3 + 3 * k] = vs[k].z; }
    o[13] = si.u; o[14] = si.v; o[15] = si.t; o[16] = (float)si.mat; o[17] = si.light >= 0 ? 1.0f : 0.0f;
    const float* q = in + 6 * (size_t)i;
    if(si.mat >= 0){
        MatCtx mc;
        mat_fetch(S, si, mc);
        Scatter sc;
        if(mat_scatter(mc, si, ro, rd, q[0], q[1], q[2], sc)){
            o[18] = 1.0f;
            o[19] = sc.f.x; o[20] = sc.f.y; o[21] = sc.f.z; o[22] = sc.pdf; o[23] = (float)sc.flags;
            o[24] = sc.o.x; o[25] = sc.o.y; o[26] = sc.o.z; o[27] = sc.d.x; o[28] = sc.d.y; o[29] = sc.d.z;
            const V3 a = mat_eval(mc, si, rd, sc.d);
            o[30] = a.x; o[31] = a.y; o[32] = a.z; o[33] = mat_pdf(mc, si, rd, sc.d);
        }
        const V3 w2 = normalize(v3(q[3], q[4], q[5]));
        const V3 a2 = mat_eval(mc, si, rd, w2);
        o[34] = a2.x; o[35] = a2.y; o[36] = a2.z; o[37] = mat_pdf(mc, si, rd, w2);
    }
    if(si.light >= 0){
        const DevLight& L = S.lights[si.light];
        const V3 le = area_L(S, L, si.n, si.ns, si.u, si.v, rd);   // what k_shade evaluates on an emissive hit
        o[38] = le.x; o[39] = le.y; o[40] = le.z; o[41] = L.pmf_hit; o[42] = area_pdf(L, si.p, si.n, ro, rd);
    }
}

__global__ void __launch_bounds__(128) k_unit_light_sample(ShadeScene S, int k, uint32_t n, const float* __restrict__ in, float* __restrict__ out, int stride, float time){
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if(i >= n)return;
    const DevLight& L = S.lights[k];
    Pcg rng = pcg_seed(1, 1 + i);
    const float* q = in + 5 * (size_t)i;
    float* o = out + (size_t)i * stride;
    LightSmp ls;
    light_sample(S, L, q[0], q[1], time, rng, ls);
    const V3 ip = v3(q[2], q[3], q[4]);
    V3 ldir; float t;
    if(is_zero(ls.n)){ ldir = ls.dir; t = CUDART_INF_F; }
    else{ ldir = ls.p - ip; t = length(ldir) - kEps; }
    const V3 sd = normalize(ldir);
    o[0] = ls.L.x; o[1] = ls.L.y; o[2] = ls.L.z; o[3] = ls.p.x; o[4] = ls.p.y; o[5] = ls.p.z; o[6] = ls.n.x; o[7] = ls.n.y; o[8] = ls.n.z;
    o[9] = ls.u; o[10] = ls.v; o[11] = ls.dir.x; o[12] = ls.dir.y; o[13] = ls.dir.z; o[14] = sd.x; o[15] = sd.y; o[16] = sd.z; o[17] = t;
    o[18] = light_pdf_sampled(S, L, ls.p, ls.n, ip, sd, time);
    const V3 Lv = light_L_sampled(S, L, ls, sd);
    o[19] = Lv.x; o[20] = Lv.y; o[21] = Lv.z;
}

__global__ void __launch_bounds__(128) k_unit_hg(float g, uint32_t n, const float* __restrict__ in, float* __restrict__ out){
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if(i >= n)return;
    const float gc = fminf(fmaxf(g, -0.99f), 0.99f);        // HenyeyGreenstein ctor clamp (PhaseFunction.hpp:20)
    const float* q = in + 5 * (size_t)i;
    const V3 wi = normalize(v3(q[0], q[1], q[2]));
    const V3 wo = hg_sample(gc, wi, q[3], q[4]);
    const float pdf = phase_hg(dot(wi, wo), gc);
    out[5 * (size_t)i] = wo.x; out[5 * (size_t)i + 1] = wo.y; out[5 * (size_t)i + 2] = wo.z; out[5 * (size_t)i + 3] = pdf; out[5 * (size_t)i + 4] = pdf;
}

namespace {
struct Tmp {       // device scratch of one probe call
    void* p = nullptr;
    ~Tmp(){ if(p)cudaFree(p); }
    bool up(const void* src, size_t bytes, const char* what){
        return cuda_ok(cudaMalloc(&p, bytes ? bytes : 16), what) && (bytes == 0 || cuda_ok(cudaMemcpy(p, src, bytes, cudaMemcpyHostToDevice), what));
    }
    bool make(size_t bytes, const char* what){ return cuda_ok(cudaMalloc(&p, bytes ? bytes : 16), what); }
};
}

} // namespace ptb

using namespace ptb;

extern "C" {

int pt_unit_shade(pt_scene* s, const pt_ray* rays, size_t n, const float* in6, float* out, int stride){
    if(!s || !s->committed){ set_error("scene not committed"); return PT_ERR_STATE; }
    if(stride < 43 || (n && (!rays || !in6 || !out))){ set_error("pt_unit_shade: stride >= 43 and non-null buffers"); return PT_ERR_INVALID; }
    if(n == 0)return PT_OK;
    if(!cuda_ok(cudaSetDevice(s->device), "cudaSetDevice"))return PT_ERR_CUDA;
    Tmp dr, di, dout;
    if(!dr.up(rays, n * sizeof(pt_ray), "rays") || !di.up(in6, n * 6 * sizeof(float), "inputs") || !dout.make(n * (size_t)stride * sizeof(float), "outputs"))return PT_ERR_CUDA;
    const unsigned grid = (unsigned)((n + 127) / 128);
    if(s->dev.fast)k_unit_shade<kWidthFast4><<<grid, 128>>>(s->dev, s->shade, (const float4*)dr.p, (uint32_t)n, (const float*)di.p, (float*)dout.p, stride);
    else switch(s->dev.width){
    case 2: k_unit_shade<2><<<grid, 128>>>(s->dev, s->shade, (const float4*)dr.p, (uint32_t)n, (const float*)di.p, (float*)dout.p, stride); break;
    case 4: k_unit_shade<4><<<grid, 128>>>(s->dev, s->shade, (const float4*)dr.p, (uint32_t)n, (const float*)di.p, (float*)dout.p, stride); break;
    default: k_unit_shade<8><<<grid, 128>>>(s->dev, s->shade, (const float4*)dr.p, (uint32_t)n, (const float*)di.p, (float*)dout.p, stride); break;
    }
    if(!cuda_ok(cudaGetLastError(), "k_unit_shade") || !cuda_ok(cudaMemcpy(out, dout.p, n * (size_t)stride * sizeof(float), cudaMemcpyDeviceToHost), "outputs"))return PT_ERR_CUDA;
    return PT_OK;
}

int pt_unit_light_count(pt_scene* s){
    if(!s || !s->committed){ set_error("scene not committed"); return PT_ERR_STATE; }
    return s->shade.n_lights;
}

int pt_unit_light_sample(pt_scene* s, int k, size_t n, const float* in5, float* out, int stride, float time){
    if(!s || !s->committed){ set_error("scene not committed"); return PT_ERR_STATE; }
    if(k < 0 || k >= s->shade.n_lights || stride < 22 || (n && (!in5 || !out))){ set_error("pt_unit_light_sample: bad light index / stride / buffers"); return PT_ERR_INVALID; }
    if(n == 0)return PT_OK;
    if(!cuda_ok(cudaSetDevice(s->device), "cudaSetDevice"))return PT_ERR_CUDA;
    Tmp di, dout;
    if(!di.up(in5, n * 5 * sizeof(float), "inputs") || !dout.make(n * (size_t)stride * sizeof(float), "outputs"))return PT_ERR_CUDA;
    k_unit_light_sample<<<(unsigned)((n + 127) / 128), 128>>>(s->shade, k, (uint32_t)n, (const float*)di.p, (float*)dout.p, stride, time);
    if(!cuda_ok(cudaGetLastError(), "k_unit_light_sample") || !cuda_ok(cudaMemcpy(out, dout.p, n * (size_t)stride * sizeof(float), cudaMemcpyDeviceToHost), "outputs"))return PT_ERR_CUDA;
    return PT_OK;
}

int pt_unit_hg(int device, float g, size_t n, const float* in5, float* out5){
    if(n && (!in5 || !out5)){ set_error("null buffer"); return PT_ERR_INVALID; }
    if(n == 0)return PT_OK;
    if(!cuda_ok(cudaSetDevice(device), "cudaSetDevice"))return PT_ERR_CUDA;
    Tmp di, dout;
    if(!di.up(in5, n * 5 * sizeof(float), "inputs") || !dout.make(n * 5 * sizeof(float), "outputs"))return PT_ERR_CUDA;
    k_unit_hg<<<(unsigned)((n + 127) / 128), 128>>>(g, (uint32_t)n, (const float*)di.p, (float*)dout.p);
    if(!cuda_ok(cudaGetLastError(), "k_unit_hg") || !cuda_ok(cudaMemcpy(out5, dout.p, n * 5 * sizeof(float), cudaMemcpyDeviceToHost), "outputs"))return PT_ERR_CUDA;
    return PT_OK;
}

int pt_unit_table(pt_scene* s, int table, void* dst, size_t cap_bytes, size_t* bytes){
    if(!s || !s->committed || !bytes){ set_error("pt_unit_table: needs a committed scene"); return PT_ERR_INVALID; }
    if(!cuda_ok(cudaSetDevice(s->device), "cudaSetDevice"))return PT_ERR_CUDA;
    const DeviceBuffers& b = s->buf;
    const bool four = s->dev.width == 4;
    const void* src = nullptr;
    size_t n = 0;
    switch(table){
        case 0: src = four ? b.n4 : s->dev.width == 2 ? b.n2 : b.n8; n = s->total_nodes * (four ? sizeof(Bvh4Node) : s->dev.width == 2 ? sizeof(Bvh2Node) : sizeof(Bvh8Node)); break;
        case 1: src = b.tris; n = s->total_tris * sizeof(Tri); break;
        case 2: src = b.n4q; n = s->dev.fast ? s->total_nodes * sizeof(Bvh4QNode) : 0; break;
        case 3: src = b.tris64; n = s->dev.fast ? s->total_tris * sizeof(Tri64) : 0; break;
        case 4: src = b.inst; n = s->n_inst_records * sizeof(Instance); break;
        // shading geometry: positions, normals, texcoords, tangents, indices, shading triangles, emissive ordinals
        case 5: src = b.vtx_pos; n = s->n_shade_vertices * 12; break;
        case 6: src = b.vtx_nrm; n = s->n_shade_vertices * 12; break;
        case 7: src = b.vtx_uv; n = s->n_shade_vertices * 8; break;
        case 8: src = b.vtx_tan; n = s->n_shade_vertices * 12; break;
        case 9: src = b.indices; n = s->n_shade_indices * 4; break;
        case 10: src = b.tri_mesh; n = s->n_shade_tris * sizeof(ShadeTri); break;
        case 11: src = b.tri_light; n = s->n_shade_tris * 4; break;
        default: set_error("pt_unit_table: unknown table"); return PT_ERR_INVALID;
    }
    *bytes = n;
    if(n == 0 || !dst || cap_bytes < n)return PT_OK;
    return cuda_ok(cudaMemcpy(dst, src, n, cudaMemcpyDeviceToHost), "table download") ? PT_OK : PT_ERR_CUDA;
}

} // extern "C"
