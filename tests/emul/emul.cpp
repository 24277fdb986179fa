// tests/emul/emul.cpp — TEST INFRASTRUCTURE ONLY.
//
// Compiles the product's device traversal header (pathtracing_b200/csrc/trace.cuh)
// for the HOST with g++ by shimming the handful of CUDA intrinsics it uses, and
// runs it over the product's own packed scene (scene_build.cpp + bvh_build.cpp).
// This lets the CPU-only test tier (-m "not gpu") check the host BVH build, the
// flattening, the instance records and the traversal logic against the
// reference before any GPU time is spent.  It is never linked into
// libptb200.so and is not a fallback: the product has no host traversal.
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>
#include <cuda_runtime.h>

#ifndef __CUDACC__
static inline float __fsub_rn(float a, float b){ return a - b; }
static inline float __fadd_rn(float a, float b){ return a + b; }
static inline float __fmul_rn(float a, float b){ return a * b; }
static inline float __fdiv_rn(float a, float b){ return a / b; }
static inline unsigned __float_as_uint(float f){ unsigned u; memcpy(&u, &f, 4); return u; }
static inline float __uint_as_float(unsigned u){ float f; memcpy(&f, &u, 4); return f; }
static inline float __int_as_float(unsigned u){ float f; memcpy(&f, &u, 4); return f; }
template <class T> static inline T __ldg(const T* p){ return *p; }
static inline float __uint2float_rn(unsigned u){ return (float)u; }
#endif

#include "../../pathtracing_b200/csrc/scene_host.hpp"
#include "../../pathtracing_b200/csrc/trace.cuh"
#include "../../pathtracing_b200/csrc/shade_tables.hpp"
#include "../../pathtracing_b200/csrc/shade_mat.cuh"
#include "../../pathtracing_b200/csrc/shade_light.cuh"
#include "../../pathtracing_b200/csrc/camera.cuh"

using namespace ptb;

struct EmulScene {
    pt_scene s;
    PackedTraversal P;
    std::vector<Bvh4QNode> n4q;      // width 44: the fast layout (PT_LAYOUT_FAST)
    std::vector<Tri64> tris64;
    DevScene D;
    ShadeTables T;          // host copies of the shading tables (emul_shade_tables)
    bool shaded = false;
    unsigned long long overflows = 0;
};

extern "C" {

PT_API void* emul_scene_create(){ return new EmulScene(); }
PT_API void emul_scene_destroy(void* p){ delete (EmulScene*)p; }
PT_API const char* emul_last_error(){ return get_error(); }
PT_API int emul_scene_add_texture(void* p, const pt_texture_desc* d){ return scene_add_texture(&((EmulScene*)p)->s, d); }
PT_API int emul_scene_add_material(void* p, const pt_material_desc* d){ return scene_add_material(&((EmulScene*)p)->s, d); }
PT_API int emul_scene_add_medium(void* p, const pt_medium_desc* d){ return scene_add_medium(&((EmulScene*)p)->s, d); }
PT_API int emul_scene_add_mesh(void* p, const pt_mesh_desc* d){ return scene_add_mesh(&((EmulScene*)p)->s, d); }
PT_API int emul_scene_add_blas(void* p, const int32_t* ids, int n){ return scene_add_blas(&((EmulScene*)p)->s, ids, n); }
PT_API int emul_scene_add_instance(void* p, int b, const float* xf){ return scene_add_instance(&((EmulScene*)p)->s, b, xf); }
PT_API int emul_scene_add_shape(void* p, const pt_shape_desc* d, const float* xf){ return scene_add_shape(&((EmulScene*)p)->s, d, xf); }
PT_API int emul_scene_animate(void* p, int id, const float* dir, const float* tb){ return scene_animate(&((EmulScene*)p)->s, id, dir, tb); }
PT_API void emul_mat4_inverse(const float* m, float* out){ mat4_inverse(m, out); }
PT_API int emul_scene_add_light(void* p, const pt_light_desc* d){ return scene_add_light(&((EmulScene*)p)->s, d); }

PT_API int emul_scene_commit(void* p, int width){
    EmulScene* e = (EmulScene*)p;
    const bool fast = width == kWidthFast4;
    if(fast)width = 4;
    int rc = pack_traversal(&e->s, width, e->P);
    if(rc != PT_OK)return rc;
    e->D.fast = fast ? 1 : 0; e->D.n4q = nullptr; e->D.tris64 = nullptr;
    if(fast){
        encode_bvh4q(e->P.n4, e->n4q);
        e->tris64.resize(e->P.tris.size());
        for(size_t i = 0; i < e->P.tris.size(); i++){ memset(&e->tris64[i], 0, sizeof(Tri64)); e->tris64[i].t = e->P.tris[i]; }
        e->D.n4q = e->n4q.data(); e->D.tris64 = e->tris64.data();
    }
    e->D.n2 = e->P.n2.data(); e->D.n4 = e->P.n4.data(); e->D.n8 = e->P.n8.data(); e->D.tris = e->P.tris.data(); e->D.inst = e->P.insts.data();
    e->D.tlas_root = e->P.tlas_root; e->D.width = width;
    e->D.single_root = e->P.single_root; e->D.single_id = e->P.single_id;
    for(int a = 0; a < 3; a++){ e->D.tlo[a] = e->P.bounds.lo[a]; e->D.thi[a] = e->P.bounds.hi[a]; }
    return PT_OK;
}

// fast layout: every decoded child box contains the exact one; returns the number of violations, the clusters
// checked and the largest growth of a child box relative to its cluster's extent (per side, any axis)
PT_API int emul_bvh4q_check(void* p, int* n_nodes, float* worst_growth){
    EmulScene* e = (EmulScene*)p;
    int bad = 0;
    float worst = 0.0f;
    for(size_t i = 0; i < e->n4q.size(); i++){
        const Bvh4Node& a = e->P.n4[i];
        const Bvh4QNode& q = e->n4q[i];
        const float* mins[3] = { a.xmin, a.ymin, a.zmin };
        const float* maxs[3] = { a.xmax, a.ymax, a.zmax };
        for(int ax = 0; ax < 3; ax++){
            const float step = __uint_as_float((unsigned)q.ex[ax] << 23);
            float lo = INFINITY, hi = -INFINITY;
            for(int c = 0; c < 4; c++)if(a.ref[c] != REF_EMPTY){ lo = std::min(lo, mins[ax][c]); hi = std::max(hi, maxs[ax][c]); }
            for(int c = 0; c < 4; c++){
                if(a.ref[c] == REF_EMPTY)continue;
                const float dl = q.org[ax] + (float)((q.q[ax] >> (8 * c)) & 255u) * step, dh = q.org[ax] + (float)((q.q[3 + ax] >> (8 * c)) & 255u) * step;
                if(!(dl <= mins[ax][c]) || !(dh >= maxs[ax][c]) || !(mins[ax][c] - dl >= 0.2f * step) || !(dh - maxs[ax][c] >= 0.2f * step))bad++;
                if(hi > lo){
                    const float ext = std::max(hi - lo, 1e-30f);
                    if(step <= ext){ worst = std::max(worst, (mins[ax][c] - dl) / ext); worst = std::max(worst, (dh - maxs[ax][c]) / ext); }
                }
            }
        }
    }
    *n_nodes = (int)e->n4q.size(); *worst_growth = worst;
    return bad;
}

// flatten_bvh on `threads` threads against the one-thread recursion, over n random item boxes: number of differing
// bytes in the node arrays + differing entries of order / leaf_last / root (0 = identical)
PT_API long long emul_flatten_check(unsigned n, unsigned seed, int width, unsigned threads, int clustered, int* n_nodes){
    std::vector<Box> boxes(n);
    uint64_t st = seed * 0x9E3779B97F4A7C15ull + 1;
    auto rnd = [&]{ st = st * 6364136223846793005ull + 1442695040888963407ull; return (float)((st >> 40) & 0xFFFFFF) / 16777216.0f; };
    for(unsigned i = 0; i < n; i++){
        float c[3] = { rnd(), rnd(), rnd() };
        if(clustered && (i & 3))for(int k = 0; k < 3; k++)c[k] = 0.5f + 0.01f * c[k];     // a dense clump: a lopsided tree
        Box b = box_empty();
        for(int v = 0; v < 3; v++){ float p[3]; for(int k = 0; k < 3; k++)p[k] = c[k] + 0.01f * (rnd() - 0.5f); box_grow(b, p); }
        boxes[i] = b;
    }
    const BinaryBvh bb = build_binary_bvh(boxes.data(), n);
    const FlatBvh a = flatten_bvh(bb, width, 1), t = flatten_bvh(bb, width, threads);
    long long bad = 0;
    auto cmp = [&](const void* x, size_t nx, const void* y, size_t ny){
        if(nx != ny){ bad += 1 + (long long)(nx > ny ? nx - ny : ny - nx); return; }
        const unsigned char* p = (const unsigned char*)x; const unsigned char* q = (const unsigned char*)y;
        for(size_t i = 0; i < nx; i++)bad += p[i] != q[i];
    };
    cmp(a.n2.data(), a.n2.size() * sizeof(Bvh2Node), t.n2.data(), t.n2.size() * sizeof(Bvh2Node));
    cmp(a.n4.data(), a.n4.size() * sizeof(Bvh4Node), t.n4.data(), t.n4.size() * sizeof(Bvh4Node));
    cmp(a.n8.data(), a.n8.size() * sizeof(Bvh8Node), t.n8.data(), t.n8.size() * sizeof(Bvh8Node));
    cmp(a.order.data(), a.order.size() * 4, t.order.data(), t.order.size() * 4);
    cmp(a.leaf_last.data(), a.leaf_last.size() * 4, t.leaf_last.data(), t.leaf_last.size() * 4);
    bad += a.root != t.root;
    *n_nodes = (int)(a.n2.size() + a.n4.size() + a.n8.size());
    return bad;
}

// rays of the last emul_trace_closest whose traversal needed more than the 96 stack rows (the product counts them and fails the call)
PT_API unsigned long long emul_stack_overflows(void* p){ return ((EmulScene*)p)->overflows; }

PT_API int emul_scene_bounds(void* p, float lo[3], float hi[3]){
    EmulScene* e = (EmulScene*)p;
    for(int a = 0; a < 3; a++){ lo[a] = e->P.bounds.lo[a]; hi[a] = e->P.bounds.hi[a]; }
    return PT_OK;
}

PT_API int emul_node_count(void* p){ EmulScene* e = (EmulScene*)p; return (int)(e->D.width == 2 ? e->P.n2.size() : e->D.width == 4 ? e->P.n4.size() : e->P.n8.size()); }

PT_API int emul_trace_closest(void* p, const pt_ray* rays, size_t n, pt_hit* hits, uint32_t* inst, unsigned long long* stats){
    EmulScene* e = (EmulScene*)p;
    const DevScene& D = e->D;
    int nt = (int)std::thread::hardware_concurrency(); if(nt < 1)nt = 1;
    std::vector<TraceCounters> cs(nt, TraceCounters{0, 0, 0, 0});
    std::vector<unsigned long long> over(nt, 0);
    auto work = [&](int tid){
        for(size_t i = tid; i < n; i += nt){
            HitRec h;
            const float4* rb = reinterpret_cast<const float4*>(rays);
            if(D.fast)trace_scene<kWidthFast4, false, true>(D, rb, i, h, &cs[tid], NoAlphaWrap());
            else if(D.width == 2)trace_scene<2, false, true>(D, rb, i, h, &cs[tid], NoAlphaWrap());
            else if(D.width == 4)trace_scene<4, false, true>(D, rb, i, h, &cs[tid], NoAlphaWrap());
            else trace_scene<8, false, true>(D, rb, i, h, &cs[tid], NoAlphaWrap());
            hits[i].t = h.t; hits[i].u = h.u; hits[i].v = h.v; hits[i].prim = h.prim;
            if(inst)inst[i] = h.inst;
            over[tid] += h.overflow;
        }
    };
    std::vector<std::thread> pool;
    for(int t = 0; t < nt; t++)pool.emplace_back(work, t);
    for(auto& t : pool)t.join();
    if(stats){
        stats[0] = stats[1] = stats[2] = stats[3] = 0;
        for(auto& c : cs){ stats[0] += c.tlas_nodes; stats[1] += c.blas_nodes; stats[2] += c.instances; stats[3] += c.triangles; }
    }
    e->overflows = 0;
    for(unsigned long long o : over)e->overflows += o;
    return PT_OK;
}

PT_API int emul_trace_any(void* p, const pt_ray* rays, size_t n, uint8_t* occ){
    EmulScene* e = (EmulScene*)p;
    const DevScene& D = e->D;
    int nt = (int)std::thread::hardware_concurrency(); if(nt < 1)nt = 1;
    auto work = [&](int tid){
        for(size_t i = tid; i < n; i += nt){
            HitRec h;
            const float4* rb = reinterpret_cast<const float4*>(rays);
            bool o = D.fast ? trace_scene<kWidthFast4, true, false>(D, rb, i, h, nullptr, NoAlphaWrap())
                   : D.width == 2 ? trace_scene<2, true, false>(D, rb, i, h, nullptr, NoAlphaWrap())
                   : D.width == 4 ? trace_scene<4, true, false>(D, rb, i, h, nullptr, NoAlphaWrap())
                                  : trace_scene<8, true, false>(D, rb, i, h, nullptr, NoAlphaWrap());
            occ[i] = o ? 1 : 0;
        }
    };
    std::vector<std::thread> pool;
    for(int t = 0; t < nt; t++)pool.emplace_back(work, t);
    for(auto& t : pool)t.join();
    return PT_OK;
}

} // extern "C"


// ---- deterministic unit probes of the product's shade*.cuh compiled for the host (IEEE math, no fast-math):
// the same record layouts as oracle/ref_driver.cpp ref_unit_* ------------------------------------------------
extern "C" {

PT_API int emul_shade_tables(void* p, int light_sampler){
    EmulScene* e = (EmulScene*)p;
    e->s.light_sampler = light_sampler;
    int rc = build_shade_tables_host(&e->s, e->P.inst_order, e->T);
    e->shaded = rc == PT_OK;
    return rc;
}

PT_API int emul_unit_shade(void* p, const pt_ray* rays, size_t n, const float* in, float* out, int stride){
    EmulScene* e = (EmulScene*)p;
    if(!e->shaded || stride < 43)return PT_ERR_STATE;
    const ShadeScene& S = e->T.view;
    const float4* rb = reinterpret_cast<const float4*>(rays);
    for(size_t i = 0; i < n; i++){
        float* o = out + i * (size_t)stride;
        for(int k = 0; k < stride; k++)o[k] = 0.0f;
        HitRec h;
        const bool hit = e->D.width == 2 ? trace_scene<2, false, false>(e->D, rb, i, h, nullptr, NoAlphaWrap())
                       : e->D.width == 4 ? trace_scene<4, false, false>(e->D, rb, i, h, nullptr, NoAlphaWrap())
                                         : trace_scene<8, false, false>(e->D, rb, i, h, nullptr, NoAlphaWrap());
        if(!hit)continue;
        const V3 ro = v3(rays[i].ox, rays[i].oy, rays[i].oz), rd = v3(rays[i].dx, rays[i].dy, rays[i].dz);
        SurfInt si;
        build_interaction(S, ro, rd, rays[i].time, h.t, h.u, h.v, h.prim, h.inst, si);
        o[0] = 1.0f;
        const V3 vs[4] = { si.p, si.n, si.ns, si.tangent };
        for(int k = 0; k < 4; k++){ o[1 + 3 * k] = vs[k].x; o[2 + 3 * k] = vs[k].y; o[3 + 3 * k] = vs[k].z; }
        o[13] = si.u; o[14] = si.v; o[15] = si.t; o[16] = (float)si.mat; o[17] = si.light >= 0 ? 1.0f : 0.0f;
        const float* q = in + 6 * i;
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
    return PT_OK;
}

PT_API int emul_unit_light_count(void* p){ return (int)((EmulScene*)p)->T.lights.size(); }

// per light: 0 power is not stored per light on the device (only the sampler's cdf): pmf_sample | 1 pmf_hit | 2 delta | 3 valid
PT_API int emul_unit_lights(void* p, float* out){
    EmulScene* e = (EmulScene*)p;
    for(size_t k = 0; k < e->T.lights.size(); k++){
        const DevLight& L = e->T.lights[k];
        out[4 * k] = L.pmf_sample; out[4 * k + 1] = L.pmf_hit; out[4 * k + 2] = light_is_delta(L) ? 1.0f : 0.0f; out[4 * k + 3] = (float)L.valid;
    }
    return PT_OK;
}

PT_API int emul_unit_light_sample(void* p, int k, size_t n, const float* in, float* out, int stride, float time){
    EmulScene* e = (EmulScene*)p;
    if(k < 0 || k >= (int)e->T.lights.size() || stride < 22)return PT_ERR_INVALID;
    const ShadeScene& S = e->T.view;
    const DevLight& L = S.lights[k];
    Pcg rng = pcg_seed(1, 1);
    for(size_t i = 0; i < n; i++){
        const float* q = in + 5 * i;
        float* o = out + i * (size_t)stride;
        LightSmp ls;
        light_sample(S, L, q[0], q[1], time, rng, ls);
        const V3 ip = v3(q[2], q[3], q[4]);
        V3 ldir; float t;
        if(is_zero(ls.n)){ ldir = ls.dir; t = INFINITY; }
        else{ ldir = ls.p - ip; t = length(ldir) - kEps; }
        const V3 sd = normalize(ldir);
        o[0] = ls.L.x; o[1] = ls.L.y; o[2] = ls.L.z; o[3] = ls.p.x; o[4] = ls.p.y; o[5] = ls.p.z; o[6] = ls.n.x; o[7] = ls.n.y; o[8] = ls.n.z;
        o[9] = ls.u; o[10] = ls.v; o[11] = ls.dir.x; o[12] = ls.dir.y; o[13] = ls.dir.z; o[14] = sd.x; o[15] = sd.y; o[16] = sd.z; o[17] = t;
        o[18] = light_pdf_sampled(S, L, ls.p, ls.n, ip, sd, time);
        const V3 Lv = light_L_sampled(S, L, ls, sd);
        o[19] = Lv.x; o[20] = Lv.y; o[21] = Lv.z;
    }
    return PT_OK;
}

PT_API int emul_unit_escape(void* p, size_t n, const float* dirs, float* out){
    EmulScene* e = (EmulScene*)p;
    const ShadeScene& S = e->T.view;
    for(size_t i = 0; i < n; i++){
        const V3 rd = normalize(v3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]));
        V3 le = v3(0.0f); float pp = 0.0f;
        for(int k = 0; k < S.n_inf_lights; k++){
            const DevLight& L = S.lights[S.inf_lights[k]];
            le += inf_Le(S, L, rd);
            pp += L.pmf_sample * light_pdf_sampled(S, L, v3(0.0f), v3(0.0f), v3(0.0f), rd, 0.0f);
        }
        out[4 * i] = le.x; out[4 * i + 1] = le.y; out[4 * i + 2] = le.z; out[4 * i + 3] = pp;
    }
    return PT_OK;
}

PT_API int emul_unit_camera(const pt_camera_desc* cd, int W, int H, size_t n, const double* in, float* out){
    CamParams C;
    make_camera(cd, W, H, C);
    for(size_t i = 0; i < n; i++){
        const double* q = in + 5 * i;
        V3 o, d; float t;
        camera_ray(C, W, H, (float)q[0], (float)q[1], (float)q[2], (float)q[3], (float)q[4], o, d, t);
        out[7 * i] = o.x; out[7 * i + 1] = o.y; out[7 * i + 2] = o.z; out[7 * i + 3] = d.x; out[7 * i + 4] = d.y; out[7 * i + 5] = d.z; out[7 * i + 6] = t;
    }
    return PT_OK;
}

PT_API uint32_t emul_unit_permutation_element(uint32_t i, uint32_t l, uint32_t p){ return permutation_element(i, l, p); }
PT_API uint64_t emul_unit_hash(uint32_t px, uint32_t py, uint64_t dim){ return hash_ppd(px, py, dim); }
PT_API int emul_unit_hg(float g, size_t n, const float* in, float* out){
    const float gc = fminf(fmaxf(g, -0.99f), 0.99f);        // HenyeyGreenstein ctor clamp (PhaseFunction.hpp:20), applied at pt_scene_add_medium
    for(size_t i = 0; i < n; i++){
        const float* q = in + 5 * i;
        const V3 wi = normalize(v3(q[0], q[1], q[2]));
        const V3 wo = hg_sample(gc, wi, q[3], q[4]);
        const float pdf = phase_hg(dot(wi, wo), gc);
        out[5 * i] = wo.x; out[5 * i + 1] = wo.y; out[5 * i + 2] = wo.z; out[5 * i + 3] = pdf; out[5 * i + 4] = pdf;
    }
    return PT_OK;
}

} // extern "C"
