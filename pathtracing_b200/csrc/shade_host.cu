// shade_host.cu — uploads the shading tables built by shade_tables.cpp and fills pt_scene::shade
// with the device pointers.
#include <cuda_runtime.h>
#include <cstdlib>
#include <cstring>
#include <cstdio>
#include <chrono>

#include "shade_tables.hpp"

namespace ptb {

namespace {

thread_local pt_scene* g_scene = nullptr;      // the scene build_shade_tables is uploading for (commit is not re-entrant)

template <class T>
bool up(void** dst, const std::vector<T>& v, const char* what){
    *dst = nullptr;
    size_t bytes = v.size() * sizeof(T);
    if(bytes == 0)bytes = 16;
    if(!cuda_ok(cudaMalloc(dst, bytes), what))return false;
    if(!v.empty() && !cuda_ok(cudaMemcpy(*dst, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice), what))return false;
    note_upload(g_scene, dst, bytes);
    return true;
}

// Pageable host memory to the device through two pinned staging buffers: a plain cudaMemcpy from a std::vector ran at 2-3 GB/s on the
// bench boxes (420 MB of vertex tables: 0.19-0.68 s); a host memcpy into pinned memory overlapped with the DMA of the previous piece
// reaches the memcpy rate.  Small copies go straight.
struct Staging {
    static constexpr size_t kPiece = (size_t)16 << 20;
    void* pin[2] = { nullptr, nullptr };
    cudaEvent_t ev[2] = { nullptr, nullptr };
    cudaStream_t st = nullptr;
    bool ok = false;
    bool init(){
        if(ok)return true;
        if(cudaMallocHost(&pin[0], kPiece) != cudaSuccess || cudaMallocHost(&pin[1], kPiece) != cudaSuccess ||
           cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming) != cudaSuccess || cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming) != cudaSuccess ||
           cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess){ cudaGetLastError(); release(); return false; }
        ok = true;
        return true;
    }
    void release(){
        for(int k = 0; k < 2; k++){ if(pin[k])cudaFreeHost(pin[k]); if(ev[k])cudaEventDestroy(ev[k]); pin[k] = nullptr; ev[k] = nullptr; }
        if(st)cudaStreamDestroy(st);
        st = nullptr; ok = false;
    }
    ~Staging(){ release(); }
    bool copy(void* dst, const void* src, size_t bytes, const char* what){
        if(bytes == 0)return true;
        if(bytes < ((size_t)2 << 20) || !init())return cuda_ok(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice), what);
        int k = 0;
        for(size_t off = 0; off < bytes; off += kPiece, k ^= 1){
            const size_t n = bytes - off < kPiece ? bytes - off : kPiece;
            if(!cuda_ok(cudaEventSynchronize(ev[k]), what))return false;        // the DMA that last used this buffer is done (a fresh event is complete)
            memcpy(pin[k], (const char*)src + off, n);
            if(!cuda_ok(cudaMemcpyAsync((char*)dst + off, pin[k], n, cudaMemcpyHostToDevice, st), what) || !cuda_ok(cudaEventRecord(ev[k], st), what))return false;
        }
        return cuda_ok(cudaStreamSynchronize(st), what);
    }
};

// per-triangle shading records of one mesh inside one BLAS: global vertex ids + mesh id (ShadeTri), and the emissive ordinals
__global__ void k_make_stri(const uint32_t* __restrict__ idx, uint32_t vtx_base, uint32_t mesh, uint32_t n, ShadeTri* __restrict__ out){
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if(t >= n)return;
    ShadeTri r;
    r.i0 = vtx_base + idx[3 * (size_t)t]; r.i1 = vtx_base + idx[3 * (size_t)t + 1]; r.i2 = vtx_base + idx[3 * (size_t)t + 2]; r.mesh = mesh;
    out[t] = r;
}
__global__ void k_fill_emis(int32_t* __restrict__ out, int32_t first, uint32_t n){
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if(t < n)out[t] = first + (int32_t)t;
}

} // namespace

bool h2d_staged(int device, void* dst, const void* src, size_t bytes, const char* what){
    static thread_local Staging stages[8];          // streams and events belong to a device
    return stages[device & 7].copy(dst, src, bytes, what);
}

namespace {

bool dev_alloc(void** dst, size_t bytes, const char* what){
    *dst = nullptr;
    if(bytes == 0)bytes = 16;
    if(!cuda_ok(cudaMalloc(dst, bytes), what))return false;
    note_upload(g_scene, dst, bytes);
    return true;
}

// The per-vertex and per-triangle tables assembled on the device from the meshes' own arrays (ShadeTables::geometry_on_device)
int geometry_tables_on_device(pt_scene* s, const ShadeTables& T){
    DeviceBuffers& b = s->buf;
    const size_t nv = T.n_vertices, ni = T.n_indices, nt = T.n_blas_tris;
    if(!dev_alloc(&b.vtx_pos, 12 * nv, "positions") || !dev_alloc(&b.vtx_nrm, 12 * nv, "normals") || !dev_alloc(&b.vtx_uv, 8 * nv, "texcoords") ||
       !dev_alloc(&b.vtx_tan, 12 * nv, "tangents") || !dev_alloc(&b.indices, 4 * ni, "indices") || !dev_alloc(&b.tri_mesh, sizeof(ShadeTri) * nt, "shading triangles") ||
       !dev_alloc(&b.tri_light, 4 * nt, "tri_emis"))return PT_ERR_CUDA;
    struct { int dev; bool copy(void* d, const void* h, size_t n, const char* w){ return h2d_staged(dev, d, h, n, w); } } stage { s->device };
    for(size_t mi = 0; mi < s->meshes.size(); mi++){
        const HostMesh& m = s->meshes[mi];
        const DevMesh& d = T.meshes[mi];
        const size_t v = m.pos.size() / 3;
        if(m.nrm.size() != 3 * v || m.uv.size() != 2 * v || (!m.tan.empty() && m.tan.size() != 3 * v)){ set_error("mesh vertex arrays of different lengths"); return PT_ERR_INVALID; }
        if(!stage.copy((float*)b.vtx_pos + 3 * (size_t)d.vtx_base, m.pos.data(), 12 * v, "H2D positions") ||
           !stage.copy((float*)b.vtx_nrm + 3 * (size_t)d.vtx_base, m.nrm.data(), 12 * v, "H2D normals") ||
           !stage.copy((float*)b.vtx_uv + 2 * (size_t)d.vtx_base, m.uv.data(), 8 * v, "H2D texcoords") ||
           !stage.copy((uint32_t*)b.indices + d.idx_base, m.idx.data(), 4 * m.idx.size(), "H2D indices"))return PT_ERR_CUDA;
        if(m.tan.empty()){ if(v && !cuda_ok(cudaMemset((float*)b.vtx_tan + 3 * (size_t)d.vtx_base, 0, 12 * v), "tangents clear"))return PT_ERR_CUDA; }
        else if(!stage.copy((float*)b.vtx_tan + 3 * (size_t)d.vtx_base, m.tan.data(), 12 * v, "H2D tangents"))return PT_ERR_CUDA;
    }
    if(nt && !cuda_ok(cudaMemset(b.tri_light, 0xFF, 4 * nt), "tri_emis clear"))return PT_ERR_CUDA;       // -1: not emissive
    for(const ShadeTables::TriRun& r : T.tri_runs){
        if(r.count == 0)continue;
        const DevMesh& d = T.meshes[r.mesh];
        k_make_stri<<<(r.count + 255) / 256, 256>>>((const uint32_t*)b.indices + d.idx_base, d.vtx_base, (uint32_t)r.mesh, r.count, (ShadeTri*)b.tri_mesh + r.first);
        if(r.emis_first >= 0)k_fill_emis<<<(r.count + 255) / 256, 256>>>((int32_t*)b.tri_light + r.first, r.emis_first, r.count);
    }
    return cuda_ok(cudaGetLastError(), "shading triangle kernels") ? PT_OK : PT_ERR_CUDA;
}

} // namespace

int build_shade_tables(pt_scene* s, const std::vector<uint32_t>& inst_order){
    ShadeTables T;
    static const bool dev_geom = []{ const char* e = getenv("PTB_GPU_PACK"); return !(e && atoi(e) == 0); }();
    T.geometry_on_device = dev_geom && s->builder == PT_BUILDER_GPU;      // the host builder keeps the all-host path (and is what the tests compare against)
    int rc = build_shade_tables_host(s, inst_order, T);
    if(rc != PT_OK)return rc;
    DeviceBuffers& b = s->buf;
    g_scene = s;
    const auto tu0 = std::chrono::steady_clock::now();
    if(T.geometry_on_device){
        rc = geometry_tables_on_device(s, T);
        if(rc != PT_OK)return rc;
    } else if(!up(&b.vtx_pos, T.pos, "positions") || !up(&b.vtx_nrm, T.nrm, "normals") || !up(&b.vtx_uv, T.uv, "texcoords") ||
              !up(&b.vtx_tan, T.tan, "tangents") || !up(&b.indices, T.idx, "indices") || !up(&b.tri_mesh, T.stri, "shading triangles") ||
              !up(&b.tri_light, T.tri_emis, "tri_emis"))return PT_ERR_CUDA;
    if(!up(&b.meshes, T.meshes, "meshes") || !up(&b.blas_info, T.blas, "blas info") || !up(&b.inst_shade, T.inst, "instances") ||
       !up(&b.textures, T.textures, "textures") || !up(&b.texels, T.texels, "texels") ||
       !up(&b.materials, T.materials, "materials") || !up(&b.media, T.media, "media") || !up(&b.lights, T.lights, "lights") ||
       !up(&b.sampler_lights, T.sampler_lights, "sampler lights") || !up(&b.light_cdf, T.sampler_cdf, "light cdf") ||
       !up(&b.inf_lights, T.inf_lights, "infinite lights") || !up(&b.env_cdf, T.env_cdf, "env cdf"))return PT_ERR_CUDA;
    if(getenv("PTB_PROFILE"))fprintf(stderr, "[ptb] shade tables: upload %.1f ms\n", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tu0).count());
    s->n_shade_vertices = T.n_vertices; s->n_shade_indices = T.n_indices; s->n_shade_tris = T.n_blas_tris;
    ShadeScene& D = s->shade;      // pointer fields: bind_device_pointers (scene_host.cu)
    D.n_lights = (int32_t)T.lights.size(); D.n_sampler_lights = (int32_t)T.sampler_lights.size(); D.n_inf_lights = (int32_t)T.inf_lights.size();
    D.sampler_kind = T.view.sampler_kind; D.total_power = T.view.total_power;
    D.scene_medium = T.view.scene_medium;
    D.env_w = T.view.env_w; D.env_h = T.view.env_h;
    return PT_OK;
}

} // namespace ptb
