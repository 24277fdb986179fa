// scene_host.cu — scene commit: host BVH build, packing into the device layout,
// one upload.  Replaces Model::BuildBlas (Model.hpp:41-60) + Scene::BuildTlas
// (Scene.hpp:20-24) + the light-sampler wiring of examples/example_1.cpp:99-101.
#include <cuda_runtime.h>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <thread>

#include "scene_host.hpp"

namespace ptb {

bool cuda_ok(int err, const char* what){
    if(err == (int)cudaSuccess)return true;
    set_error(std::string(what) + ": " + cudaGetErrorString((cudaError_t)err));
    return false;
}

void note_upload(pt_scene* s, void** slot, size_t bytes){
    s->uploads.push_back({ (size_t)((char*)slot - (char*)&s->buf), bytes });
}

void bind_device_pointers(pt_scene* s){
    const DeviceBuffers& b = s->buf;
    s->dev.n2 = (const Bvh2Node*)b.n2; s->dev.n4 = (const Bvh4Node*)b.n4; s->dev.n8 = (const Bvh8Node*)b.n8;
    s->dev.tris = (const Tri*)b.tris; s->dev.inst = (const Instance*)b.inst;
    s->dev.n4q = (const Bvh4QNode*)b.n4q; s->dev.tris64 = (const Tri64*)b.tris64;
    ShadeScene& D = s->shade;
    D.meshes = (const DevMesh*)b.meshes; D.blas = (const DevBlasInfo*)b.blas_info; D.inst = (const DevInstShade*)b.inst_shade;
    D.pos = (const float*)b.vtx_pos; D.nrm = (const float*)b.vtx_nrm; D.uv = (const float*)b.vtx_uv; D.tan = (const float*)b.vtx_tan;
    D.idx = (const uint32_t*)b.indices; D.stri = (const ShadeTri*)b.tri_mesh; D.tri_emis = (const int32_t*)b.tri_light;
    D.textures = (const DevTexture*)b.textures; D.texels = (const uint8_t*)b.texels;
    D.materials = (const DevMaterial*)b.materials; D.media = (const DevMedium*)b.media;
    D.lights = (const DevLight*)b.lights; D.sampler_lights = (const int32_t*)b.sampler_lights; D.sampler_cdf = (const float*)b.light_cdf;
    D.inf_lights = (const int32_t*)b.inf_lights; D.env_cdf = (const float*)b.env_cdf;
}

// a scene-wide array of `total` records: the host segments go up here, the device pieces are placed by the caller
template <class T>
static bool upload_segs(pt_scene* s, void** dst, const std::vector<T>& v, const std::vector<HostSeg>& segs, size_t total, const char* what){
    *dst = nullptr;
    size_t bytes = total * sizeof(T);
    if(bytes == 0)bytes = 16;
    if(!cuda_ok(cudaMalloc(dst, bytes), what))return false;
    for(const HostSeg& g : segs){
        if(g.count == 0)continue;
        if(g.host_off + g.count > v.size() || g.global_base + g.count > total){ set_error(std::string(what) + ": segment out of range"); return false; }
        if(!cuda_ok(cudaMemcpy((T*)*dst + g.global_base, v.data() + g.host_off, g.count * sizeof(T), cudaMemcpyHostToDevice), what))return false;
    }
    note_upload(s, dst, bytes);
    return true;
}

template <class T>
static bool upload(pt_scene* s, void** dst, const std::vector<T>& v, const char* what){
    *dst = nullptr;
    size_t bytes = v.size() * sizeof(T);
    if(bytes == 0)bytes = 16;   // keep pointers valid
    if(!cuda_ok(cudaMalloc(dst, bytes), what))return false;
    if(!v.empty() && !cuda_ok(cudaMemcpy(*dst, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice), what))return false;
    note_upload(s, dst, bytes);
    return true;
}

int build_shade_tables(pt_scene* s, const std::vector<uint32_t>& inst_order);   // shade_host.cu

void scene_free_device(pt_scene* s){
    DeviceBuffers& b = s->buf;
    void** ptrs[] = { &b.n2, &b.n4, &b.n8, &b.tris, &b.inst, &b.n4q, &b.tris64, &b.tile_list, &b.meshes, &b.blas_info, &b.inst_shade, &b.vtx_pos, &b.vtx_nrm, &b.vtx_uv, &b.vtx_tan,
        &b.indices, &b.tri_light, &b.textures, &b.texels, &b.materials, &b.media, &b.lights, &b.light_cdf, &b.env_cdf, &b.film, &b.work, &b.adapt,
        &b.stage_rays, &b.stage_hits, &b.stage_inst, &b.stage_occ, &b.stage_rays24, &b.stage_hits8, &b.stats, &b.tri_mesh, &b.inst_light_base, &b.sampler_lights, &b.inf_lights, &b.cursors, &b.film_sq, &b.stream_aux };
    for(void** p : ptrs){ if(*p){ cudaFree(*p); *p = nullptr; } }
    if(b.h_counters){ cudaFreeHost(b.h_counters); b.h_counters = nullptr; }
    for(void*& e : b.wave_events)if(e){ cudaEventDestroy((cudaEvent_t)e); e = nullptr; }
    if(b.h_stream){ cudaFreeHost(b.h_stream); b.h_stream = nullptr; }
    b.work_bytes = 0; b.adapt_bytes = 0; b.tile_list_bytes = 0; b.stage_cap = 0; b.stage24_cap = 0; b.stream_aux_bytes = 0; b.h_stream_bytes = 0;
}

int scene_commit(pt_scene* s, int width, int light_sampler){
    if(s->committed){ set_error("scene already committed"); return PT_ERR_STATE; }
    if(!cuda_ok(cudaSetDevice(s->device), "cudaSetDevice"))return PT_ERR_CUDA;
    s->width = width;
    s->light_sampler = light_sampler;

    const auto t0 = std::chrono::steady_clock::now();
    s->build_stats = pt_build_stats {};
    s->build_stats.builder = s->builder;
    GpuBuildTimes gt;
    // small hierarchies (a TLAS of a few instances, a two-triangle light) are not worth the launches
    static const uint32_t gpu_min = []{ const char* e = getenv("PTB_GPU_BUILD_MIN"); return e ? (uint32_t)atoll(e) : 4096u; }();
    auto gpu_builder = [](const Box* boxes, uint32_t n, BinaryBvh& out, void* user) -> int{
        if(n < gpu_min){ out = build_binary_bvh(boxes, n); return PT_OK; }
        return build_binary_bvh_gpu(boxes, n, 2, out, (GpuBuildTimes*)user);
    };
    PackedTraversal P;
    // PT_BUILDER_GPU: large BVH4 BLASes are packed on the device from the mesh arrays (scene_pack_gpu.cu: no tree download, no host
    // flatten, no per-triangle host array); everything else builds its binary tree on the device and flattens it on the host
    int prc = s->builder == PT_BUILDER_GPU ? pack_traversal(s, width, P, gpu_builder, &gt, pack_blas_on_device) : pack_traversal(s, width, P);
    if(prc != PT_OK)return prc;
    const auto t1 = std::chrono::steady_clock::now();
    std::vector<Bvh2Node>& n2 = P.n2;
    std::vector<Bvh4Node>& n4 = P.n4;
    std::vector<Tri>& tris = P.tris;
    std::vector<Instance>& insts = P.insts;

    // ---- upload ---------------------------------------------------------------------
    DeviceBuffers& b = s->buf;
    s->uploads.clear();
    static const std::vector<HostSeg> no_segs;
    if(!upload_segs(s, &b.n2, n2, width == 2 ? P.node_segs : no_segs, width == 2 ? P.total_nodes : 0, "upload bvh2 nodes") ||
       !upload_segs(s, &b.n4, n4, width == 4 ? P.node_segs : no_segs, width == 4 ? P.total_nodes : 0, "upload bvh4 clusters") ||
       !upload_segs(s, &b.n8, P.n8, width == 8 ? P.node_segs : no_segs, width == 8 ? P.total_nodes : 0, "upload bvh8 clusters") ||
       !upload_segs(s, &b.tris, tris, P.tri_segs, P.total_tris, "upload triangles") || !upload(s, &b.inst, insts, "upload instances")){
        scene_free_device(s);
        return PT_ERR_CUDA;
    }
    for(const DevBlasPiece& pc : P.dev_pieces)
        if(place_blas_piece(pc, width, width == 2 ? b.n2 : width == 4 ? b.n4 : b.n8, b.tris) != PT_OK){ scene_free_device(s); return PT_ERR_CUDA; }
    s->dev.fast = 0;
    if(s->layout == PT_LAYOUT_FAST){
        if(width != 4){ set_error("PT_LAYOUT_FAST needs bvh_width 4"); scene_free_device(s); return PT_ERR_INVALID; }
        std::vector<Bvh4QNode> n4q;
        encode_bvh4q(n4, n4q);
        std::vector<Tri64> t64(tris.size());
        for(size_t i = 0; i < tris.size(); i++){ t64[i].t = tris[i]; t64[i].pad[0] = t64[i].pad[1] = t64[i].pad[2] = t64[i].pad[3] = 0u; }
        if(!upload_segs(s, &b.n4q, n4q, P.node_segs, P.total_nodes, "upload quantised bvh4 clusters") ||
           !upload_segs(s, &b.tris64, t64, P.tri_segs, P.total_tris, "upload padded triangles")){ scene_free_device(s); return PT_ERR_CUDA; }
        for(const DevBlasPiece& pc : P.dev_pieces)
            if(encode_fast_range(b.n4, b.n4q, pc.node_base, pc.n_nodes, b.tris, b.tris64, pc.tri_base, pc.n_tris) != PT_OK){ scene_free_device(s); return PT_ERR_CUDA; }
        s->dev.fast = 1;
    }
    P.dev_pieces.clear();       // the pieces are in place: their buffers go back
    s->dev.tlas_root = P.tlas_root;
    s->dev.width = width;
    s->dev.single_root = P.single_root; s->dev.single_id = P.single_id;
    for(int a = 0; a < 3; a++){ s->dev.tlo[a] = P.bounds.lo[a]; s->dev.thi[a] = P.bounds.hi[a]; }
    s->total_tris = P.total_tris;
    s->total_nodes = P.total_nodes;
    s->n_inst_records = insts.size();

    const auto t2 = std::chrono::steady_clock::now();
    int rc = build_shade_tables(s, P.inst_order);
    if(rc != PT_OK){ scene_free_device(s); return rc; }
    bind_device_pointers(s);

    cudaStream_t st, st2, st3, st4;
    if(!cuda_ok(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking), "cudaStreamCreate") ||
       !cuda_ok(cudaStreamCreateWithFlags(&st2, cudaStreamNonBlocking), "cudaStreamCreate") ||
       !cuda_ok(cudaStreamCreateWithFlags(&st3, cudaStreamNonBlocking), "cudaStreamCreate") ||
       !cuda_ok(cudaStreamCreateWithFlags(&st4, cudaStreamNonBlocking), "cudaStreamCreate")){ scene_free_device(s); return PT_ERR_CUDA; }
    s->stream = st; s->stream2 = st2; s->stream3 = st3; s->stream4 = st4;
    if(!cuda_ok(cudaMalloc(&b.stats, 256), "cudaMalloc stats") || !cuda_ok(cudaMemset(b.stats, 0, 256), "stats clear") ||
       !cuda_ok(cudaMalloc(&b.cursors, kCursorRing * sizeof(unsigned long long)), "cudaMalloc cursors")){ scene_free_device(s); return PT_ERR_CUDA; }
    // the tables went up as pageable cudaMemcpy on the legacy stream; the non-blocking streams created above do not
    // order against it: make everything land before the first kernel can read it
    if(!cuda_ok(cudaDeviceSynchronize(), "scene commit")){ scene_free_device(s); return PT_ERR_CUDA; }
    s->committed = true;
    const auto t3 = std::chrono::steady_clock::now();
    auto sec = [](auto a, auto b){ return std::chrono::duration<double>(b - a).count(); };
    pt_build_stats& bs = s->build_stats;
    bs.total_s = sec(t0, t3); bs.pack_s = sec(t0, t1); bs.upload_s = sec(t1, t2); bs.shade_tables_s = sec(t2, t3);
    bs.gpu_build_ms = gt.build_ms; bs.gpu_upload_ms = gt.upload_ms; bs.gpu_download_ms = gt.download_ms;
    bs.gpu_levels = gt.levels; bs.gpu_subtrees = gt.subtrees; bs.gpu_items = gt.items;
    return PT_OK;
}

} // namespace ptb
