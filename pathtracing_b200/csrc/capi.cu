// capi.cu — the extern "C" surface declared in include/pt_b200.h.
#include <cuda_runtime.h>
#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>

#include "scene_host.hpp"

using namespace ptb;

#define PT_CHECK_SCENE(s) do{ if(!(s)){ set_error("null scene"); return PT_ERR_INVALID; } }while(0)
#define PT_CHECK_OPEN(s) do{ PT_CHECK_SCENE(s); if((s)->committed){ set_error("scene already committed"); return PT_ERR_STATE; } }while(0)
#define PT_CHECK_READY(s) do{ PT_CHECK_SCENE(s); if(!(s)->committed){ set_error("scene not committed"); return PT_ERR_STATE; } \
    if(!cuda_ok(cudaSetDevice((s)->device), "cudaSetDevice"))return PT_ERR_CUDA; }while(0)

static int ensure_stage(pt_scene* s, size_t n){
    DeviceBuffers& b = s->buf;
    if(n <= b.stage_cap)return PT_OK;
    void** ptrs[] = { &b.stage_rays, &b.stage_hits, &b.stage_inst, &b.stage_occ };
    for(void** p : ptrs){ if(*p){ cudaFree(*p); *p = nullptr; } }
    b.stage_cap = 0;
    if(!cuda_ok(cudaMalloc(&b.stage_rays, n * sizeof(pt_ray)), "cudaMalloc rays") ||
       !cuda_ok(cudaMalloc(&b.stage_hits, n * sizeof(pt_hit)), "cudaMalloc hits") ||
       !cuda_ok(cudaMalloc(&b.stage_inst, n * sizeof(uint32_t)), "cudaMalloc inst") ||
       !cuda_ok(cudaMalloc(&b.stage_occ, n), "cudaMalloc occ"))return PT_ERR_CUDA;
    b.stage_cap = n;
    return PT_OK;
}

static int ensure_stage24(pt_scene* s, size_t n){
    DeviceBuffers& b = s->buf;
    if(n <= b.stage24_cap)return PT_OK;
    void** ptrs[] = { &b.stage_rays24, &b.stage_hits8 };
    for(void** p : ptrs){ if(*p){ cudaFree(*p); *p = nullptr; } }
    b.stage24_cap = 0;
    if(!cuda_ok(cudaMalloc(&b.stage_rays24, n * sizeof(pt_ray24)), "cudaMalloc compact rays") ||
       !cuda_ok(cudaMalloc(&b.stage_hits8, n * sizeof(pt_hit8)), "cudaMalloc compact hits"))return PT_ERR_CUDA;
    b.stage24_cap = n;
    return PT_OK;
}

// Host-buffer entry points: a three-stage pipeline over three streams.  Chunk k is copied in on the H2D
// stream, traced on the compute stream once its copy has landed, and copied out on the D2H stream once its
// kernel is done — so the kernels run back to back while both PCIe directions stay busy (full duplex).  The
// first chunks are small so that the compute stream starts early; the staging buffers hold the whole batch,
// so no chunk ever waits for a buffer.
static size_t chunk_rays(){
    static size_t v = 0;
    if(!v){
        v = 2u << 20;      // 2 Mi rays: best of the 0.75-8 Mi sweep with two trace streams (scratch/e2e_sweep.py)
        if(const char* e = getenv("PTB_CHUNK")){ const long long x = atoll(e); if(x >= 1024)v = (size_t)x; }
    }
    return v;
}

static cudaEvent_t pipeline_event(pt_scene* s, size_t i){
    while(s->events.size() <= i){
        cudaEvent_t e = nullptr;
        if(!cuda_ok(cudaEventCreateWithFlags(&e, cudaEventDisableTiming), "cudaEventCreate"))return nullptr;
        s->events.push_back(e);
    }
    return (cudaEvent_t)s->events[i];
}

// launch(off, m, stream) queues the kernel for rays [off, off+m); copy_out(off, m, stream) queues its D2H copies
// rays / ray_bytes / d_in: the host records, their size and where they are staged (pt_ray into stage_rays, or pt_ray24 into stage_rays24)
template <class Launch, class CopyOut>
static int trace_pipelined(pt_scene* s, const void* rays, size_t n, Launch launch, CopyOut copy_out, size_t ray_bytes = sizeof(pt_ray), void* d_in = nullptr){
    int rc = ensure_stage(s, n);
    if(rc != PT_OK)return rc;
    // Chunks alternate between two trace streams: the persistent kernel of chunk k+1 moves onto the SMs that
    // chunk k's tail (its longest rays) leaves idle.  PTB_TRACE_STREAMS=1 serialises them again.
    static const int n_cs = []{ const char* e = getenv("PTB_TRACE_STREAMS"); return e && atoi(e) == 1 ? 1 : 2; }();
    cudaStream_t cs[2] = { (cudaStream_t)s->stream, n_cs == 2 ? (cudaStream_t)s->stream4 : (cudaStream_t)s->stream };
    cudaStream_t in = (cudaStream_t)s->stream2, out = (cudaStream_t)s->stream3;
    char* d_rays = (char*)(d_in ? d_in : s->buf.stage_rays);
    const size_t full = chunk_rays();
    size_t k = 0;
    for(size_t off = 0; off < n; k++){
        size_t m = k == 0 ? full / 8 : k == 1 ? full / 2 : full;      // ramp up: 1/8, 1/2, then full chunks
        if(m < 1024)m = 1024;
        const size_t left = n - off;
        if(m >= left)m = left > full / 2 ? left - full / 4 : left;    // ramp down: the last D2H is a quarter chunk at most
        cudaStream_t compute = cs[k & 1];
        cudaEvent_t landed = pipeline_event(s, 2 * k), traced = pipeline_event(s, 2 * k + 1);
        if(!landed || !traced)return PT_ERR_CUDA;
        if(!cuda_ok(cudaMemcpyAsync(d_rays + off * ray_bytes, (const char*)rays + off * ray_bytes, m * ray_bytes, cudaMemcpyHostToDevice, in), "H2D rays") ||
           !cuda_ok(cudaEventRecord(landed, in), "event") || !cuda_ok(cudaStreamWaitEvent(compute, landed, 0), "wait"))return PT_ERR_CUDA;
        rc = launch(off, m, compute);
        if(rc != PT_OK)return rc;
        if(!cuda_ok(cudaEventRecord(traced, compute), "event") || !cuda_ok(cudaStreamWaitEvent(out, traced, 0), "wait"))return PT_ERR_CUDA;
        rc = copy_out(off, m, out);
        if(rc != PT_OK)return rc;
        off += m;
    }
    if(!cuda_ok(cudaStreamSynchronize(out), "sync") || !cuda_ok(cudaStreamSynchronize(cs[0]), "sync") || !cuda_ok(cudaStreamSynchronize(cs[1]), "sync") ||
       !cuda_ok(cudaStreamSynchronize(in), "sync"))return PT_ERR_CUDA;
    return PT_OK;
}

// Traversal-stack overflow (trace.cuh TF_OVERFLOW) is counted on the device; the synchronous entry points fail the call
// whose rays moved the counter (the asynchronous *_dev ones cannot: pt_scene_stack_overflows reads it).
static int check_overflow(pt_scene* s, const char* what){
    unsigned long long now = 0;
    const int rc = scene_overflow_count(s, &now);
    if(rc != PT_OK)return rc;
    if(now != s->overflow_seen){
        set_error(std::string(what) + ": traversal stack overflow on " + std::to_string(now - s->overflow_seen) + " rays (more than 96 stack entries needed); their hits may be wrong");
        s->overflow_seen = now;
        return PT_ERR_OVERFLOW;
    }
    return PT_OK;
}

constexpr int kStreamGaveUp = 1000;    // internal: the streamed kernel timed out waiting for rays -> the caller runs the chunked pipeline

// Streamed form of the same pipeline, for plain scenes and large batches: ONE persistent kernel over the whole
// batch, launched before the first ray has arrived.  The H2D stream copies the rays in chunks and moves a watermark
// in device memory after each; a warp that grabs rays beyond the watermark waits for it (trace.cu feed_grab<true>).
// Every retired ray bumps the counter of its 256 Ki-ray granule (after a fence), the host polls the counters and
// sends the finished prefix back on the D2H stream.  Against one kernel per chunk this removes every chunk's tail
// (a 2 Mi-ray launch runs at 0.91 Grays/s, the 16 Mi-ray one at 1.06: profiles/r01_trace4_v7.md).
// copy_out(off, m, stream) queues the D2H copies of rays [off, off + m).
// compact: the rays are pt_ray24 records staged in stage_rays24, the kernel writes pt_hit8 into stage_hits8
template <class CopyOut>
static int trace_streamed(pt_scene* s, const void* rays, size_t n, bool any, CopyOut copy_out, bool compact = false){
    int rc = ensure_stage(s, n);
    if(rc != PT_OK)return rc;
    const size_t ray_bytes = compact ? sizeof(pt_ray24) : sizeof(pt_ray);
    DeviceBuffers& b = s->buf;
    const unsigned shift = 18;
    const size_t ng = (n + ((size_t)1 << shift) - 1) >> shift;
    const size_t aux_bytes = 64 + 1024 * ng;       // per granule 32 sub-counters, 32 B apart
    // Chunk sizes of the H2D stream (measured on the C2 workload, scratch/r02_stream_sweep.py).  The 32 B form is bound by the
    // link: 1 Mi-ray copies (every copy costs the stream about 18 us, smaller ones lose more than they smooth), and the last
    // 2 Mi rays in 256 Ki-ray pieces so that little work is left when the last piece lands (12.01 -> 11.85 ms).  The 24 B form is
    // bound by the kernel: 256 Ki-ray copies feed it more evenly (11.27 -> 10.94 ms).  PTB_STREAM_CHUNK_LOG2 / PTB_STREAM_TAPER_LOG2 override both.
    static const int chunk_env = []{ const char* e = getenv("PTB_STREAM_CHUNK_LOG2"); const int v = e ? atoi(e) : 0; return v == 0 ? 0 : v < 16 ? 16 : v > 24 ? 24 : v; }();
    static const unsigned taper_log2 = []{ const char* e = getenv("PTB_STREAM_TAPER_LOG2"); const int v = e ? atoi(e) : 21; return (unsigned)(v < 0 ? 0 : v > 30 ? 30 : v); }();
    const unsigned chunk_log2 = chunk_env ? (unsigned)chunk_env : compact ? 18u : 20u;
    const size_t chunk = (size_t)1 << chunk_log2;
    const size_t small_chunk = std::min(chunk, (size_t)1 << 18);
    const size_t taper_from = taper_log2 >= 18 && n > ((size_t)1 << taper_log2) ? n - ((size_t)1 << taper_log2) : n;
    const size_t n_chunks = (n + chunk - 1) / chunk + (n - taper_from) / small_chunk + 4;
    if(aux_bytes > b.stream_aux_bytes){
        if(b.stream_aux){ cudaFree(b.stream_aux); b.stream_aux = nullptr; b.stream_aux_bytes = 0; }
        if(!cuda_ok(cudaMalloc(&b.stream_aux, aux_bytes), "cudaMalloc stream ctl"))return PT_ERR_CUDA;
        b.stream_aux_bytes = aux_bytes;
    }
    const size_t h_bytes = 8 * n_chunks + aux_bytes + 64;
    if(h_bytes > b.h_stream_bytes){
        if(b.h_stream){ cudaFreeHost(b.h_stream); b.h_stream = nullptr; b.h_stream_bytes = 0; }
        if(!cuda_ok(cudaMallocHost(&b.h_stream, h_bytes), "pinned stream ctl"))return PT_ERR_CUDA;
        b.h_stream_bytes = h_bytes;
    }
    unsigned long long* h_wm = (unsigned long long*)b.h_stream;
    unsigned char* h_poll = (unsigned char*)b.h_stream + 8 * n_chunks;
    h_poll += (64 - ((uintptr_t)h_poll & 63)) & 63;
    cudaStream_t compute = (cudaStream_t)s->stream, in = (cudaStream_t)s->stream2, out = (cudaStream_t)s->stream3, poll = (cudaStream_t)s->stream4;
    cudaEvent_t ready = pipeline_event(s, 0);
    if(!ready)return PT_ERR_CUDA;
    StreamCtl ctl;
    ctl.landed = (const unsigned long long*)b.stream_aux;
    ctl.abort = (unsigned*)((char*)b.stream_aux + 8);
    ctl.done = (unsigned*)((char*)b.stream_aux + 64);
    ctl.shift = shift;
    if(!cuda_ok(cudaMemsetAsync(b.stream_aux, 0, aux_bytes, compute), "stream ctl reset") || !cuda_ok(cudaEventRecord(ready, compute), "event"))return PT_ERR_CUDA;
    static const bool prof = getenv("PTB_PROFILE") != nullptr;
    cudaEvent_t pe0 = nullptr, pe1 = nullptr, pe2 = nullptr, pe3 = nullptr;
    if(prof){ cudaEventCreate(&pe0); cudaEventCreate(&pe1); cudaEventCreate(&pe2); cudaEventCreate(&pe3); cudaEventRecord(pe0, compute); }
    rc = compact ? launch_trace_streamed(s, false, b.stage_rays24, n, b.stage_hits8, b.stage_inst, nullptr, ctl, compute, true)
                 : launch_trace_streamed(s, any, b.stage_rays, n, b.stage_hits, b.stage_inst, b.stage_occ, ctl, compute);
    if(rc != PT_OK)return rc;
    if(prof)cudaEventRecord(pe1, compute);
    // rays in: a small first chunk so that the kernel has work at once, then 1 Mi-ray chunks, the watermark after each
    if(!cuda_ok(cudaStreamWaitEvent(in, ready, 0), "wait") || !cuda_ok(cudaStreamWaitEvent(poll, ready, 0), "wait"))return PT_ERR_CUDA;
    char* d_rays = (char*)(compact ? b.stage_rays24 : b.stage_rays);
    size_t k = 0;
    bool enqueue_failed = false;
    for(size_t off = 0; off < n && !enqueue_failed; k++){
        size_t m = k == 0 ? chunk / 16 : k == 1 ? chunk / 4 : chunk;
        if(off >= taper_from)m = small_chunk;
        else if(m > taper_from - off)m = taper_from - off;
        if(m > n - off)m = n - off;
        h_wm[k] = (unsigned long long)(off + m);
        if(!cuda_ok(cudaMemcpyAsync(d_rays + off * ray_bytes, (const char*)rays + off * ray_bytes, m * ray_bytes, cudaMemcpyHostToDevice, in), "H2D rays") ||
           !cuda_ok(cudaMemcpyAsync(b.stream_aux, h_wm + k, 8, cudaMemcpyHostToDevice, in), "H2D watermark"))enqueue_failed = true;
        off += m;
    }
    if(prof)cudaEventRecord(pe3, in);
    // results out: poll the granule counters, send every finished prefix
    const unsigned* h_done = (const unsigned*)(h_poll + 64);
    const unsigned* h_abort = (const unsigned*)(h_poll + 8);
    size_t sent = 0, g_ready = 0;
    bool failed = enqueue_failed, gave_up = false;
    const auto t0 = std::chrono::steady_clock::now();
    auto t_progress = t0;                   // the 20 s limit counts from the last granule that finished, not from the start of the call
    while(sent < n && !failed){
        if(!cuda_ok(cudaMemcpyAsync(h_poll, b.stream_aux, aux_bytes, cudaMemcpyDeviceToHost, poll), "poll") || !cuda_ok(cudaStreamSynchronize(poll), "poll")){ failed = true; break; }
        if(*h_abort){ set_error("streamed ray cast: the kernel gave up waiting for its rays"); failed = true; gave_up = true; break; }
        while(g_ready < ng){
            const size_t g_lo = g_ready << shift, g_n = std::min(n - g_lo, (size_t)1 << shift);
            unsigned long long got = 0;
            for(int q = 0; q < 32; q++)got += h_done[(g_ready * 32 + q) * 8];
            if(got != (unsigned long long)g_n)break;
            g_ready++;
        }
        const size_t upto = std::min(n, g_ready << shift);
        if(upto > sent){
            if(copy_out(sent, upto - sent, out) != PT_OK){ failed = true; break; }
            sent = upto;
            t_progress = std::chrono::steady_clock::now();
        } else if(std::chrono::duration<double>(std::chrono::steady_clock::now() - t_progress).count() > 20.0){
            set_error("streamed ray cast: timed out"); failed = true; break;
        }
    }
    if(failed){
        // make sure the kernel ends: raise the abort flag it polls while waiting
        static const unsigned one = 1u;
        cudaMemcpyAsync((char*)b.stream_aux + 8, &one, 4, cudaMemcpyHostToDevice, poll);
        cudaStreamSynchronize(poll); cudaStreamSynchronize(in); cudaStreamSynchronize(compute); cudaStreamSynchronize(out);
        return gave_up ? kStreamGaveUp : (int)PT_ERR_CUDA;
    }
    if(prof)cudaEventRecord(pe2, out);
    if(!cuda_ok(cudaStreamSynchronize(out), "sync") || !cuda_ok(cudaStreamSynchronize(compute), "sync") || !cuda_ok(cudaStreamSynchronize(in), "sync"))return PT_ERR_CUDA;
    if(prof){
        float k = 0, t = 0, h = 0; cudaEventElapsedTime(&k, pe0, pe1); cudaEventElapsedTime(&t, pe0, pe2); cudaEventElapsedTime(&h, pe0, pe3);
        cudaEventDestroy(pe3);
        fprintf(stderr, "[ptb] streamed: H2D done at %.3f ms, kernel %.3f ms, last D2H done at %.3f ms, host %.3f ms\n", h, k, t,
                std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() * 1e3);
        cudaEventDestroy(pe0); cudaEventDestroy(pe1); cudaEventDestroy(pe2);
    }
    return PT_OK;
}

// which form a host-buffer call takes: PTB_E2E_STREAMED=0 keeps one kernel per chunk everywhere
static bool use_streamed(pt_scene* s, size_t n){
    static const bool on = []{ const char* e = getenv("PTB_E2E_STREAMED"); return !(e && atoi(e) == 0); }();
    if(!on || s->streamed_off || n < ((size_t)1 << 20) || s->has_alpha_tex || s->has_wrappers)return false;
    if(!s->streamed_probed){
        // once per scene: a process whose launch calls block (profiler, CUDA_LAUNCH_BLOCKING) cannot feed a running kernel
        s->streamed_probed = true;
        if(!probe_async_launch(s, s->buf.stats, s->stream, s->stream2)){ cudaGetLastError(); s->streamed_off = true; return false; }
    }
    return true;
}

extern "C" {

int pt_device_count(void){
    int n = 0;
    if(cudaGetDeviceCount(&n) != cudaSuccess){ cudaGetLastError(); return 0; }
    return n;
}

pt_scene* pt_scene_create(int device){
    int n = pt_device_count();
    if(n <= 0){ set_error("no CUDA device: pathtracing_b200 has no CPU fallback"); return nullptr; }
    if(device < 0 || device >= n){ set_error("bad device index"); return nullptr; }
    pt_scene* s = new (std::nothrow) pt_scene();
    if(!s){ set_error("out of memory"); return nullptr; }
    s->device = device;
    if(const char* e = getenv("PTB_BUILDER"))s->builder = e[0] == 'h' ? PT_BUILDER_HOST : PT_BUILDER_GPU;
    return s;
}

void pt_scene_destroy(pt_scene* s){
    if(!s)return;
    pt_dist_finalize(s);
    if(s->committed || s->buf.n2 || s->buf.n4){
        cudaSetDevice(s->device);
        cudaDeviceSynchronize();
        scene_free_device(s);
        if(s->stream)cudaStreamDestroy((cudaStream_t)s->stream);
        if(s->stream2)cudaStreamDestroy((cudaStream_t)s->stream2);
        if(s->stream3)cudaStreamDestroy((cudaStream_t)s->stream3);
        if(s->stream4)cudaStreamDestroy((cudaStream_t)s->stream4);
        for(void* e : s->events)cudaEventDestroy((cudaEvent_t)e);
    }
    delete s;
}

const char* pt_last_error(void){ return get_error(); }

int pt_scene_add_texture(pt_scene* s, const pt_texture_desc* d){
    PT_CHECK_OPEN(s);
    return scene_add_texture(s, d);
}

int pt_scene_add_material(pt_scene* s, const pt_material_desc* d){
    PT_CHECK_OPEN(s);
    return scene_add_material(s, d);
}

int pt_scene_add_medium(pt_scene* s, const pt_medium_desc* d){
    PT_CHECK_OPEN(s);
    return scene_add_medium(s, d);
}

int pt_scene_add_mesh(pt_scene* s, const pt_mesh_desc* d){
    PT_CHECK_OPEN(s);
    return scene_add_mesh(s, d);
}

int pt_scene_add_blas(pt_scene* s, const int32_t* mesh_ids, int n){
    PT_CHECK_OPEN(s);
    return scene_add_blas(s, mesh_ids, n);
}

int pt_scene_add_instance(pt_scene* s, int blas_id, const float* xf){
    PT_CHECK_OPEN(s);
    return scene_add_instance(s, blas_id, xf);
}

int pt_scene_add_shape(pt_scene* s, const pt_shape_desc* d){
    PT_CHECK_OPEN(s);
    return scene_add_shape(s, d);
}

int pt_scene_add_shape_xf(pt_scene* s, const pt_shape_desc* d, const float* xf){
    PT_CHECK_OPEN(s);
    if(!xf){ set_error("null transform"); return PT_ERR_INVALID; }
    return scene_add_shape(s, d, xf);
}

int pt_scene_animate(pt_scene* s, int prim_id, const float* dir, const float* tb){
    PT_CHECK_OPEN(s);
    return scene_animate(s, prim_id, dir, tb);
}

int pt_scene_add_light(pt_scene* s, const pt_light_desc* d){
    PT_CHECK_OPEN(s);
    return scene_add_light(s, d);
}

int pt_scene_set_medium(pt_scene* s, int medium_id){
    PT_CHECK_OPEN(s);
    if(medium_id >= (int)s->media.size()){ set_error("medium id out of range"); return PT_ERR_INVALID; }
    s->scene_medium = medium_id;
    return PT_OK;
}

int pt_scene_commit(pt_scene* s, int bvh_width, int light_sampler){
    PT_CHECK_SCENE(s);
    return scene_commit(s, bvh_width, light_sampler);
}

int pt_scene_set_builder(pt_scene* s, int builder){
    PT_CHECK_OPEN(s);
    if(builder != PT_BUILDER_HOST && builder != PT_BUILDER_GPU){ set_error("unknown builder"); return PT_ERR_INVALID; }
    s->builder = builder;
    return PT_OK;
}

int pt_scene_set_layout(pt_scene* s, int layout){
    PT_CHECK_OPEN(s);
    if(layout != PT_LAYOUT_EXACT && layout != PT_LAYOUT_FAST){ set_error("unknown layout"); return PT_ERR_INVALID; }
    s->layout = layout;
    return PT_OK;
}

int pt_scene_set_allreduce(pt_scene* s, pt_allreduce_fn sum, void* user){
    PT_CHECK_SCENE(s);
    s->allreduce = sum; s->allreduce_user = user;
    return PT_OK;
}

int pt_scene_build_stats(const pt_scene* s, pt_build_stats* out){
    PT_CHECK_SCENE(s);
    if(!out){ set_error("null stats"); return PT_ERR_INVALID; }
    if(!s->committed){ set_error("scene not committed"); return PT_ERR_STATE; }
    *out = s->build_stats;
    return PT_OK;
}

int pt_build_bvh2(int device, int builder, const float* boxes6, uint32_t n, pt_bvh2_node* nodes, uint32_t node_capacity,
                  uint32_t* n_nodes, uint32_t* order, float* gpu_build_ms){
    static_assert(sizeof(pt_bvh2_node) == sizeof(BuildNode), "pt_bvh2_node must match BuildNode");
    if(!boxes6 || !nodes || !n_nodes || !order){ set_error("null buffer"); return PT_ERR_INVALID; }
    if(n == 0){ *n_nodes = 0; return PT_OK; }
    if((uint64_t)node_capacity < 2ull * n - 1){ set_error("node_capacity must be >= 2n - 1"); return PT_ERR_INVALID; }
    BinaryBvh b;
    if(builder == PT_BUILDER_HOST)b = build_binary_bvh((const Box*)boxes6, n);
    else if(builder == PT_BUILDER_GPU){
        const int nd = pt_device_count();
        if(nd <= 0){ set_error("no CUDA device: pathtracing_b200 has no CPU fallback"); return PT_ERR_NO_DEVICE; }
        if(device < 0 || device >= nd){ set_error("bad device index"); return PT_ERR_INVALID; }
        if(!cuda_ok(cudaSetDevice(device), "cudaSetDevice"))return PT_ERR_CUDA;
        GpuBuildTimes gt;
        const int rc = build_binary_bvh_gpu((const Box*)boxes6, n, 2, b, &gt);
        if(rc != PT_OK)return rc;
        if(gpu_build_ms)*gpu_build_ms = gt.build_ms;
    } else{ set_error("unknown builder"); return PT_ERR_INVALID; }
    const BinaryBvh c = canonical_bvh(b);
    if(c.nodes.size() > node_capacity){ set_error("internal: node count exceeds 2n - 1"); return PT_ERR_INVALID; }
    memcpy(nodes, c.nodes.data(), c.nodes.size() * sizeof(BuildNode));
    memcpy(order, c.order.data(), (size_t)n * 4);
    *n_nodes = (uint32_t)c.nodes.size();
    return PT_OK;
}

int pt_scene_bounds(const pt_scene* s, float bmin[3], float bmax[3]){
    PT_CHECK_SCENE(s);
    if(!s->committed){ set_error("scene not committed"); return PT_ERR_STATE; }
    for(int a = 0; a < 3; a++){ bmin[a] = s->bounds.lo[a]; bmax[a] = s->bounds.hi[a]; }
    return PT_OK;
}

// ---- ray casting -------------------------------------------------------------------

int pt_trace_closest(pt_scene* s, const pt_ray* rays, size_t n, pt_hit* hits, uint32_t* inst){
    PT_CHECK_READY(s);
    if(n == 0)return PT_OK;
    if(!rays || !hits){ set_error("null buffer"); return PT_ERR_INVALID; }
    DeviceBuffers& b = s->buf;
    auto out_hits = [&](size_t off, size_t m, cudaStream_t st){
        if(!cuda_ok(cudaMemcpyAsync(hits + off, (pt_hit*)b.stage_hits + off, m * sizeof(pt_hit), cudaMemcpyDeviceToHost, st), "D2H hits"))return (int)PT_ERR_CUDA;
        if(inst && !cuda_ok(cudaMemcpyAsync(inst + off, (uint32_t*)b.stage_inst + off, m * sizeof(uint32_t), cudaMemcpyDeviceToHost, st), "D2H inst"))return (int)PT_ERR_CUDA;
        return (int)PT_OK;
    };
    if(use_streamed(s, n)){
        const int rc = trace_streamed(s, rays, n, false, out_hits);
        if(rc != kStreamGaveUp)return rc != PT_OK ? rc : check_overflow(s, "pt_trace_closest");
        s->streamed_off = true;      // this process blocks in kernel launches: stay with one kernel per chunk
    }
    const int prc = trace_pipelined(s, rays, n,
        [&](size_t off, size_t m, cudaStream_t st){
            return launch_trace_closest(s, (pt_ray*)b.stage_rays + off, m, (pt_hit*)b.stage_hits + off, inst ? (uint32_t*)b.stage_inst + off : nullptr, st);
        },
        [&](size_t off, size_t m, cudaStream_t st){
            if(!cuda_ok(cudaMemcpyAsync(hits + off, (pt_hit*)b.stage_hits + off, m * sizeof(pt_hit), cudaMemcpyDeviceToHost, st), "D2H hits"))return (int)PT_ERR_CUDA;
            if(inst && !cuda_ok(cudaMemcpyAsync(inst + off, (uint32_t*)b.stage_inst + off, m * sizeof(uint32_t), cudaMemcpyDeviceToHost, st), "D2H inst"))return (int)PT_ERR_CUDA;
            return (int)PT_OK;
        });
    return prc != PT_OK ? prc : check_overflow(s, "pt_trace_closest");
}

int pt_trace_closest_compact(pt_scene* s, const pt_ray24* rays, size_t n, pt_hit8* hits, uint32_t* inst){
    PT_CHECK_READY(s);
    if(n == 0)return PT_OK;
    if(!rays || !hits){ set_error("null buffer"); return PT_ERR_INVALID; }
    int rc = ensure_stage24(s, n);
    if(rc != PT_OK)return rc;
    DeviceBuffers& b = s->buf;
    auto out_hits = [&](size_t off, size_t m, cudaStream_t st){
        if(!cuda_ok(cudaMemcpyAsync(hits + off, (pt_hit8*)b.stage_hits8 + off, m * sizeof(pt_hit8), cudaMemcpyDeviceToHost, st), "D2H hits"))return (int)PT_ERR_CUDA;
        if(inst && !cuda_ok(cudaMemcpyAsync(inst + off, (uint32_t*)b.stage_inst + off, m * sizeof(uint32_t), cudaMemcpyDeviceToHost, st), "D2H inst"))return (int)PT_ERR_CUDA;
        return (int)PT_OK;
    };
    if(use_streamed(s, n)){
        rc = trace_streamed(s, rays, n, false, out_hits, true);
        if(rc != kStreamGaveUp)return rc != PT_OK ? rc : check_overflow(s, "pt_trace_closest_compact");
        s->streamed_off = true;
    }
    // one kernel per chunk: expand the chunk to pt_ray on the device, the ordinary kernel, pack its hits to 8 B
    const int prc = trace_pipelined(s, rays, n,
        [&](size_t off, size_t m, cudaStream_t st){
            int r = launch_expand_rays24((pt_ray24*)b.stage_rays24 + off, (pt_ray*)b.stage_rays + off, m, st);
            if(r == PT_OK)r = launch_trace_closest(s, (pt_ray*)b.stage_rays + off, m, (pt_hit*)b.stage_hits + off, inst ? (uint32_t*)b.stage_inst + off : nullptr, st);
            if(r == PT_OK)r = launch_pack_hits8((pt_hit*)b.stage_hits + off, (pt_hit8*)b.stage_hits8 + off, m, st);
            return r;
        }, out_hits, sizeof(pt_ray24), b.stage_rays24);
    return prc != PT_OK ? prc : check_overflow(s, "pt_trace_closest_compact");
}

int pt_trace_any(pt_scene* s, const pt_ray* rays, size_t n, uint8_t* occluded){
    PT_CHECK_READY(s);
    if(n == 0)return PT_OK;
    if(!rays || !occluded){ set_error("null buffer"); return PT_ERR_INVALID; }
    DeviceBuffers& b = s->buf;
    if(use_streamed(s, n)){
        const int rc = trace_streamed(s, rays, n, true, [&](size_t off, size_t m, cudaStream_t st){
            return cuda_ok(cudaMemcpyAsync(occluded + off, (uint8_t*)b.stage_occ + off, m, cudaMemcpyDeviceToHost, st), "D2H occ") ? (int)PT_OK : (int)PT_ERR_CUDA; });
        if(rc != kStreamGaveUp)return rc != PT_OK ? rc : check_overflow(s, "pt_trace_any");
        s->streamed_off = true;      // this process blocks in kernel launches: stay with one kernel per chunk
    }
    const int prc = trace_pipelined(s, rays, n,
        [&](size_t off, size_t m, cudaStream_t st){ return launch_trace_any(s, (pt_ray*)b.stage_rays + off, m, (uint8_t*)b.stage_occ + off, st); },
        [&](size_t off, size_t m, cudaStream_t st){
            return cuda_ok(cudaMemcpyAsync(occluded + off, (uint8_t*)b.stage_occ + off, m, cudaMemcpyDeviceToHost, st), "D2H occ") ? (int)PT_OK : (int)PT_ERR_CUDA;
        });
    return prc != PT_OK ? prc : check_overflow(s, "pt_trace_any");
}

int pt_trace_closest_dev(pt_scene* s, const void* d_rays, size_t n, void* d_hits, void* d_inst, void* stream){
    PT_CHECK_READY(s);
    if(n && (!d_rays || !d_hits)){ set_error("null buffer"); return PT_ERR_INVALID; }
    return launch_trace_closest(s, d_rays, n, d_hits, d_inst, stream);
}

int pt_trace_any_dev(pt_scene* s, const void* d_rays, size_t n, void* d_occ, void* stream){
    PT_CHECK_READY(s);
    if(n && (!d_rays || !d_occ)){ set_error("null buffer"); return PT_ERR_INVALID; }
    return launch_trace_any(s, d_rays, n, d_occ, stream);
}

int pt_scene_stack_overflows(pt_scene* s, uint64_t* count){
    PT_CHECK_READY(s);
    if(!count){ set_error("null count"); return PT_ERR_INVALID; }
    unsigned long long now = 0;
    const int rc = scene_overflow_count(s, &now);
    if(rc == PT_OK)*count = now;
    return rc;
}

int pt_scene_set_counting(pt_scene* s, int on){
    PT_CHECK_READY(s);
    if(on && s->dev.width != 4){ set_error("counted renders need bvh_width 4"); return PT_ERR_INVALID; }
    s->counting = on != 0;
    if(!cuda_ok(cudaDeviceSynchronize(), "sync") || !cuda_ok(cudaMemset((unsigned long long*)s->buf.stats + 16, 0, 128), "counters clear"))return PT_ERR_CUDA;
    return PT_OK;
}

int pt_scene_render_counts(pt_scene* s, pt_trace_stats* closest, pt_trace_stats* shadow){
    PT_CHECK_READY(s);
    if(!closest || !shadow){ set_error("null stats"); return PT_ERR_INVALID; }
    unsigned long long h[16];
    if(!cuda_ok(cudaDeviceSynchronize(), "sync") || !cuda_ok(cudaMemcpy(h, (unsigned long long*)s->buf.stats + 16, 128, cudaMemcpyDeviceToHost), "D2H counters") ||
       !cuda_ok(cudaMemset((unsigned long long*)s->buf.stats + 16, 0, 128), "counters clear"))return PT_ERR_CUDA;
    pt_trace_stats* out[2] = { closest, shadow };
    for(int k = 0; k < 2; k++){
        const unsigned long long* c = h + 8 * k;
        out[k]->rays = 0; out[k]->tlas_nodes = c[0]; out[k]->blas_nodes = c[1]; out[k]->instances = c[2]; out[k]->triangles = c[3]; out[k]->hits = c[4];
    }
    return PT_OK;
}

int pt_trace_closest_stats(pt_scene* s, const void* d_rays, size_t n, pt_trace_stats* out){
    PT_CHECK_READY(s);
    if(!out || (n && !d_rays)){ set_error("null buffer"); return PT_ERR_INVALID; }
    cudaStream_t st = (cudaStream_t)s->stream;
    if(!cuda_ok(cudaMemsetAsync(s->buf.stats, 0, 64, st), "memset"))return PT_ERR_CUDA;
    int rc = launch_trace_stats(s, d_rays, n, s->buf.stats, st);
    if(rc != PT_OK)return rc;
    unsigned long long h[8];
    if(!cuda_ok(cudaMemcpyAsync(h, s->buf.stats, 64, cudaMemcpyDeviceToHost, st), "D2H stats") || !cuda_ok(cudaStreamSynchronize(st), "sync"))return PT_ERR_CUDA;
    out->rays = n; out->tlas_nodes = h[0]; out->blas_nodes = h[1]; out->instances = h[2]; out->triangles = h[3]; out->hits = h[4];
    return PT_OK;
}

// ---- rendering ---------------------------------------------------------------------

int pt_render(pt_scene* s, const pt_camera_desc* cam, const pt_render_params* rp, int clear, pt_render_stats* stats){
    PT_CHECK_READY(s);
    if(!cam || !rp){ set_error("null desc"); return PT_ERR_INVALID; }
    return render_passes(s, cam, rp, clear, stats);
}

int pt_film_device_ptr(pt_scene* s, void** d_film, size_t* n_doubles){
    PT_CHECK_READY(s);
    if(!s->buf.film){ set_error("no film yet: call pt_render first"); return PT_ERR_STATE; }
    if(d_film)*d_film = s->buf.film;
    if(n_doubles)*n_doubles = 4 * (size_t)s->film_w * s->film_h;
    return PT_OK;
}

int pt_film_download(pt_scene* s, double* rgbw){
    PT_CHECK_READY(s);
    if(!s->buf.film || !rgbw){ set_error("no film"); return PT_ERR_STATE; }
    if(!cuda_ok(cudaDeviceSynchronize(), "sync"))return PT_ERR_CUDA;
    if(!cuda_ok(cudaMemcpy(rgbw, s->buf.film, sizeof(double) * 4 * (size_t)s->film_w * s->film_h, cudaMemcpyDeviceToHost), "D2H film"))return PT_ERR_CUDA;
    return PT_OK;
}

int pt_film_download_sumsq(pt_scene* s, double* sumsq){
    PT_CHECK_READY(s);
    if(!s->buf.film_sq || !sumsq){ set_error("no film"); return PT_ERR_STATE; }
    if(!cuda_ok(cudaDeviceSynchronize(), "sync"))return PT_ERR_CUDA;
    if(!cuda_ok(cudaMemcpy(sumsq, s->buf.film_sq, sizeof(double) * 3 * (size_t)s->film_w * s->film_h, cudaMemcpyDeviceToHost), "D2H film_sq"))return PT_ERR_CUDA;
    return PT_OK;
}

int pt_film_upload(pt_scene* s, const double* rgbw, int width, int height){
    PT_CHECK_READY(s);
    if(!rgbw || width <= 0 || height <= 0){ set_error("bad film"); return PT_ERR_INVALID; }
    if(s->buf.film && (s->film_w != width || s->film_h != height)){ cudaFree(s->buf.film); s->buf.film = nullptr; }
    if(!s->buf.film){
        if(s->buf.film_sq){ cudaFree(s->buf.film_sq); s->buf.film_sq = nullptr; }
        if(!cuda_ok(cudaMalloc(&s->buf.film, sizeof(double) * 4 * (size_t)width * height), "cudaMalloc film") ||
           !cuda_ok(cudaMalloc(&s->buf.film_sq, sizeof(double) * 3 * (size_t)width * height), "cudaMalloc film_sq") ||
           !cuda_ok(cudaMemset(s->buf.film_sq, 0, sizeof(double) * 3 * (size_t)width * height), "memset film_sq"))return PT_ERR_CUDA;
    }
    s->film_w = width; s->film_h = height;
    if(!cuda_ok(cudaMemcpy(s->buf.film, rgbw, sizeof(double) * 4 * (size_t)width * height, cudaMemcpyHostToDevice), "H2D film"))return PT_ERR_CUDA;
    return PT_OK;
}

} // extern "C"
