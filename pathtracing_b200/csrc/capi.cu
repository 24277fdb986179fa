// capi.cu — the extern "C" surface declared in include/pt_b200.h.
#include <cuda_runtime.h>
#include <cstring>
#include <new>

#include "scene_host.hpp"

using namespace ptb;

#define PT_CHECK_SCENE(s) do{ if(!(s)){ set_error("null scene"); return PT_ERR_INVALID; } }while(0)
#define PT_CHECK_OPEN(s) do{ PT_CHECK_SCENE(s); if((s)->committed){ set_error("scene already committed"); return PT_ERR_STATE; } }while(0)
#define PT_CHECK_READY(s) do{ PT_CHECK_SCENE(s); if(!(s)->committed){ set_error("scene not committed"); return PT_ERR_STATE; } \
    if(!cuda_ok(cudaSetDevice((s)->device), "cudaSetDevice"))return PT_ERR_CUDA; }while(0)

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
    return s;
}

void pt_scene_destroy(pt_scene* s){
    if(!s)return;
    if(s->committed || s->buf.n2 || s->buf.n4){
        cudaSetDevice(s->device);
        cudaDeviceSynchronize();
        scene_free_device(s);
        if(s->stream)cudaStreamDestroy((cudaStream_t)s->stream);
        if(s->stream2)cudaStreamDestroy((cudaStream_t)s->stream2);
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

int pt_scene_bounds(const pt_scene* s, float bmin[3], float bmax[3]){
    PT_CHECK_SCENE(s);
    if(!s->committed){ set_error("scene not committed"); return PT_ERR_STATE; }
    for(int a = 0; a < 3; a++){ bmin[a] = s->bounds.lo[a]; bmax[a] = s->bounds.hi[a]; }
    return PT_OK;
}

// ---- ray casting -------------------------------------------------------------------

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

// Host-buffer entry points.  The batch is cut into chunks that ping-pong over
// two streams so that the H2D copy of chunk k+1, the kernel of chunk k and the
// D2H copy of chunk k-1 overlap (full-duplex PCIe + compute).
static const size_t kChunk = 1u << 20;

int pt_trace_closest(pt_scene* s, const pt_ray* rays, size_t n, pt_hit* hits, uint32_t* inst){
    PT_CHECK_READY(s);
    if(n == 0)return PT_OK;
    if(!rays || !hits){ set_error("null buffer"); return PT_ERR_INVALID; }
    int rc = ensure_stage(s, n);
    if(rc != PT_OK)return rc;
    DeviceBuffers& b = s->buf;
    cudaStream_t st[2] = { (cudaStream_t)s->stream, (cudaStream_t)s->stream2 };
    int k = 0;
    for(size_t off = 0; off < n; off += kChunk, k ^= 1){
        const size_t m = n - off < kChunk ? n - off : kChunk;
        pt_ray* dr = (pt_ray*)b.stage_rays + off;
        pt_hit* dh = (pt_hit*)b.stage_hits + off;
        uint32_t* di = (uint32_t*)b.stage_inst + off;
        if(!cuda_ok(cudaMemcpyAsync(dr, rays + off, m * sizeof(pt_ray), cudaMemcpyHostToDevice, st[k]), "H2D rays"))return PT_ERR_CUDA;
        rc = launch_trace_closest(s, dr, m, dh, inst ? di : nullptr, st[k]);
        if(rc != PT_OK)return rc;
        if(!cuda_ok(cudaMemcpyAsync(hits + off, dh, m * sizeof(pt_hit), cudaMemcpyDeviceToHost, st[k]), "D2H hits"))return PT_ERR_CUDA;
        if(inst && !cuda_ok(cudaMemcpyAsync(inst + off, di, m * sizeof(uint32_t), cudaMemcpyDeviceToHost, st[k]), "D2H inst"))return PT_ERR_CUDA;
    }
    if(!cuda_ok(cudaStreamSynchronize(st[0]), "sync") || !cuda_ok(cudaStreamSynchronize(st[1]), "sync"))return PT_ERR_CUDA;
    return PT_OK;
}

int pt_trace_any(pt_scene* s, const pt_ray* rays, size_t n, uint8_t* occluded){
    PT_CHECK_READY(s);
    if(n == 0)return PT_OK;
    if(!rays || !occluded){ set_error("null buffer"); return PT_ERR_INVALID; }
    int rc = ensure_stage(s, n);
    if(rc != PT_OK)return rc;
    DeviceBuffers& b = s->buf;
    cudaStream_t st[2] = { (cudaStream_t)s->stream, (cudaStream_t)s->stream2 };
    int k = 0;
    for(size_t off = 0; off < n; off += kChunk, k ^= 1){
        const size_t m = n - off < kChunk ? n - off : kChunk;
        pt_ray* dr = (pt_ray*)b.stage_rays + off;
        uint8_t* dc = (uint8_t*)b.stage_occ + off;
        if(!cuda_ok(cudaMemcpyAsync(dr, rays + off, m * sizeof(pt_ray), cudaMemcpyHostToDevice, st[k]), "H2D rays"))return PT_ERR_CUDA;
        rc = launch_trace_any(s, dr, m, dc, st[k]);
        if(rc != PT_OK)return rc;
        if(!cuda_ok(cudaMemcpyAsync(occluded + off, dc, m, cudaMemcpyDeviceToHost, st[k]), "D2H occ"))return PT_ERR_CUDA;
    }
    if(!cuda_ok(cudaStreamSynchronize(st[0]), "sync") || !cuda_ok(cudaStreamSynchronize(st[1]), "sync"))return PT_ERR_CUDA;
    return PT_OK;
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
