// multi.cu — multi-GPU behind the C-ABI (SURVEY.md §8e, §8b pt_multi_render).
//
// The reference's Integrator::Render(threadCount) (Integrators.hpp:13, Integrators.cpp:23-41,117-123) owns its worker
// threads, every worker adds its tiles into ONE film (Film::Merge, Film.hpp:125-132, atomic<double> fetch_add).
// Here a worker is a GPU: the committed scene is replicated device to device (no second build), rank r of G renders
// the sample passes r, r + G, ... of the frame into its own film (streams of random numbers are keyed by sample id, so
// the sum does not depend on G), and ONE reduce(sum) of the fp64 film accumulators brings the frame to rank 0 — the
// only collective of the path (plus one all-reduce of the per-pixel estimator sums per adaptive round).
//
//   pt_multi_*  : one process drives G GPUs, one host thread per GPU (ncclCommInitAll).
//   pt_dist_*   : one process per GPU (torchrun / mpirun); the caller carries the 128-byte NCCL id from rank 0 to the
//                 others (ncclCommInitRank).
//
// NCCL is bound at run time (dlopen, the copy already in the process if there is one: a process that also holds
// PyTorch must not end up with two NCCLs), and only when a communicator over more than one device is asked for.
// Ranks that share a device ("virtual ranks", the 1-GPU test of the sharding logic) are reduced by a plain add
// kernel instead: NCCL refuses two ranks on one GPU.
#include <cuda_runtime.h>
#include <nccl.h>
#include <dlfcn.h>

#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "scene_host.hpp"

namespace ptb {

namespace {

struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string where;
};

NcclApi* nccl_api(){
    static NcclApi api;
    static std::once_flag once;
    static bool ok = false;
    std::call_once(once, []{
        // a copy that is already mapped (e.g. PyTorch's) wins: two NCCLs in one process do not share a bootstrap
        const char* names[] = { "libnccl.so.2", "libnccl.so" };
        for(const char* n : names){ api.lib = dlopen(n, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL); if(api.lib){ api.where = std::string(n) + " (already loaded)"; break; } }
        if(!api.lib)for(const char* n : names){ api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if(api.lib){ api.where = n; break; } }
        if(!api.lib)return;
        bool all = true;
        auto sym = [&](const char* name){ void* p = dlsym(api.lib, name); if(!p)all = false; return p; };
        api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
        api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
        api.CommInitAll = (decltype(api.CommInitAll))sym("ncclCommInitAll");
        api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
        api.Reduce = (decltype(api.Reduce))sym("ncclReduce");
        api.AllReduce = (decltype(api.AllReduce))sym("ncclAllReduce");
        api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
        api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
        api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
        ok = all;
    });
    if(!ok){ set_error("NCCL is not available: libnccl.so.2 could not be loaded (multi-GPU rendering needs it)"); return nullptr; }
    return &api;
}

bool nccl_ok(NcclApi* N, ncclResult_t r, const char* what){
    if(r == ncclSuccess)return true;
    set_error(std::string(what) + ": " + (N && N->GetErrorString ? N->GetErrorString(r) : "NCCL error"));
    return false;
}

// dst += src (virtual ranks on one device; also the stand-in when NCCL is absent and every rank sits on device 0)
__global__ void k_add_f64(double* __restrict__ dst, const double* __restrict__ src, size_t n){
    for(size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)dst[i] += src[i];
}

// pt_allreduce_fn of a scene that owns an NCCL communicator (adaptive rounds, render.cu merge_round)
int nccl_allreduce_cb(void* user, void* d_buf, size_t n_doubles, void* stream){
    pt_scene* s = (pt_scene*)user;
    NcclApi* N = nccl_api();
    if(!N || !s->comm)return 1;
    return nccl_ok(N, N->AllReduce(d_buf, d_buf, n_doubles, ncclFloat64, ncclSum, (ncclComm_t)s->comm, (cudaStream_t)stream), "ncclAllReduce") ? 0 : 1;
}

// film (4 W H doubles) and per-pixel sums of squares (3 W H) of this rank to `root`, on the scene's stream
int reduce_film_nccl(pt_scene* s, int root){
    NcclApi* N = nccl_api();
    if(!N)return PT_ERR_STATE;
    const size_t npix = (size_t)s->film_w * s->film_h;
    cudaStream_t st = (cudaStream_t)s->stream;
    ncclComm_t c = (ncclComm_t)s->comm;
    if(!nccl_ok(N, N->GroupStart(), "ncclGroupStart") ||
       !nccl_ok(N, N->Reduce(s->buf.film, s->buf.film, 4 * npix, ncclFloat64, ncclSum, root, c, st), "ncclReduce(film)") ||
       !nccl_ok(N, N->Reduce(s->buf.film_sq, s->buf.film_sq, 3 * npix, ncclFloat64, ncclSum, root, c, st), "ncclReduce(film_sq)") ||
       !nccl_ok(N, N->GroupEnd(), "ncclGroupEnd"))return PT_ERR_CUDA;
    return PT_OK;
}

pt_render_params shard_params(const pt_render_params& rp, int rank, int world){
    pt_render_params q = rp;
    q.pass_begin = rp.pass_begin + rank * rp.pass_stride;
    q.pass_stride = rp.pass_stride * world;
    return q;
}

// Passes are the unit of sharding while there are enough of them: every rank traces whole frames, perfectly balanced.  A frame
// with fewer passes than ranks would leave GPUs idle; it is split by interleaved 32 x 32 tiles instead.
// PTB_SHARD=tiles|passes forces one.  Adaptive frames always shard passes (the stopping rule needs every pixel on every rank).
bool shard_by_tiles(const pt_render_params& rp, int world){
    if(world <= 1 || rp.adaptive)return false;
    if(const char* e = getenv("PTB_SHARD"))return e[0] == 't';
    const int passes = rp.pass_begin >= rp.spp ? 0 : (rp.spp - rp.pass_begin + rp.pass_stride - 1) / rp.pass_stride;
    return passes < world;
}

// one rank's share of a frame, by passes or by tiles
int render_share(pt_scene* s, const pt_camera_desc* cam, const pt_render_params* rp, int rank, int world, int clear, pt_render_stats* st){
    if(shard_by_tiles(*rp, world)){
        s->tile_rank = rank; s->tile_world = world;
        const int rc = render_passes(s, cam, rp, clear, st);
        s->tile_rank = 0; s->tile_world = 1;
        return rc;
    }
    const pt_render_params q = shard_params(*rp, rank, world);
    return render_passes(s, cam, &q, clear, st);
}

} // namespace

// ---- replica of a committed scene on another device ---------------------------------------------------------------
int scene_replicate(pt_scene* src, int device, pt_scene** out){
    if(!src || !out){ set_error("null scene"); return PT_ERR_INVALID; }
    if(!src->committed){ set_error("scene not committed"); return PT_ERR_STATE; }
    int nd = 0;
    if(cudaGetDeviceCount(&nd) != cudaSuccess || device < 0 || device >= nd){ cudaGetLastError(); set_error("bad device index"); return PT_ERR_INVALID; }
    pt_scene* r = new (std::nothrow) pt_scene();
    if(!r){ set_error("out of memory"); return PT_ERR_NOMEM; }
    r->device = device; r->width = src->width; r->light_sampler = src->light_sampler;
    r->materials = src->materials; r->media = src->media; r->lights = src->lights;     // small POD lists the host-side checks look at
    r->scene_medium = src->scene_medium; r->bounds = src->bounds;
    r->dev = src->dev; r->shade = src->shade;                                          // scalars; pointers are re-bound below
    r->total_tris = src->total_tris; r->total_nodes = src->total_nodes; r->n_inst_records = src->n_inst_records; r->n_shade_vertices = src->n_shade_vertices; r->n_shade_indices = src->n_shade_indices; r->n_shade_tris = src->n_shade_tris;
    r->has_wrappers = src->has_wrappers; r->has_alpha_tex = src->has_alpha_tex; r->has_boundary = src->has_boundary;
    r->builder = src->builder; r->build_stats = src->build_stats;
    if(!cuda_ok(cudaSetDevice(device), "cudaSetDevice")){ delete r; return PT_ERR_CUDA; }
    bool ok = true;
    for(const auto& u : src->uploads){
        void** dst = (void**)((char*)&r->buf + u.first);
        void* const* from = (void* const*)((const char*)&src->buf + u.first);
        if(!cuda_ok(cudaMalloc(dst, u.second), "cudaMalloc replica")){ ok = false; break; }
        // device to device: over NVLink where peers are connected, staged through the host by the runtime otherwise
        if(!cuda_ok(cudaMemcpyPeer(*dst, device, *from, src->device, u.second), "cudaMemcpyPeer replica")){ ok = false; break; }
        note_upload(r, dst, u.second);
    }
    cudaStream_t st[4] = { nullptr, nullptr, nullptr, nullptr };
    for(int k = 0; ok && k < 4; k++)ok = cuda_ok(cudaStreamCreateWithFlags(&st[k], cudaStreamNonBlocking), "cudaStreamCreate");
    r->stream = st[0]; r->stream2 = st[1]; r->stream3 = st[2]; r->stream4 = st[3];
    ok = ok && cuda_ok(cudaMalloc(&r->buf.stats, 256), "cudaMalloc stats") && cuda_ok(cudaMemset(r->buf.stats, 0, 256), "stats clear") &&
         cuda_ok(cudaMalloc(&r->buf.cursors, kCursorRing * sizeof(unsigned long long)), "cudaMalloc cursors") && cuda_ok(cudaDeviceSynchronize(), "replica");
    if(!ok){
        scene_free_device(r);
        for(cudaStream_t q : st)if(q)cudaStreamDestroy(q);
        delete r;
        return PT_ERR_CUDA;
    }
    bind_device_pointers(r);
    r->committed = true;
    *out = r;
    return PT_OK;
}

} // namespace ptb

using namespace ptb;

// ---- one process, G GPUs ----------------------------------------------------------------------------------------------
struct pt_multi {
    std::vector<pt_scene*> ranks;          // ranks[0] = the primary (not owned), the others are replicas (owned)
    std::vector<ncclComm_t> comms;         // empty when every rank shares one device
    bool virtual_ranks = false;

    struct Worker {
        std::thread th;
        std::function<int()> job;
        bool has_job = false, stop = false;
        int rc = 0;
        std::string err;
    };
    std::vector<Worker> workers;
    std::mutex mu;
    std::condition_variable cv_job, cv_done;
    int pending = 0;

    void loop(int r){
        Worker& w = workers[r];
        cudaSetDevice(ranks[r]->device);
        for(;;){
            std::function<int()> job;
            {
                std::unique_lock<std::mutex> lk(mu);
                cv_job.wait(lk, [&]{ return w.has_job || w.stop; });
                if(w.stop)return;
                job = w.job;
            }
            const int rc = job();
            const std::string err = rc != PT_OK ? get_error() : "";
            {
                std::lock_guard<std::mutex> lk(mu);
                w.rc = rc; w.err = err; w.has_job = false;
                if(--pending == 0)cv_done.notify_all();
            }
        }
    }
    // run fn(rank) on every rank's thread, wait for all; the first failure is reported
    int run(const std::function<int(int)>& fn){
        {
            std::lock_guard<std::mutex> lk(mu);
            pending = (int)workers.size();
            for(size_t r = 0; r < workers.size(); r++){ workers[r].job = [fn, r]{ return fn((int)r); }; workers[r].has_job = true; }
        }
        cv_job.notify_all();
        std::unique_lock<std::mutex> lk(mu);
        cv_done.wait(lk, [&]{ return pending == 0; });
        for(auto& w : workers)if(w.rc != PT_OK){ set_error(w.err); return w.rc; }
        return PT_OK;
    }
};

extern "C" {

int pt_scene_replicate(pt_scene* src, int device, pt_scene** out){ return scene_replicate(src, device, out); }

pt_multi* pt_multi_create(pt_scene* primary, const int* devices, int n_devices){
    if(!primary || !primary->committed){ set_error("pt_multi_create: the primary scene must be committed"); return nullptr; }
    int nd = 0;
    if(cudaGetDeviceCount(&nd) != cudaSuccess || nd <= 0){ cudaGetLastError(); set_error("no CUDA device"); return nullptr; }
    std::vector<int> devs;
    if(devices && n_devices > 0)devs.assign(devices, devices + n_devices);
    else{ devs.push_back(primary->device); for(int d = 0; d < nd; d++)if(d != primary->device)devs.push_back(d); }
    if(devs[0] != primary->device){ set_error("pt_multi_create: devices[0] must be the primary scene's device"); return nullptr; }
    if(devs.size() > 64){ set_error("pt_multi_create: at most 64 ranks"); return nullptr; }
    bool distinct = true, same = true;
    for(size_t i = 0; i < devs.size(); i++){
        if(devs[i] < 0 || devs[i] >= nd){ set_error("pt_multi_create: bad device index"); return nullptr; }
        if(devs[i] != devs[0])same = false;
        for(size_t j = 0; j < i; j++)if(devs[i] == devs[j])distinct = false;
    }
    if(devs.size() > 1 && !distinct && !same){ set_error("pt_multi_create: devices must be all different (NCCL) or all the same (virtual ranks on one GPU)"); return nullptr; }
    pt_multi* m = new (std::nothrow) pt_multi();
    if(!m){ set_error("out of memory"); return nullptr; }
    m->virtual_ranks = devs.size() > 1 && same;
    m->ranks.push_back(primary);
    auto destroy = [&](){ for(size_t r = 1; r < m->ranks.size(); r++)pt_scene_destroy(m->ranks[r]); delete m; return (pt_multi*)nullptr; };
    for(size_t r = 1; r < devs.size(); r++){
        pt_scene* rep = nullptr;
        if(scene_replicate(primary, devs[r], &rep) != PT_OK)return destroy();
        m->ranks.push_back(rep);
    }
    if(devs.size() > 1 && distinct){
        NcclApi* N = nccl_api();
        if(!N)return destroy();
        m->comms.resize(devs.size());
        if(!nccl_ok(N, N->CommInitAll(m->comms.data(), (int)devs.size(), devs.data()), "ncclCommInitAll")){ m->comms.clear(); return destroy(); }
        for(size_t r = 0; r < devs.size(); r++){
            pt_scene* s = m->ranks[r];
            s->comm = m->comms[r]; s->comm_rank = (int)r; s->comm_world = (int)devs.size(); s->comm_owned = false;
        }
    }
    cudaSetDevice(primary->device);
    m->workers = std::vector<pt_multi::Worker>(devs.size());
    for(size_t r = 0; r < devs.size(); r++)m->workers[r].th = std::thread([m, r]{ m->loop((int)r); });
    return m;
}

void pt_multi_destroy(pt_multi* m){
    if(!m)return;
    {
        std::lock_guard<std::mutex> lk(m->mu);
        for(auto& w : m->workers)w.stop = true;
    }
    m->cv_job.notify_all();
    for(auto& w : m->workers)if(w.th.joinable())w.th.join();
    NcclApi* N = m->comms.empty() ? nullptr : nccl_api();
    for(size_t r = 0; r < m->ranks.size(); r++){
        pt_scene* s = m->ranks[r];
        if(N && r < m->comms.size()){ cudaSetDevice(s->device); cudaDeviceSynchronize(); N->CommDestroy(m->comms[r]); }
        s->comm = nullptr; s->comm_rank = 0; s->comm_world = 1;
        if(r > 0)pt_scene_destroy(s);
    }
    if(!m->ranks.empty())cudaSetDevice(m->ranks[0]->device);
    delete m;
}

int pt_multi_gpus(const pt_multi* m){ return m ? (int)m->ranks.size() : 0; }

pt_scene* pt_multi_scene(pt_multi* m, int rank){
    if(!m || rank < 0 || rank >= (int)m->ranks.size()){ set_error("pt_multi_scene: bad rank"); return nullptr; }
    return m->ranks[rank];
}

// Integrator::Render over all ranks: the frame ends up in the primary scene's film (pt_film_download(primary)).
int pt_multi_render(pt_multi* m, const pt_camera_desc* cam, const pt_render_params* rp, int clear, pt_render_stats* stats){
    if(!m || !cam || !rp){ set_error("null argument"); return PT_ERR_INVALID; }
    const int G = (int)m->ranks.size();
    std::vector<pt_render_stats> st(G);
    const bool sharded_adaptive = rp->adaptive && G > 1;
    if(sharded_adaptive && m->virtual_ranks){ set_error("adaptive sampling over virtual ranks (several ranks on one GPU) is not supported: the per-round all-reduce needs NCCL"); return PT_ERR_INVALID; }
    int rc = m->run([&](int r) -> int{
        pt_scene* s = m->ranks[r];
        if(!cuda_ok(cudaSetDevice(s->device), "cudaSetDevice"))return PT_ERR_CUDA;
        if(sharded_adaptive){ s->allreduce = nccl_allreduce_cb; s->allreduce_user = s; }
        // every rank but the primary starts from an empty film: the reduce adds them onto whatever the primary keeps
        int rc = render_share(s, cam, rp, r, G, r == 0 ? clear : 1, &st[r]);
        if(rc != PT_OK)return rc;
        if(G > 1 && !m->virtual_ranks){
            rc = reduce_film_nccl(s, 0);
            if(rc != PT_OK)return rc;
            if(!cuda_ok(cudaStreamSynchronize((cudaStream_t)s->stream), "film reduce"))return PT_ERR_CUDA;
        }
        return PT_OK;
    });
    if(rc != PT_OK)return rc;
    if(m->virtual_ranks){
        pt_scene* p = m->ranks[0];
        if(!cuda_ok(cudaSetDevice(p->device), "cudaSetDevice"))return PT_ERR_CUDA;
        const size_t npix = (size_t)p->film_w * p->film_h;
        cudaStream_t ps = (cudaStream_t)p->stream;
        for(int r = 1; r < G; r++){
            k_add_f64<<<1184, 256, 0, ps>>>((double*)p->buf.film, (const double*)m->ranks[r]->buf.film, 4 * npix);
            k_add_f64<<<1184, 256, 0, ps>>>((double*)p->buf.film_sq, (const double*)m->ranks[r]->buf.film_sq, 3 * npix);
        }
        if(!cuda_ok(cudaStreamSynchronize(ps), "film add"))return PT_ERR_CUDA;
    }
    if(stats){
        pt_render_stats t {};
        for(const pt_render_stats& x : st){
            t.samples += x.samples; t.closest_rays += x.closest_rays; t.shadow_rays += x.shadow_rays; t.kernel_launches += x.kernel_launches;
            t.nonfinite_samples += x.nonfinite_samples; if(x.gpu_ms > t.gpu_ms)t.gpu_ms = x.gpu_ms;
        }
        *stats = t;
    }
    return PT_OK;
}

// ---- one process per GPU ---------------------------------------------------------------------------------------------
int pt_dist_unique_id(void* id128){
    if(!id128){ set_error("null id"); return PT_ERR_INVALID; }
    NcclApi* N = nccl_api();
    if(!N)return PT_ERR_STATE;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    if(!nccl_ok(N, N->GetUniqueId(&id), "ncclGetUniqueId"))return PT_ERR_CUDA;
    memcpy(id128, &id, 128);
    return PT_OK;
}

int pt_dist_init(pt_scene* s, const void* id128, int rank, int world){
    if(!s || !s->committed){ set_error("pt_dist_init: scene not committed"); return PT_ERR_STATE; }
    if(world < 1 || rank < 0 || rank >= world || (world > 1 && !id128)){ set_error("pt_dist_init: bad rank / world / id"); return PT_ERR_INVALID; }
    if(s->comm){ set_error("pt_dist_init: the scene already belongs to a communicator"); return PT_ERR_STATE; }
    s->comm_rank = rank; s->comm_world = world;
    if(world == 1)return PT_OK;
    NcclApi* N = nccl_api();
    if(!N)return PT_ERR_STATE;
    if(!cuda_ok(cudaSetDevice(s->device), "cudaSetDevice"))return PT_ERR_CUDA;
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    ncclComm_t c = nullptr;
    if(!nccl_ok(N, N->CommInitRank(&c, world, id, rank), "ncclCommInitRank")){ s->comm_rank = 0; s->comm_world = 1; return PT_ERR_CUDA; }
    s->comm = c; s->comm_owned = true;
    return PT_OK;
}

int pt_dist_finalize(pt_scene* s){
    if(!s)return PT_OK;
    if(s->comm && s->comm_owned){
        NcclApi* N = nccl_api();
        cudaSetDevice(s->device);
        cudaDeviceSynchronize();
        if(N)N->CommDestroy((ncclComm_t)s->comm);
    }
    s->comm = nullptr; s->comm_owned = false; s->comm_rank = 0; s->comm_world = 1;
    return PT_OK;
}

// This rank's share of the frame (passes rank, rank + world, ...) and the one reduce of the film to rank 0.
int pt_dist_render(pt_scene* s, const pt_camera_desc* cam, const pt_render_params* rp, int clear, pt_render_stats* stats){
    if(!s || !cam || !rp){ set_error("null argument"); return PT_ERR_INVALID; }
    if(!s->committed){ set_error("scene not committed"); return PT_ERR_STATE; }
    if(!cuda_ok(cudaSetDevice(s->device), "cudaSetDevice"))return PT_ERR_CUDA;
    const int G = s->comm_world, r = s->comm_rank;
    if(G > 1 && !s->comm){ set_error("pt_dist_render: call pt_dist_init first"); return PT_ERR_STATE; }
    if(G > 1 && rp->adaptive){ s->allreduce = nccl_allreduce_cb; s->allreduce_user = s; }
    int rc = render_share(s, cam, rp, r, G, (r == 0 ? clear : 1), stats);
    if(rc != PT_OK)return rc;
    if(G > 1){
        rc = reduce_film_nccl(s, 0);
        if(rc != PT_OK)return rc;
        if(!cuda_ok(cudaStreamSynchronize((cudaStream_t)s->stream), "film reduce"))return PT_ERR_CUDA;
    }
    return PT_OK;
}

} // extern "C"
