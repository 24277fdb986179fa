// trace.cu — ray-cast kernels: batched Scene::Intersect / Scene::IntersectPred
// (Scene.cpp:4-6,31-33) for BVH2 and BVH4.
//
// Persistent warps: the grid is (#SMs x resident CTAs) and never larger; every
// warp pulls chunks of kChunk consecutive rays from a global cursor and keeps
// its 32 lanes busy by handing the next ray of the chunk to a lane the moment
// its previous ray finishes (trace.cuh: one trav_step round per loop
// iteration).  Rays are read as two 128-bit loads, hits written as one 128-bit
// store.  The first 16 stack entries of every lane live in shared memory (128 KB per
// SM at 8 CTAs), deeper ones spill to local memory; the remaining ~90 KB stay L1 for
// the upper tree levels (PTB_STACK_SMEM=0|8|16|24 selects the split; 16 measured best).
#include <cstdio>
#include <cstdlib>
#include <string>
#include "scene_host.hpp"
#include "trace.cuh"
#include "alpha.cuh"

namespace ptb {

constexpr int kTraceBlock = 128;
constexpr unsigned kChunk = 64;        // rays per cursor grab (one atomic per 64 rays)

// Warp-level ray dispenser.  base/avail are warp-uniform.
struct RayFeed {
    unsigned long long base;   // next ray of the current chunk
    unsigned avail;            // rays left in the current chunk
    bool exhausted;
};

// FEED: the rays are still arriving (host-buffer pipeline, StreamCtl in scene_host.hpp): a grabbed chunk is waited
// for until the H2D watermark covers it.  A wait of about a second means the copy is not coming (a host that blocks in the launch call, e.g. under a
// profiler or CUDA_LAUNCH_BLOCKING; the caller then falls back to one kernel per chunk): raise the abort flag
// and let every warp run dry instead of hanging the device.
__device__ __forceinline__ unsigned long long global_ns(){
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long* p){
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

template <bool FEED>
__device__ __forceinline__ void feed_grab(RayFeed& f, unsigned long long* cursor, size_t n, unsigned lane, const StreamCtl& ctl){
    unsigned long long b = 0;
    if(lane == 0){
        b = atomicAdd(cursor, (unsigned long long)kChunk);
        if(FEED && b < n){
            const unsigned long long need = b + kChunk < n ? b + kChunk : (unsigned long long)n;
            // acquire loads of the watermark: the rays below it were written (by the copy engine) before it moved.
            // The wait is bounded in wall time (%globaltimer, nanoseconds), not in SM clocks.
            const unsigned long long t0 = global_ns();
            while(ld_acquire_u64(ctl.landed) < need){
                if(*(volatile const unsigned*)ctl.abort != 0u || global_ns() - t0 > 1000000000ull){ atomicExch(ctl.abort, 1u); b = n; break; }
                __nanosleep(100);
            }
            __threadfence();
        }
    }
    b = __shfl_sync(0xffffffffu, b, 0);
    if(b >= n){ f.exhausted = true; f.avail = 0; return; }
    f.base = b;
    const unsigned long long left = n - b;
    f.avail = left < kChunk ? (unsigned)left : kChunk;
}

// dynamic shared memory of a k_trace CTA: KS stack rows of 8 B per lane + the nine world-ray rows of 4 B per lane (trace.cuh)
static constexpr size_t trace_smem_bytes(int ks){ return ks > 0 ? (size_t)ks * kTraceBlock * sizeof(uint2) + (size_t)kWorldRayRows * kTraceBlock * sizeof(float) : 0; }
// COMPACT (with FEED, closest hit only): rays are 24 B {o, d} records, results 8 B {t, prim} (pt_trace_closest_compact)
// resident CTAs the register allocation has to leave room for: the plain fast-layout kernels ran at 56 registers (9 CTAs of 128 threads per SM)
// until the instance world-box test took them to 64 (8 CTAs: C2 1631 -> 1516 Mrays/s); everything else takes what it needs
#ifndef PTB_SINGLE_CTAS
#define PTB_SINGLE_CTAS 9       // 10 (48 registers) compiles without spills and runs at 1000 instead of 1654 Mrays/s
#endif
template <int WIDTH, bool STATS, class Alpha>
constexpr int trace_min_ctas(){ return (WIDTH == kWidthFast4 && !STATS && !Alpha::kWrap) ? (Alpha::kSingle ? PTB_SINGLE_CTAS : 9) : 1; }

template <int WIDTH, bool ANY, bool STATS, int KS, class Alpha, bool FEED = false, bool COMPACT = false>
__global__ void __launch_bounds__(kTraceBlock, trace_min_ctas<WIDTH, STATS, Alpha>()) k_trace(DevScene S, const float4* __restrict__ rays, size_t n,
                                                       float4* __restrict__ hits, uint32_t* __restrict__ inst, uint8_t* __restrict__ occ,
                                                       unsigned long long* __restrict__ cursor, unsigned long long* __restrict__ stats,
                                                       unsigned refill_min, const Alpha alpha, const StreamCtl ctl, const DynBatch dyn,
                                                       unsigned long long* __restrict__ overflow, const ConnectAcc acc){
    if(dyn.n_dev){
        // wavefront loop: the batch size and which of the two ray queues is current live on the device
        n = *dyn.n_dev;
        if(dyn.sel_dev)rays += (size_t)(*dyn.sel_dev) * dyn.sel_stride;
        if(n == 0)return;
    }
    const unsigned leaf_min = (refill_min >> 8) & 0xFFu;       // packed by the launchers: rounds << 16 | leaf_min << 8 | refill_min
    const int rounds = (int)(refill_min >> 16);                // traversal rounds between two looks at the refill condition
    refill_min &= 0xFFu;
    extern __shared__ uint2 smem_stack[];            // KS rows of kTraceBlock entries
    uint2 spill[kStack > KS ? kStack - KS : 1];
    TravStack<KS, KS ? kTraceBlock : 1> stack;
    if constexpr(KS > 0){
        stack.sm = smem_stack + threadIdx.x;
        stack.wr = reinterpret_cast<float*>(smem_stack + KS * kTraceBlock) + threadIdx.x;     // nine rows of kTraceBlock floats behind the stack rows
    }
    stack.lm = spill;
    TravState s;
    s.flags = TF_DONE;
    size_t idx = (size_t)-1;
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt = (1u << lane) - 1u;
    RayFeed feed = { 0, 0, false };
    TraceCounters c = { 0, 0, 0, 0 };
    unsigned long long hitc = 0;

    for(;;){
        const unsigned need = __ballot_sync(0xffffffffu, (s.flags & TF_DONE) != 0);
        // Refill in batches: the retire + fetch + init path runs with only the finished lanes
        // active, so it is entered once `refill_min` lanes are waiting (or nothing else is left).
        if(need && (__popc(need) >= (int)refill_min || feed.exhausted || need == 0xffffffffu)){
            if(s.flags & TF_DONE){
                if(idx != (size_t)-1){
                    // retire the finished ray
                    if(s.flags & TF_OVERFLOW)atomicAdd(overflow, 1ull);      // never on any scene seen so far: 96 entries
                    if(ANY){
                        if(acc.ps){
                            if(!(s.flags & TF_OCCLUDED)){
                                float4* d = (float4*)acc.ps + ((size_t)acc.slot[idx] * acc.ps_slot + acc.ps_off);
                                const float4 c = ((const float4*)acc.contrib)[idx];
                                float4 r = *d;
                                r.x += c.x; r.y += c.y; r.z += c.z;
                                *d = r;
                            }
                        } else occ[idx] = (s.flags & TF_OCCLUDED) ? 1 : 0;
                    } else{
                        const bool found = s.hprim != 0xFFFFFFFFu;
                        if constexpr(COMPACT)((float2*)hits)[idx] = make_float2(found ? s.tmax : CUDART_INF_F, __uint_as_float(s.hprim));
                        else if(!STATS || hits)hits[idx] = make_float4(found ? s.tmax : CUDART_INF_F, s.hu, s.hv, __uint_as_float(s.hprim));
                        if(inst)inst[idx] = s.hinst;
                        if(STATS && found)hitc++;
                    }
                    if(FEED){
                        // release: this ray's result is written before its granule's count says so
                        // 32 sub-counters per granule, one sector each: every SM works on the same one or two granules
                        // at any time, a single counter would serialise all retirements on one L2 atomic unit
                        unsigned* c = ctl.done + ((((idx >> ctl.shift) << 5) | (blockIdx.x & 31u)) << 3);
#if defined(__CUDA_ARCH__)
                        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" :: "l"(c) : "memory");
#endif
                    }
                    idx = (size_t)-1;
                }
            }
            if(feed.avail == 0 && !feed.exhausted)feed_grab<FEED>(feed, cursor, n, lane, ctl);
            const unsigned rank = __popc(need & lt);
            if((s.flags & TF_DONE) && rank < feed.avail){
                idx = (size_t)(feed.base + rank);
                if constexpr(COMPACT){
                    float4 ra, rb;
                    fetch_ray(CgRays24 { (const float2*)rays }, idx, ra, rb);
                    trav_init<WIDTH>(s, S, ra, rb);
                } else if constexpr(FEED)trav_init<WIDTH>(s, S, __ldcg(rays + 2 * idx), __ldcg(rays + 2 * idx + 1));
                else trav_init<WIDTH>(s, S, __ldg(rays + 2 * idx), __ldg(rays + 2 * idx + 1));
            }
            const unsigned want = __popc(need);
            const unsigned taken = want < feed.avail ? want : feed.avail;
            feed.base += taken; feed.avail -= taken;
            if(feed.exhausted && __all_sync(0xffffffffu, idx == (size_t)-1))break;
        }
        // postponed leaf processing: the leaf code (triangle test, instance entry / exit) runs in a round only when at least
        // leaf_min lanes hold a leaf item, or when no live lane has anything else to do
        bool do_leaf = true;
        if(leaf_min > 1u){
            const bool live = !(s.flags & TF_DONE);
            const bool at_leaf = live && (s.cur & REF_LEAF) && s.cur != REF_EMPTY;
            const unsigned nl = __popc(__ballot_sync(0xffffffffu, at_leaf)), nv = __popc(__ballot_sync(0xffffffffu, live));
            do_leaf = nl >= leaf_min || nl == nv;
        }
#pragma unroll 1
        for(int rep = 0; rep < rounds; rep++){
        if(!(s.flags & TF_DONE)){
            if constexpr(COMPACT)trav_step<WIDTH, ANY, STATS>(s, S, stack, CgRays24 { (const float2*)rays }, idx, STATS ? &c : nullptr, alpha, do_leaf);
            else if constexpr(FEED)trav_step<WIDTH, ANY, STATS>(s, S, stack, CgRays { rays }, idx, STATS ? &c : nullptr, alpha, do_leaf);
            else trav_step<WIDTH, ANY, STATS>(s, S, stack, rays, idx, STATS ? &c : nullptr, alpha, do_leaf);
        }
        }
    }
    if(STATS){
        // warp reduce, one atomic per warp per counter
        unsigned long long v[5] = { c.tlas_nodes, c.blas_nodes, c.instances, c.triangles, hitc };
#pragma unroll
        for(int k = 0; k < 5; k++){
            unsigned long long x = v[k];
            for(int o = 16; o > 0; o >>= 1)x += __shfl_down_sync(0xffffffffu, x, o);
            if(lane == 0 && x)atomicAdd(stats + k, x);
        }
    }
}

static int check_launch(const char* what){
    cudaError_t e = cudaGetLastError();
    if(e != cudaSuccess){ set_error(std::string(what) + ": " + cudaGetErrorString(e)); return PT_ERR_CUDA; }
    return PT_OK;
}

// persistent grid size: resident CTAs of this kernel on this device (cached per instantiation)
template <class K>
static unsigned persistent_grid(K kernel, int device, size_t smem, int& cache){
    if(cache > 0)return (unsigned)cache;
    int sms = 0, per_sm = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if(smem == 0)cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 0);   // all L1
    else cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kTraceBlock, smem);
    if(sms <= 0)sms = 148;
    if(per_sm <= 0)per_sm = 1;
    cache = sms * per_sm;
    return (unsigned)cache;
}

static int env_int(const char* name, int dflt, int lo, int hi){
    const char* e = getenv(name);
    int v = e ? atoi(e) : dflt;
    return v < lo ? lo : (v > hi ? hi : v);
}
// refill threshold in the low byte, postponed-leaf threshold (PTB_LEAF_MIN, 1 = leaf items are never postponed) above it
// and the traversal rounds per refill check (PTB_ROUNDS: 1 looks every round; 4 measured best, profiles/r02_trace.md) above that
static unsigned refill_min(){ static int v = env_int("PTB_REFILL_MIN", 4, 1, 32) | (env_int("PTB_LEAF_MIN", 1, 1, 32) << 8) | (env_int("PTB_ROUNDS", 4, 1, 16) << 16); return (unsigned)v; }
// shared-memory stack rows per lane: 0 (all local), 8, 16 or 24
static int stack_rows(){ static int v = env_int("PTB_STACK_SMEM", 16, 0, 24); return v; }

// s->buf.stats: 8 x u64 of the counting build, then the scene's traversal-stack overflow counter
static unsigned long long* overflow_counter(pt_scene* sc){ return (unsigned long long*)sc->buf.stats + 8; }

int scene_overflow_count(pt_scene* s, unsigned long long* out){
    if(!cuda_ok(cudaDeviceSynchronize(), "sync") ||
       !cuda_ok(cudaMemcpy(out, overflow_counter(s), sizeof(unsigned long long), cudaMemcpyDeviceToHost), "D2H overflow counter"))return PT_ERR_CUDA;
    return PT_OK;
}

// wavefront form of launch_ks (render.cu): always the full persistent grid (the batch size is only known on the
// device), the cursor belongs to the caller, who clears it between launches
template <int WIDTH, bool ANY, int KS, class Alpha>
static int launch_dyn_ks(pt_scene* sc, const void* d_rays, const DynBatch& dyn, void* d_hits, void* d_inst, void* d_occ, const ConnectAcc& acc, unsigned long long* cursor, void* stream, const Alpha& alpha){
    static int cache[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
    cudaStream_t st = (cudaStream_t)stream;
    auto kernel = k_trace<WIDTH, ANY, false, KS, Alpha>;
    const size_t smem = trace_smem_bytes(KS);
    const unsigned grid = persistent_grid(kernel, sc->device, smem, cache[sc->device & 7]);
    kernel<<<grid, kTraceBlock, smem, st>>>(sc->dev, (const float4*)d_rays, 0, (float4*)d_hits, (uint32_t*)d_inst, (uint8_t*)d_occ,
                                            cursor, nullptr, refill_min(), alpha, StreamCtl(), dyn, overflow_counter(sc), acc);
    return check_launch("k_trace (wavefront)");
}

// counting build of the wavefront launches (pt_scene_set_counting): node / triangle / instance visits of a render's
// closest-hit and shadow rays, for the per-sample algorithmic bytes of SURVEY.md section 8d.  BVH4 layouts only.
template <int WIDTH, bool ANY, class Alpha>
static int launch_dyn_count(pt_scene* sc, const void* d_rays, const DynBatch& dyn, void* d_hits, void* d_inst, void* d_occ, const ConnectAcc& acc, unsigned long long* cursor, void* stream, const Alpha& alpha){
    static int cache[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
    auto kernel = k_trace<WIDTH, ANY, true, 16, Alpha>;
    const size_t smem = trace_smem_bytes(16);
    const unsigned grid = persistent_grid(kernel, sc->device, smem, cache[sc->device & 7]);
    unsigned long long* stats = (unsigned long long*)sc->buf.stats + (ANY ? 24 : 16);
    kernel<<<grid, kTraceBlock, smem, (cudaStream_t)stream>>>(sc->dev, (const float4*)d_rays, 0, (float4*)d_hits, (uint32_t*)d_inst, (uint8_t*)d_occ,
                                                             cursor, stats, refill_min(), alpha, StreamCtl(), dyn, overflow_counter(sc), acc);
    return check_launch("k_trace (wavefront, counting)");
}

template <int WIDTH, bool ANY>
static int launch_dyn(pt_scene* sc, const void* d_rays, const DynBatch& dyn, void* d_hits, void* d_inst, void* d_occ, const ConnectAcc& acc, unsigned long long* cursor, void* stream){
    if constexpr(WIDTH == 4 || WIDTH == kWidthFast4){
        if(sc->counting){
            if(sc->has_alpha_tex || sc->has_wrappers)
                return launch_dyn_count<WIDTH, ANY>(sc, d_rays, dyn, d_hits, d_inst, d_occ, acc, cursor, stream, AlphaTex { sc->shade, sc->alpha_salt += 0x9E3779B97F4A7C15ull });
            return launch_dyn_count<WIDTH, ANY>(sc, d_rays, dyn, d_hits, d_inst, d_occ, acc, cursor, stream, NoAlpha());
        }
    }
    if(sc->has_alpha_tex || sc->has_wrappers){
        const AlphaTex at = { sc->shade, sc->alpha_salt += 0x9E3779B97F4A7C15ull };
        return launch_dyn_ks<WIDTH, ANY, 16>(sc, d_rays, dyn, d_hits, d_inst, d_occ, acc, cursor, stream, at);
    }
    if(sc->dev.single_root != REF_EMPTY)return launch_dyn_ks<WIDTH, ANY, 16>(sc, d_rays, dyn, d_hits, d_inst, d_occ, acc, cursor, stream, NoAlphaSingle());
    return launch_dyn_ks<WIDTH, ANY, 16>(sc, d_rays, dyn, d_hits, d_inst, d_occ, acc, cursor, stream, NoAlpha());
}

int launch_trace_closest_dyn(pt_scene* sc, const void* d_rays, const DynBatch& dyn, void* d_hits, void* d_inst, unsigned long long* cursor, void* stream){
    if(sc->dev.fast)return launch_dyn<kWidthFast4, false>(sc, d_rays, dyn, d_hits, d_inst, nullptr, ConnectAcc(), cursor, stream);
    return sc->dev.width == 2 ? launch_dyn<2, false>(sc, d_rays, dyn, d_hits, d_inst, nullptr, ConnectAcc(), cursor, stream)
         : sc->dev.width == 4 ? launch_dyn<4, false>(sc, d_rays, dyn, d_hits, d_inst, nullptr, ConnectAcc(), cursor, stream)
                              : launch_dyn<8, false>(sc, d_rays, dyn, d_hits, d_inst, nullptr, ConnectAcc(), cursor, stream);
}

int launch_trace_any_dyn(pt_scene* sc, const void* d_rays, const DynBatch& dyn, void* d_occ, const ConnectAcc& acc, unsigned long long* cursor, void* stream){
    if(sc->dev.fast)return launch_dyn<kWidthFast4, true>(sc, d_rays, dyn, nullptr, nullptr, d_occ, acc, cursor, stream);
    return sc->dev.width == 2 ? launch_dyn<2, true>(sc, d_rays, dyn, nullptr, nullptr, d_occ, acc, cursor, stream)
         : sc->dev.width == 4 ? launch_dyn<4, true>(sc, d_rays, dyn, nullptr, nullptr, d_occ, acc, cursor, stream)
                              : launch_dyn<8, true>(sc, d_rays, dyn, nullptr, nullptr, d_occ, acc, cursor, stream);
}

template <int WIDTH, bool ANY, bool STATS, int KS, class Alpha>
static int launch_ks(pt_scene* sc, const void* d_rays, size_t n, void* d_hits, void* d_inst, void* d_occ, void* d_stats, void* stream, const Alpha& alpha){
    static int cache[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };   // per device ordinal (one process drives at most 8 GPUs)
    cudaStream_t st = (cudaStream_t)stream;
    auto kernel = k_trace<WIDTH, ANY, STATS, KS, Alpha>;
    const size_t smem = trace_smem_bytes(KS);
    unsigned grid = persistent_grid(kernel, sc->device, smem, cache[sc->device & 7]);
    const size_t need_ctas = (n + kTraceBlock - 1) / kTraceBlock;
    if(need_ctas < grid)grid = (unsigned)need_ctas;
    // one cursor per launch from a small ring, so launches on different streams never share one
    unsigned long long* cursor = (unsigned long long*)sc->buf.cursors + (sc->cursor_next++ % kCursorRing);
    if(!cuda_ok(cudaMemsetAsync(cursor, 0, sizeof(unsigned long long), st), "cursor reset"))return PT_ERR_CUDA;
    kernel<<<grid, kTraceBlock, smem, st>>>(sc->dev, (const float4*)d_rays, n, (float4*)d_hits, (uint32_t*)d_inst, (uint8_t*)d_occ,
                                            cursor, (unsigned long long*)d_stats, refill_min(), alpha, StreamCtl(), DynBatch(), overflow_counter(sc), ConnectAcc());
    return check_launch("k_trace");
}

template <int WIDTH, bool ANY, bool STATS>
static int launch(pt_scene* sc, const void* d_rays, size_t n, void* d_hits, void* d_inst, void* d_occ, void* d_stats, void* stream){
    if constexpr(STATS)return launch_ks<WIDTH, ANY, STATS, 0>(sc, d_rays, n, d_hits, d_inst, d_occ, d_stats, stream, NoAlpha());
    else{
    const int ks = stack_rows();
    if(sc->has_alpha_tex || sc->has_wrappers){
        // some triangle's material can reject a hit, or some top-level primitive is a wrapper chain: the
        // general kernel variant (evaluates Material::Alpha where a triangle asks for it, AlphaTex::kWrap)
        const AlphaTex at = { sc->shade, sc->alpha_salt += 0x9E3779B97F4A7C15ull };
        if(ks >= 16)return launch_ks<WIDTH, ANY, STATS, 16>(sc, d_rays, n, d_hits, d_inst, d_occ, d_stats, stream, at);
        return launch_ks<WIDTH, ANY, STATS, 0>(sc, d_rays, n, d_hits, d_inst, d_occ, d_stats, stream, at);
    }
    const NoAlpha na;
    if(ks >= 16 && sc->dev.single_root != REF_EMPTY)return launch_ks<WIDTH, ANY, STATS, 16>(sc, d_rays, n, d_hits, d_inst, d_occ, d_stats, stream, NoAlphaSingle());
    if constexpr(WIDTH == 8 || WIDTH == kWidthFast4){
        if(ks >= 16)return launch_ks<WIDTH, ANY, STATS, 16>(sc, d_rays, n, d_hits, d_inst, d_occ, d_stats, stream, na);
        return launch_ks<WIDTH, ANY, STATS, 0>(sc, d_rays, n, d_hits, d_inst, d_occ, d_stats, stream, na);
    }
    if(ks >= 24)return launch_ks<WIDTH, ANY, STATS, 24>(sc, d_rays, n, d_hits, d_inst, d_occ, d_stats, stream, na);
    if(ks >= 16)return launch_ks<WIDTH, ANY, STATS, 16>(sc, d_rays, n, d_hits, d_inst, d_occ, d_stats, stream, na);
    if(ks >= 8)return launch_ks<WIDTH, ANY, STATS, 8>(sc, d_rays, n, d_hits, d_inst, d_occ, d_stats, stream, na);
    return launch_ks<WIDTH, ANY, STATS, 0>(sc, d_rays, n, d_hits, d_inst, d_occ, d_stats, stream, na);
    }
}

// One launch over a ray buffer that is still being filled (capi.cu trace_streamed): plain scenes only
template <int WIDTH, bool ANY, bool COMPACT, class Alpha>
static int launch_feed_as(pt_scene* sc, const void* d_rays, size_t n, void* d_hits, void* d_inst, void* d_occ, const StreamCtl& ctl, void* stream){
    static int cache[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
    cudaStream_t st = (cudaStream_t)stream;
    auto kernel = k_trace<WIDTH, ANY, false, 16, Alpha, true, COMPACT>;
    const size_t smem = trace_smem_bytes(16);
    unsigned grid = persistent_grid(kernel, sc->device, smem, cache[sc->device & 7]);
    const size_t need_ctas = (n + kTraceBlock - 1) / kTraceBlock;
    if(need_ctas < grid)grid = (unsigned)need_ctas;
    unsigned long long* cursor = (unsigned long long*)sc->buf.cursors + (sc->cursor_next++ % kCursorRing);
    if(!cuda_ok(cudaMemsetAsync(cursor, 0, sizeof(unsigned long long), st), "cursor reset"))return PT_ERR_CUDA;
    kernel<<<grid, kTraceBlock, smem, st>>>(sc->dev, (const float4*)d_rays, n, (float4*)d_hits, (uint32_t*)d_inst, (uint8_t*)d_occ,
                                            cursor, nullptr, refill_min(), Alpha(), ctl, DynBatch(), overflow_counter(sc), ConnectAcc());
    return check_launch("k_trace (streamed)");
}
template <int WIDTH, bool ANY, bool COMPACT = false>
static int launch_feed(pt_scene* sc, const void* d_rays, size_t n, void* d_hits, void* d_inst, void* d_occ, const StreamCtl& ctl, void* stream){
    if(sc->dev.single_root != REF_EMPTY)return launch_feed_as<WIDTH, ANY, COMPACT, NoAlphaSingle>(sc, d_rays, n, d_hits, d_inst, d_occ, ctl, stream);
    return launch_feed_as<WIDTH, ANY, COMPACT, NoAlpha>(sc, d_rays, n, d_hits, d_inst, d_occ, ctl, stream);
}

// Can the host enqueue work while a kernel runs?  (Not under a profiler that serialises launches, not with
// CUDA_LAUNCH_BLOCKING.)  A one-thread kernel waits up to ~2 ms for a flag that the host sets right after the launch call
// returns; it reports whether the flag came.
__global__ void k_probe_async(volatile const unsigned* flag, unsigned* seen){
    const long long t0 = clock64();
    unsigned ok = 0u;
    while(clock64() - t0 < 4000000LL){ if(*flag != 0u){ ok = 1u; break; } }
    *seen = ok;
}

bool probe_async_launch(pt_scene* sc, void* d_words, void* compute_stream, void* copy_stream){
    unsigned* w = (unsigned*)d_words;          // w[0] flag, w[1] result
    cudaStream_t cs = (cudaStream_t)compute_stream, xs = (cudaStream_t)copy_stream;
    static const unsigned one = 1u;
    unsigned seen = 0u;
    if(cudaMemsetAsync(w, 0, 8, cs) != cudaSuccess || cudaStreamSynchronize(cs) != cudaSuccess)return false;
    k_probe_async<<<1, 1, 0, cs>>>(w, w + 1);
    if(cudaMemcpyAsync(w, &one, 4, cudaMemcpyHostToDevice, xs) != cudaSuccess)return false;
    if(cudaStreamSynchronize(cs) != cudaSuccess || cudaStreamSynchronize(xs) != cudaSuccess)return false;
    if(cudaMemcpy(&seen, w + 1, 4, cudaMemcpyDeviceToHost) != cudaSuccess)return false;
    (void)sc;
    return seen == 1u;
}

int launch_trace_streamed(pt_scene* sc, bool any, const void* d_rays, size_t n, void* d_hits, void* d_inst, void* d_occ, const StreamCtl& ctl, void* stream, bool compact){
    if(n == 0)return PT_OK;
    const int w = sc->dev.width;
    if(compact){
        // 24 B rays in, 8 B hits out (pt_trace_closest_compact): closest hit only
        if(sc->dev.fast)return launch_feed<kWidthFast4, false, true>(sc, d_rays, n, d_hits, d_inst, nullptr, ctl, stream);
        return w == 2 ? launch_feed<2, false, true>(sc, d_rays, n, d_hits, d_inst, nullptr, ctl, stream)
             : w == 4 ? launch_feed<4, false, true>(sc, d_rays, n, d_hits, d_inst, nullptr, ctl, stream)
                      : launch_feed<8, false, true>(sc, d_rays, n, d_hits, d_inst, nullptr, ctl, stream);
    }
    if(sc->dev.fast)return any ? launch_feed<kWidthFast4, true>(sc, d_rays, n, nullptr, nullptr, d_occ, ctl, stream)
                               : launch_feed<kWidthFast4, false>(sc, d_rays, n, d_hits, d_inst, nullptr, ctl, stream);
    if(any)return w == 2 ? launch_feed<2, true>(sc, d_rays, n, nullptr, nullptr, d_occ, ctl, stream)
                : w == 4 ? launch_feed<4, true>(sc, d_rays, n, nullptr, nullptr, d_occ, ctl, stream)
                         : launch_feed<8, true>(sc, d_rays, n, nullptr, nullptr, d_occ, ctl, stream);
    return w == 2 ? launch_feed<2, false>(sc, d_rays, n, d_hits, d_inst, nullptr, ctl, stream)
         : w == 4 ? launch_feed<4, false>(sc, d_rays, n, d_hits, d_inst, nullptr, ctl, stream)
                  : launch_feed<8, false>(sc, d_rays, n, d_hits, d_inst, nullptr, ctl, stream);
}

// chunked form of pt_trace_closest_compact (small batches, scenes with alpha / wrappers, processes that cannot feed a
// running kernel): the 24 B rays are expanded to pt_ray on the device, the hits packed to 8 B after the ordinary kernel
__global__ void k_expand_rays24(const float2* __restrict__ in, float4* __restrict__ out, size_t n){
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if(i >= n)return;
    const float2 x = in[3 * i], y = in[3 * i + 1], z = in[3 * i + 2];
    out[2 * i] = make_float4(x.x, x.y, y.x, CUDART_INF_F);
    out[2 * i + 1] = make_float4(y.y, z.x, z.y, 0.0f);
}
__global__ void k_pack_hits8(const float4* __restrict__ in, float2* __restrict__ out, size_t n){
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if(i >= n)return;
    const float4 h = in[i];
    out[i] = make_float2(h.x, h.w);
}
int launch_expand_rays24(const void* d_rays24, void* d_rays, size_t n, void* stream){
    if(n == 0)return PT_OK;
    k_expand_rays24<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const float2*)d_rays24, (float4*)d_rays, n);
    return check_launch("k_expand_rays24");
}
int launch_pack_hits8(const void* d_hits, void* d_hits8, size_t n, void* stream){
    if(n == 0)return PT_OK;
    k_pack_hits8<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const float4*)d_hits, (float2*)d_hits8, n);
    return check_launch("k_pack_hits8");
}

int launch_trace_closest(pt_scene* sc, const void* d_rays, size_t n, void* d_hits, void* d_inst, void* stream){
    if(n == 0)return PT_OK;
    if(sc->dev.fast)return launch<kWidthFast4, false, false>(sc, d_rays, n, d_hits, d_inst, nullptr, nullptr, stream);
    return sc->dev.width == 2 ? launch<2, false, false>(sc, d_rays, n, d_hits, d_inst, nullptr, nullptr, stream)
         : sc->dev.width == 4 ? launch<4, false, false>(sc, d_rays, n, d_hits, d_inst, nullptr, nullptr, stream)
                              : launch<8, false, false>(sc, d_rays, n, d_hits, d_inst, nullptr, nullptr, stream);
}

int launch_trace_any(pt_scene* sc, const void* d_rays, size_t n, void* d_occ, void* stream){
    if(n == 0)return PT_OK;
    if(sc->dev.fast)return launch<kWidthFast4, true, false>(sc, d_rays, n, nullptr, nullptr, d_occ, nullptr, stream);
    return sc->dev.width == 2 ? launch<2, true, false>(sc, d_rays, n, nullptr, nullptr, d_occ, nullptr, stream)
         : sc->dev.width == 4 ? launch<4, true, false>(sc, d_rays, n, nullptr, nullptr, d_occ, nullptr, stream)
                              : launch<8, true, false>(sc, d_rays, n, nullptr, nullptr, d_occ, nullptr, stream);
}

// counting build of the same closest-hit traversal
int launch_trace_stats(pt_scene* sc, const void* d_rays, size_t n, void* d_counters, void* stream){
    if(n == 0)return PT_OK;
    if(sc->dev.fast)return launch<kWidthFast4, false, true>(sc, d_rays, n, nullptr, nullptr, nullptr, d_counters, stream);
    return sc->dev.width == 2 ? launch<2, false, true>(sc, d_rays, n, nullptr, nullptr, nullptr, d_counters, stream)
         : sc->dev.width == 4 ? launch<4, false, true>(sc, d_rays, n, nullptr, nullptr, nullptr, d_counters, stream)
                              : launch<8, false, true>(sc, d_rays, n, nullptr, nullptr, nullptr, d_counters, stream);
}

// ---- pt_trace_closest_static: the host-buffer closest-hit ray cast without any allocation -------------------------
// Requested by the run's operators; no counterpart in the reference.  pt_trace_closest stages the batch in device buffers
// it cudaMallocs (and grows) on demand; callers that must not allocate (serving loops, harnesses that forbid cudaMalloc)
// get the same kernel (k_trace, the same instantiation pt_trace_closest_dev launches) over a FIXED scratch area that
// lives in the library image: the ray-dispenser cursor and the staged rays / hits / instance ids of at most
// PT_STATIC_MAX_RAYS rays.  Every GPU that loads the library holds the array (3.4 MB) until exit.  All calls on one
// device share it and it keeps the previous call's contents: the cursor is cleared on the stream before every launch,
// the staging areas are fully overwritten for [0, n).  Calls that overlap on one GPU race on it.
struct StaticScratch {
    unsigned long long cursor;
    unsigned char pad[248];
    float4 rays[2 * PT_STATIC_MAX_RAYS];
    float4 hits[PT_STATIC_MAX_RAYS];
    uint32_t inst[PT_STATIC_MAX_RAYS];
};
static __device__ __align__(256) unsigned char g_static_scratch[sizeof(StaticScratch)];

// In host code `g_static_scratch` is only a handle for the runtime, not a device address: the address is asked for
// once per device, with that device current, and kept per ordinal.
static StaticScratch* static_scratch_ptr(int device){
    static StaticScratch* ptr[64] = {};
    if(device < 0 || device >= 64){ set_error("pt_trace_closest_static: device ordinal out of range"); return nullptr; }
    if(!ptr[device]){
        void* p = nullptr;
        if(!cuda_ok(cudaGetSymbolAddress(&p, g_static_scratch), "cudaGetSymbolAddress(static scratch)"))return nullptr;
        ptr[device] = (StaticScratch*)p;
    }
    return ptr[device];
}

template <int WIDTH, class Alpha>
static int launch_static(pt_scene* sc, StaticScratch* d, size_t n, bool want_inst, cudaStream_t st, const Alpha& alpha){
    static int cache[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
    auto kernel = k_trace<WIDTH, false, false, 16, Alpha>;
    const size_t smem = trace_smem_bytes(16);
    unsigned grid = persistent_grid(kernel, sc->device, smem, cache[sc->device & 7]);
    const size_t need_ctas = (n + kTraceBlock - 1) / kTraceBlock;
    if(need_ctas < grid)grid = (unsigned)need_ctas;
    if(!cuda_ok(cudaMemsetAsync(&d->cursor, 0, sizeof(unsigned long long), st), "cursor reset"))return PT_ERR_CUDA;
    kernel<<<grid, kTraceBlock, smem, st>>>(sc->dev, d->rays, n, d->hits, want_inst ? d->inst : nullptr, nullptr,
                                            &d->cursor, nullptr, refill_min(), alpha, StreamCtl(), DynBatch(), overflow_counter(sc), ConnectAcc());
    return check_launch("k_trace (static scratch)");
}

} // namespace ptb

extern "C" int pt_trace_closest_static(pt_scene* s, const pt_ray* rays, size_t n, pt_hit* hits, uint32_t* inst){
    using namespace ptb;
    if(!s){ set_error("null scene"); return PT_ERR_INVALID; }
    if(!s->committed){ set_error("scene not committed"); return PT_ERR_STATE; }
    if(n > (size_t)PT_STATIC_MAX_RAYS){
        set_error("pt_trace_closest_static: " + std::to_string(n) + " rays exceed PT_STATIC_MAX_RAYS = " + std::to_string(PT_STATIC_MAX_RAYS) +
                  " (the static scratch is fixed; split the batch or use pt_trace_closest)");
        return PT_ERR_INVALID;
    }
    if(n == 0)return PT_OK;
    if(!rays || !hits){ set_error("null buffer"); return PT_ERR_INVALID; }
    if(!cuda_ok(cudaSetDevice(s->device), "cudaSetDevice"))return PT_ERR_CUDA;
    StaticScratch* d = static_scratch_ptr(s->device);
    if(!d)return PT_ERR_CUDA;
    cudaStream_t st = (cudaStream_t)s->stream;
    if(!cuda_ok(cudaMemcpyAsync(d->rays, rays, n * sizeof(pt_ray), cudaMemcpyHostToDevice, st), "H2D rays"))return PT_ERR_CUDA;
    int rc;
    if(s->dev.fast){
        if(s->has_alpha_tex || s->has_wrappers)rc = launch_static<kWidthFast4>(s, d, n, inst != nullptr, st, AlphaTex { s->shade, s->alpha_salt += 0x9E3779B97F4A7C15ull });
        else rc = launch_static<kWidthFast4>(s, d, n, inst != nullptr, st, NoAlpha());
    } else if(s->has_alpha_tex || s->has_wrappers){
        const AlphaTex at = { s->shade, s->alpha_salt += 0x9E3779B97F4A7C15ull };
        rc = s->dev.width == 2 ? launch_static<2>(s, d, n, inst != nullptr, st, at) : s->dev.width == 4 ? launch_static<4>(s, d, n, inst != nullptr, st, at)
                                                                                                      : launch_static<8>(s, d, n, inst != nullptr, st, at);
    } else{
        const NoAlpha na;
        rc = s->dev.width == 2 ? launch_static<2>(s, d, n, inst != nullptr, st, na) : s->dev.width == 4 ? launch_static<4>(s, d, n, inst != nullptr, st, na)
                                                                                                      : launch_static<8>(s, d, n, inst != nullptr, st, na);
    }
    if(rc != PT_OK)return rc;
    if(!cuda_ok(cudaMemcpyAsync(hits, d->hits, n * sizeof(pt_hit), cudaMemcpyDeviceToHost, st), "D2H hits"))return PT_ERR_CUDA;
    if(inst && !cuda_ok(cudaMemcpyAsync(inst, d->inst, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, st), "D2H inst"))return PT_ERR_CUDA;
    if(!cuda_ok(cudaStreamSynchronize(st), "pt_trace_closest_static"))return PT_ERR_CUDA;
    return PT_OK;
}
