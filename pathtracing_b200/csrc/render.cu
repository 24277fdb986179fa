// render.cu — wavefront path integrator: TileIntegrator::Render (Integrators.cpp:23-129) with
// {Simple,,Vol}PathIntegrator::Li (:131-479) unrolled into kernels over ray queues.
//
// One batch = up to `cap` camera samples ("slots").  Per bounce:
//   extend  : persistent closest-hit kernel over the compact ray queue (trace.cu)
//   shade   : one thread per queue entry: rebuild the interaction, miss / emission with MIS,
//             medium event, BSDF sample, NEE sample -> shadow queue, Russian roulette,
//             continuation ray -> next queue (warp-aggregated compaction)
//   connect : persistent any-hit kernel over the shadow queue, then the unoccluded
//             contributions are added to their slot's radiance
// and after the last bounce the slots are splatted into the double RGBW film
// (FilmTile::Add, Film.hpp:65-82).  Path state lives in SoA arrays indexed by slot; queues
// carry slot ids.  Every sample owns a PCG32 stream keyed by (seed, pixel, sample index),
// so the image does not depend on batch size or on how passes are sharded over GPUs.
#include <cuda_runtime.h>
#include <math_constants.h>
#include <chrono>
#include <vector>
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <string>

#include "scene_host.hpp"
#include "shade_mat.cuh"
#include "shade_light.cuh"
#include "camera.cuh"
#include "trace.cuh"
#include "alpha.cuh"

#ifndef PTB_SHADE_MINB
#define PTB_SHADE_MINB 6      // 80 registers, 24 warps / SM: shade stage 5-10 % faster than at 96 registers (C1 2.14 -> 1.94, C3 66.2 -> 59.7, C4 191 -> 182 ms); 8 (64 registers) spills more and gains nothing further
#endif
namespace ptb {

struct FilterParams {   // Filter.hpp:29-144
    int kind, rx, ry;   // integer splat radius = ceil(radius - 0.5)
    float radx, rady;
    double p0, p1, gx, gy, inv_integral;
};

// Path state of one slot: four 16 B quarters
//   0 thr   throughput rgb, prevPDF           1 rad   radiance rgb
//   2 rng (pcg state, 8 B; inc is rebuilt from the sample id), depth (low 16) | rr_depth (high 16), mflag (bit0 spec, bits 1..: medium id + 1)
//   3 gid (which camera sample lives in the slot, 8 B), spos (fractional sample position inside the pixel)
// laid out one of two ways (ps_at):
//  * records (quarter k of slot i at 4 i + k): a call whose samples outnumber the pool regenerates finished slots,
//    and after a few rounds queue order is a random order over the pool: every access is a gather / scatter.  As
//    seven arrays that cost seven partial-sector writes per regenerated sample (DRAM read-modify-write each:
//    k_generate ran at a tenth of its sequential speed, 7 % of a C3 frame); a record is two whole sectors.
//    C3 335 -> 371, C4 285 -> 307 Msamples/s.
//  * planes (quarter k of slot i at k cap + i): a call that fits the pool (C1) keeps queue order == slot order, and
//    planes keep a warp's access to one quarter on 4 cache lines instead of 16 (records cost C1 8 %).
struct Arena {
    float4* rays[2]; uint32_t* qslot[2];
    float4* hits; uint32_t* hinst;
    float4* srays; uint32_t* sslot; float4* scontrib; uint8_t* socc;
    float4* ps;         // per-slot path state, 4 quarters per slot (see above)
    uint32_t ps_slot, ps_quarter;   // index strides: (4, 1) records | (1, cap) planes
    uint8_t* skey; uint32_t* perm; uint32_t* shist;   // shade-order sort: class per queue entry, sorted order, per-block histograms
    uint32_t* done;     // slots whose path ended in the last shade pass: splatted, then handed to new samples
    struct WaveCtl* ctl; // queue sizes and everything else the loop decides per bounce: on the device, never read by the host inside the loop
};

// Control block of the wavefront loop.  The host enqueues whole bounces without knowing any queue size: every kernel
// is a persistent grid that reads its batch from here, the last block of k_finish closes a bounce (which camera samples refill the slots
// that ended, which queue is current next, cursors back to zero).  The host only watches n_cur a few bounces late to
// find out when to stop (render_passes).  Round 1 read three counters back and synchronised the stream every bounce:
// 18 us x 1250 bounces on C5, and most of a C1 frame on 8 GPUs.
struct WaveCtl {
    uint32_t n_cur;        // rays in the current queue
    uint32_t n_next;       // continuation rays appended by k_shade: the next current queue
    uint32_t n_sh;         // shadow rays of this bounce
    uint32_t n_done;       // paths that ended in this bounce
    uint32_t cur;          // which of the two ray queues is current
    uint32_t n_gen;        // camera samples k_generate starts after this bounce
    uint32_t q_off;        // where they go in the (new) current queue
    uint32_t nonfinite;    // non-finite samples of the call (counted, not splatted)
    unsigned long long gen_first, next_gid, count;
    unsigned long long cursor[2];                         // ray dispensers of the closest-hit / any-hit launches
    unsigned long long closest_rays, shadow_rays, samples, bounces;   // totals of the call
    uint32_t ticket;       // blocks of k_finish that have ended (the last one plans the next bounce)
};

__device__ __forceinline__ float4* ps_at(const Arena& A, uint32_t slot, uint32_t k){
    return A.ps + ((size_t)slot * A.ps_slot + (size_t)k * A.ps_quarter);
}

struct BatchInfo {     // how a sample id maps to (pixel, sample index): id = pass_k * W*H + pixel
    int W, H, pass_begin, pass_stride, max_depth;
    int sampler, sx, sy;   // 0 UniformSampler | 1 StratifiedSampler(sx, sy)
    uint64_t seed;
    // adaptive rounds after the first: only the pixels still listed get samples; id = k * n_list + j
    // -> pixel list[j], sample index sample_base + k
    const uint32_t* list;
    uint32_t n_list, sample_base;
    int tr_walk;           // VolPath in a scene with material-less medium boundaries: shadow rays take the IntersectTr walk
};

__device__ __forceinline__ uint64_t stream_of(const BatchInfo& B, uint64_t g, uint32_t& pixel, uint32_t& sample){
    const uint64_t wh = (uint64_t)B.W * B.H;
    if(B.list){
        pixel = B.list[g % B.n_list];
        sample = B.sample_base + (uint32_t)B.pass_begin + (uint32_t)(g / B.n_list) * (uint32_t)B.pass_stride;
    } else{
        pixel = (uint32_t)(g % wh);
        sample = (uint32_t)(B.pass_begin + (g / wh) * B.pass_stride);
    }
    return (uint64_t)sample * wh + pixel;
}

// warp-aggregated queue append: one atomic per warp
__device__ __forceinline__ uint32_t queue_slot(uint32_t* counter, bool want){
    const unsigned m = __ballot_sync(0xffffffffu, want);
    if(!m)return 0;
    const unsigned lane = threadIdx.x & 31u;
    const int leader = __ffs(m) - 1;
    uint32_t base = 0;
    if((int)lane == leader)base = atomicAdd(counter, (uint32_t)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    return base + __popc(m & ((1u << lane) - 1u));
}

// ---- generate: StartPixelSample / getPixel2D / get1D / get2D / Camera::GenerateRay (Integrators.cpp:61-64) ----
// n_new camera samples gid_first.. go into the slots free[i] (or free_base + i) and onto the end of queue `cur`
// one camera sample: gid -> slot `slot`, ray -> position q of queue `cur`
__device__ __forceinline__ void generate_one(const Arena& A, const BatchInfo& B, const CamParams& C, uint64_t gid, uint32_t slot, size_t q, int cur){
    uint32_t pixel, sample;
    const uint64_t stream = stream_of(B, gid, pixel, sample);
    Pcg g = pcg_seed(B.seed, stream);
    const int x = (int)(pixel % (uint32_t)B.W), y = (int)(pixel / (uint32_t)B.W);
    double jx, jy; float time, lx, ly;
    if(B.sampler == 0){
        jx = rnd_double(g); jy = rnd_double(g);                 // UniformSampler::getPixel2D
        time = (float)rnd_double(g);
        lx = (float)rnd_double(g); ly = (float)rnd_double(g);
    } else{
        // StratifiedSampler (Sampler.hpp:73-151): dimensions 0-1 pixel, 2 time, 3-4 lens; the stratum of
        // sample `index` is a hash-seeded permutation, the jitter inside it is random_float()
        const uint32_t nS = (uint32_t)(B.sx * B.sy);
        const uint32_t index = sample - B.sample_base;           // sample_index restarts every adaptive round (Integrators.cpp:60)
        uint32_t st = permutation_element(index, nS, (uint32_t)hash_ppd((uint32_t)x, (uint32_t)y, 0));
        jx = ((double)(st % (uint32_t)B.sx) + (double)rnd_float(g)) / (double)B.sx;
        jy = ((double)(st / (uint32_t)B.sx) + (double)rnd_float(g)) / (double)B.sy;
        st = permutation_element(index, nS, (uint32_t)hash_ppd((uint32_t)x, (uint32_t)y, 2));
        time = (float)(((double)st + (double)rnd_float(g)) / (double)nS);
        st = permutation_element(index, nS, (uint32_t)hash_ppd((uint32_t)x, (uint32_t)y, 3));
        lx = (float)(((double)(st % (uint32_t)B.sx) + (double)rnd_float(g)) / (double)B.sx);
        ly = (float)(((double)(st / (uint32_t)B.sx) + (double)rnd_float(g)) / (double)B.sy);
    }
    const float px = (float)((double)x + jx), py = (float)((double)y + jy);   // GenerateRay takes a glm::vec2
    V3 org, dir; float t;
    camera_ray(C, B.W, B.H, px, py, time, lx, ly, org, dir, t);
    float4* __restrict__ qr = cur ? A.rays[1] : A.rays[0];
    qr[2 * q] = make_float4(org.x, org.y, org.z, CUDART_INF_F);
    qr[2 * q + 1] = make_float4(dir.x, dir.y, dir.z, t);
    (cur ? A.qslot[1] : A.qslot[0])[q] = slot;
    *ps_at(A, slot, 0) = make_float4(1.0f, 1.0f, 1.0f, 1.0f);
    *ps_at(A, slot, 1) = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    // depth 1: `depth++ < maxDepth` already taken for bounce 1
    *ps_at(A, slot, 2) = make_float4(__uint_as_float((uint32_t)g.state), __uint_as_float((uint32_t)(g.state >> 32)), __uint_as_float(1u),
                       __uint_as_float(1u | ((uint32_t)(C.medium + 1) << 1)));
    *ps_at(A, slot, 3) = make_float4(__uint_as_float((uint32_t)gid), __uint_as_float((uint32_t)(gid >> 32)), (float)jx, (float)jy);
}

// the first samples of a run: slots 0 .. n_gen - 1, queue 0 (k_wave_init)
__global__ void __launch_bounds__(256) k_generate(Arena A, BatchInfo B, CamParams C){
    const WaveCtl& W = *A.ctl;
    const uint32_t n_new = W.n_gen;
    const uint64_t gid_first = W.gen_first;
    for(uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_new; i += gridDim.x * blockDim.x)
        generate_one(A, B, C, gid_first + i, i, (size_t)i, 0);
}

// ---- plan: close one bounce on the device (run by ONE thread, after every other block of k_finish has ended) ----
__device__ __forceinline__ void plan_bounce(WaveCtl& W){
    const unsigned long long left = W.count - W.next_gid;
    const uint32_t m = (uint32_t)(left < (unsigned long long)W.n_done ? left : (unsigned long long)W.n_done);
    W.closest_rays += W.n_cur; W.shadow_rays += W.n_sh; W.samples += m; W.bounces += W.n_cur ? 1ull : 0ull;
    W.n_gen = m; W.gen_first = W.next_gid; W.next_gid += m;
    W.q_off = W.n_next;
    W.cur ^= 1u;
    W.n_cur = W.n_next + m;
    W.n_next = 0u; W.n_sh = 0u; W.n_done = 0u;
    W.cursor[0] = 0ull; W.cursor[1] = 0ull;
    W.ticket = 0u;
}

// start of a run of `count` camera samples: the first min(cap, count) go straight into slots 0.. of queue 0
__global__ void k_wave_init(Arena A, unsigned long long gid_base, unsigned long long count, unsigned long long cap, int reset_totals){
    WaveCtl& W = *A.ctl;
    const uint32_t n = (uint32_t)(count - gid_base < cap ? count - gid_base : cap);
    W.n_cur = n; W.n_next = 0u; W.n_sh = 0u; W.n_done = 0u; W.cur = 0u;
    W.n_gen = n; W.q_off = 0u; W.gen_first = gid_base; W.next_gid = gid_base + n; W.count = count;
    W.cursor[0] = 0ull; W.cursor[1] = 0ull; W.ticket = 0u;
    if(reset_totals){ W.nonfinite = 0u; W.closest_rays = 0ull; W.shadow_rays = 0ull; W.samples = 0ull; W.bounces = 0ull; }
    W.samples += n;
}

// ---- connect: the unoccluded NEE contributions (Integrators.cpp:243-244, 356, 398) are added by the any-hit kernel
// itself when a shadow ray retires (trace.cu, ConnectAcc): no separate pass, no occlusion flags in memory.

// ---- connect through medium boundaries: Scene::IntersectTr (Scene.cpp:8-29) ---------------------------
// VolPathIntegrator::SampleLd (Integrators.cpp:466) does not ask "is anything in the way" but walks the
// shadow ray with closest hits: a primitive without a material (a medium boundary) lets it through and
// switches the medium, transmittance is collected per segment, the first primitive WITH a material blocks.
// One lane walks one shadow ray start to end (whole-ray traversal, local stack): only scenes that hold
// such boundaries take this path, everything else keeps the any-hit kernel (which adds the contribution itself).
template <int WIDTH, class Alpha>
__global__ void __launch_bounds__(128) k_shadow_tr(Arena A, DevScene D, ShadeScene S, Alpha alpha, unsigned long long* __restrict__ overflow){
    const uint32_t n = A.ctl->n_sh;
    for(uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x){
    RayRegs ray; ray.a = A.srays[2 * (size_t)k]; ray.b = A.srays[2 * (size_t)k + 1];
    const float4 c = A.scontrib[k];
    int med = __float_as_int(c.w);
    const V3 d = v3(ray.b.x, ray.b.y, ray.b.z);
    V3 tr = v3(1.0f);
    bool blocked = false;
    for(int seg = 0; seg < 256 && ray.a.w > 0.0f; seg++){
        HitRec h;
        const bool found = trace_scene<WIDTH, false, false>(D, ray, k, h, nullptr, alpha);
        if(h.overflow)atomicAdd(overflow, 1ull);
        if(!found){
            if(med >= 0)tr *= medium_Tr(S.media[med], ray.a.w);
            break;
        }
        if(med >= 0)tr *= medium_Tr(S.media[med], h.t);
        SurfInt si;
        const V3 o = v3(ray.a.x, ray.a.y, ray.a.z);
        build_interaction(S, o, d, ray.b.w, h.t, h.u, h.v, h.prim, h.inst, si);
        if(si.mat >= 0){ blocked = true; break; }
        const V3 q = o + h.t * d;
        med = dot(d, si.n) < 0.0f ? si.medium : -1;                   // interaction.getMedium(ray.dir)
        ray.a = make_float4(q.x, q.y, q.z, ray.a.w - h.t);
    }
    if(blocked)continue;
    const uint32_t slot = A.sslot[k];
    float4 r = *ps_at(A, slot, 1);
    r.x += c.x * tr.x; r.y += c.y * tr.y; r.z += c.z * tr.z;
    *ps_at(A, slot, 1) = r;
    }
}

template <class Alpha>
static void launch_shadow_tr(const Arena& A, const pt_scene* s, unsigned grid, cudaStream_t st, const Alpha& alpha){
    unsigned long long* ovf = (unsigned long long*)s->buf.stats + 8;
    if(s->dev.fast){ k_shadow_tr<kWidthFast4><<<grid, 128, 0, st>>>(A, s->dev, s->shade, alpha, ovf); return; }
    switch(s->dev.width){
    case 2: k_shadow_tr<2><<<grid, 128, 0, st>>>(A, s->dev, s->shade, alpha, ovf); break;
    case 4: k_shadow_tr<4><<<grid, 128, 0, st>>>(A, s->dev, s->shade, alpha, ovf); break;
    default: k_shadow_tr<8><<<grid, 128, 0, st>>>(A, s->dev, s->shade, alpha, ovf); break;
    }
}

// ---- film: FilmTile::Add + Film::Merge (Film.hpp:65-82,125-132) ------------------------------
__device__ inline double filter_eval(const FilterParams& F, float x, float y){
    switch(F.kind){
    case PT_FILTER_BOX: return (fabsf(x) <= F.radx && fabsf(y) <= F.rady) ? 1.0 : 0.0;
    case PT_FILTER_GAUSSIAN: {
        const double s = F.p0, k = 0.5641895835477563 / (s * 1.4142135623730951);
        const double gx = k * exp(-((double)x * x) / (2 * s * s)), gy = k * exp(-((double)y * y) / (2 * s * s));
        return fmax(0.0, gx - F.gx) * fmax(0.0, gy - F.gy);
    }
    case PT_FILTER_MITCHELL: {
        const double b = F.p0, c = F.p1;
        double r = 1.0;
        const double xs[2] = { 2.0 * x / F.radx, 2.0 * y / F.rady };
        for(int k = 0; k < 2; k++){
            const double a = fabs(xs[k]);
            double m;
            if(a <= 1.0)m = 1.0 / 6.0 * ((12 - 9 * b - 6 * c) * a * a * a + (-18 + 12 * b + 6 * c) * a * a + (6 - 2 * b));
            else if(a <= 2.0)m = 1.0 / 6.0 * ((-b - 6 * c) * a * a * a + (6 * b + 30 * c) * a * a + (-12 * b - 48 * c) * a + (8 * b + 24 * c));
            else m = 0.0;
            r *= m;
        }
        return r;
    }
    default: {  // Lanczos: WindowedSinc(x, radius, tau)
        double r = 1.0;
        const double xs[2] = { x, y }, rs[2] = { F.radx, F.rady };
        for(int k = 0; k < 2; k++){
            const double a = xs[k];
            double w;
            if(fabs(a) > rs[k])w = 0.0;
            else{
                const double pi = 3.141592653589793;
                const double s1 = (1.0 - a * a == 1.0) ? 1.0 : sin(pi * a) / (pi * a);
                const double at = a / F.p0;
                const double s2 = (1.0 - at * at == 1.0) ? 1.0 : sin(pi * at) / (pi * at);
                w = s1 * s2;
            }
            r *= w;
        }
        return r;
    }
    }
}

// est: this call's per-pixel estimator sums for the adaptive loop, 6 doubles per pixel {sum rgb, sum of squares rgb}, or null
// ---- finish a bounce: film + regeneration + plan in one launch ---------------------------------------------------------
// Thread i takes done[i]: splats that slot's finished sample (FilmTile::Add + Film::Merge), and -- while camera samples are
// left -- starts the next one in the same slot, its ray appended behind the continuation rays k_shade queued.  The block
// that ends last (a ticket in WaveCtl) then closes the bounce for everybody (plan_bounce).  Round 1 ran these as
// k_film, a counters read-back on the host and k_generate; as three launches they cost a C1 shard a third of its frame.
// whole_pool: splat slots 0 .. n_cur - 1 instead (max_depth == 0: black samples, nothing else runs).
// est: this call's per-pixel estimator sums for the adaptive loop, 6 doubles per pixel {sum rgb, sum of squares rgb}, or null
__global__ void __launch_bounds__(256) k_finish(Arena A, BatchInfo B, FilterParams F, CamParams C, double* __restrict__ film, double* __restrict__ film_sq,
                                                double* __restrict__ est, int whole_pool){
    WaveCtl& W = *A.ctl;
    const uint32_t n_done = whole_pool ? W.n_cur : W.n_done;
    const uint32_t* __restrict__ done = whole_pool ? nullptr : A.done;
    const unsigned long long left = W.count - W.next_gid;
    const uint32_t m = whole_pool ? 0u : (uint32_t)(left < (unsigned long long)n_done ? left : (unsigned long long)n_done);
    const uint64_t gid_first = W.next_gid;
    const uint32_t q_off = W.n_next;
    const int ncur = (int)(W.cur ^ 1u);
    for(uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_done; i += gridDim.x * blockDim.x){
    const uint32_t slot = done ? done[i] : i;
    uint32_t pixel, sample;
    const float4 q3 = *ps_at(A, slot, 3);
    stream_of(B, (uint64_t)__float_as_uint(q3.x) | ((uint64_t)__float_as_uint(q3.y) << 32), pixel, sample);
    const int x = (int)(pixel % (uint32_t)B.W), y = (int)(pixel / (uint32_t)B.W);
    const float4 r = *ps_at(A, slot, 1);
    const float2 sp = make_float2(q3.z, q3.w);
    if(!(fabsf(r.x) <= 3.0e38f && fabsf(r.y) <= 3.0e38f && fabsf(r.z) <= 3.0e38f)){
        atomicAdd(&W.nonfinite, 1u);         // non-finite radiance: counted, not splatted (one NaN would own the pixel)
    } else{
        {   // per-pixel sum of squared sample radiance (owner pixel): the variance the 3-sigma parity test
            // and VarianceEstimator (Util.hpp:8-43) need
            double* q = film_sq + 3 * ((size_t)y * B.W + x);
            atomicAdd(q, (double)r.x * r.x); atomicAdd(q + 1, (double)r.y * r.y); atomicAdd(q + 2, (double)r.z * r.z);
            if(est){
                double* e = est + 6 * ((size_t)y * B.W + x);
                atomicAdd(e, (double)r.x); atomicAdd(e + 1, (double)r.y); atomicAdd(e + 2, (double)r.z);
                atomicAdd(e + 3, (double)r.x * r.x); atomicAdd(e + 4, (double)r.y * r.y); atomicAdd(e + 5, (double)r.z * r.z);
            }
        }
        for(int dy = -F.ry; dy <= F.ry; dy++)for(int dx = -F.rx; dx <= F.rx; dx++){
            const float fx = (float)((double)dx + 0.5 - (double)sp.x), fy = (float)((double)dy + 0.5 - (double)sp.y);   // Filter::Evaluate takes a glm::vec2
            const double w = filter_eval(F, fx, fy) * F.inv_integral;
            const int qx = x + dx, qy = y + dy;
            if(w <= 0.0 || qx < 0 || qy < 0 || qx >= B.W || qy >= B.H)continue;
            double* q = film + 4 * ((size_t)qy * B.W + qx);
            atomicAdd(q, (double)r.x * w); atomicAdd(q + 1, (double)r.y * w); atomicAdd(q + 2, (double)r.z * w); atomicAdd(q + 3, w);
        }
    }
    if(i < m)generate_one(A, B, C, gid_first + i, slot, (size_t)q_off + i, ncur);
    }
    if(whole_pool)return;
    // last block out plans the next bounce: every other block has finished reading WaveCtl by then
    __shared__ bool last;
    __syncthreads();
    if(threadIdx.x == 0){
        __threadfence();
        last = atomicAdd(&W.ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if(last && threadIdx.x == 0){ __threadfence(); plan_bounce(W); }
}

// ---- adaptive loop: the per-pixel stopping rule of TileIntegrator::Render (Integrators.cpp:59-86) ----
// After a round of spp samples a pixel goes on while any channel's RelativeVariance
// (1.96 * sqrt(Var / n) / mean, 0 when mean == 0; Util.hpp:38-40) exceeds 1.5 and n < 128 * spp.  The
// reference weights the channels by the luminance coefficients first; the ratio does not change.
// in == null: every pixel was sampled this round.  n = samples per pixel so far.
__global__ void __launch_bounds__(256) k_adapt(const double* __restrict__ est, const uint32_t* __restrict__ in, uint32_t n_in,
                                               uint32_t* __restrict__ out, uint32_t* __restrict__ n_out, double n){
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    bool more = false;
    uint32_t pixel = 0;
    if(j < n_in){
        pixel = in ? in[j] : j;
        const double* e = est + 6 * (size_t)pixel;
        for(int c = 0; c < 3; c++){
            const double mean = e[c] / n;
            const double var = n > 1.0 ? fmax(0.0, e[3 + c] - n * mean * mean) / (n - 1.0) : 0.0;
            const double rel = mean == 0.0 ? 0.0 : 1.96 * sqrt(var / n) / mean;
            more = more || !(rel <= 1.5);
        }
    }
    const uint32_t at = queue_slot(n_out, more);
    if(more)out[at] = pixel;
}

// est += round; round = 0 (sharded adaptive rounds, after the all-reduce of `round`)
__global__ void __launch_bounds__(256) k_est_merge(double* __restrict__ est, double* __restrict__ round, size_t n){
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if(i < n){ est[i] += round[i]; round[i] = 0.0; }
}

// ---- shade order: queue entries grouped by what k_shade will do with them -----------------------------------
// A warp of k_shade that holds a miss, a medium boundary, a Lambertian, a rough dielectric and a mirror runs all
// five code paths one after the other (profiles/r01_shade.md: 10.9 of 32 lanes per instruction).  A stable
// counting sort of the queue by class -- key pass with per-block histograms, one scan, one scatter pass --
// gives k_shade an order in which neighbouring lanes take the same branch; it reads its entry through `perm`.
constexpr int kSortBlock = 256;
constexpr int kSortKeys = 8;       // 0 miss | 1 medium boundary | 2.. material kind (+4 when the hit carries an area light)

__device__ __forceinline__ uint32_t shade_class(const Arena& A, const ShadeScene& S, uint32_t j){
    const uint32_t prim = __float_as_uint(A.hits[j].w);
    if(prim == 0xFFFFFFFFu)return 0u;
    const DevInstShade& I = S.inst[A.hinst[j]];
    int mat;
    if(I.shape >= 0)mat = I.material;
    else mat = S.meshes[S.stri[I.tri_base + prim].mesh].material;
    if(mat < 0)return 1u;
    return 2u + (uint32_t)S.materials[mat].kind;      // pt_mat_kind 0..3
}

// The batch size lives on the device: a fixed grid of kSortGrid blocks, block b owns the contiguous range
// [b * per, (b + 1) * per) of the queue (per = a multiple of kSortBlock), so block order == queue order and the sort stays stable.
constexpr unsigned kSortGrid = 1184;      // 148 SMs x 8

__device__ __forceinline__ uint32_t sort_per(uint32_t n){
    const uint32_t per = (n + kSortGrid - 1) / kSortGrid;
    return (per + kSortBlock - 1) / kSortBlock * kSortBlock;
}

__global__ void __launch_bounds__(kSortBlock) k_sort_count(Arena A, ShadeScene S, uint32_t sort_min){
    const uint32_t n = A.ctl->n_cur;
    if(n < sort_min)return;
    __shared__ uint32_t h[kSortKeys];
    if(threadIdx.x < kSortKeys)h[threadIdx.x] = 0u;
    __syncthreads();
    const uint32_t per = sort_per(n), lo = blockIdx.x * per, hi = min(n, lo + per);
    for(uint32_t j = lo + threadIdx.x; j < hi; j += kSortBlock){
        const uint32_t k = shade_class(A, S, j);
        A.skey[j] = (uint8_t)k;
        atomicAdd(&h[k], 1u);
    }
    __syncthreads();
    if(threadIdx.x < kSortKeys)A.shist[threadIdx.x * kSortGrid + blockIdx.x] = h[threadIdx.x];   // key-major: one scan gives final offsets
}

// exclusive scan of the kSortKeys x kSortGrid counters in place, one block
__global__ void __launch_bounds__(1024) k_sort_scan(Arena A, uint32_t sort_min){
    if(A.ctl->n_cur < sort_min)return;
    uint32_t* __restrict__ v = A.shist;
    const uint32_t m = kSortGrid * kSortKeys;
    __shared__ uint32_t part[1024];
    const uint32_t per = (m + 1023u) / 1024u;
    const uint32_t lo = threadIdx.x * per, hi = min(lo + per, m);
    uint32_t sum = 0;
    for(uint32_t i = lo; i < hi; i++)sum += v[i];
    part[threadIdx.x] = sum;
    __syncthreads();
    for(int o = 1; o < 1024; o <<= 1){
        const uint32_t x = threadIdx.x >= (unsigned)o ? part[threadIdx.x - o] : 0u;
        __syncthreads();
        part[threadIdx.x] += x;
        __syncthreads();
    }
    uint32_t run = part[threadIdx.x] - sum;
    for(uint32_t i = lo; i < hi; i++){ const uint32_t c = v[i]; v[i] = run; run += c; }
}

__global__ void __launch_bounds__(kSortBlock) k_sort_scatter(Arena A, uint32_t sort_min){
    const uint32_t n = A.ctl->n_cur;
    if(n < sort_min)return;
    __shared__ uint32_t wc[kSortBlock / 32][kSortKeys];
    __shared__ uint32_t base[kSortKeys];
    const unsigned lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    if(threadIdx.x < kSortKeys)base[threadIdx.x] = A.shist[threadIdx.x * kSortGrid + blockIdx.x];
    const uint32_t per = sort_per(n), lo = blockIdx.x * per, hi = min(n, lo + per);
    for(uint32_t j0 = lo; j0 < hi; j0 += kSortBlock){
        const uint32_t j = j0 + threadIdx.x;
        const uint32_t k = j < hi ? (uint32_t)A.skey[j] : 0xFFu;
        uint32_t rank = 0;
#pragma unroll
        for(uint32_t c = 0; c < (uint32_t)kSortKeys; c++){
            const unsigned m = __ballot_sync(0xffffffffu, k == c);
            if(k == c)rank = __popc(m & ((1u << lane) - 1u));
            if(lane == 0)wc[w][c] = __popc(m);
        }
        __syncthreads();
        if(j < hi){
            uint32_t off = base[k] + rank;
            for(unsigned q = 0; q < w; q++)off += wc[q][k];
            A.perm[off] = j;
        }
        __syncthreads();
        if(threadIdx.x < kSortKeys){
            uint32_t t = 0;
            for(unsigned q = 0; q < kSortBlock / 32; q++)t += wc[q][threadIdx.x];
            base[threadIdx.x] += t;
        }
        __syncthreads();
    }
}

// ---- shade: one bounce of Li() for every queue entry ------------------------------------------

struct NeeOut { bool want; V3 o, d; float tmax; V3 contrib; int medium; };

// PathIntegrator::SampleLd (Integrators.cpp:260-294) / VolPathIntegrator::SampleLd (:426-479) up to the
// visibility test, which the connect stage resolves.  `medium_pt`: the interaction is a medium event.
template <int INTEG>
__device__ inline void sample_ld(const ShadeScene& S, const SurfInt& si, const MatCtx& mc, bool medium_pt, V3 ip, V3 rd, float time, float hg_g, int cur_medium,
                                 float u_sel, float ux, float uy, Pcg& rng, bool tr_walk, NeeOut& out){
    out.want = false; out.medium = cur_medium;
    const int li = pick_light(S, u_sel);
    if(li < 0)return;
    const DevLight& L = S.lights[li];
    LightSmp ls;
    light_sample(S, L, ux, uy, time, rng, ls);
    V3 ldir; float t;
    if(is_zero(ls.n)){ ldir = ls.dir; t = CUDART_INF_F; }
    else{
        ldir = ls.p - ip;
        t = length(ldir) - kEps;
        if(INTEG == PT_INTEGRATOR_VOLPATH)t -= kEps;
    }
    const V3 sd = normalize(ldir);
    float pmf = L.pmf_sample;
    if(pmf <= 0.0f)return;
    V3 f; float spdf;
    if(medium_pt){
        spdf = phase_hg(dot(rd, sd), hg_g);
        f = v3(spdf);
    } else{
        const float d = dot(si.ns, sd);
        if(d * dot(rd, si.ns) >= 0.0f)return;
        f = mat_eval(mc, si, rd, sd) * fabsf(d);
        spdf = mat_pdf(mc, si, rd, sd);
    }
    if(INTEG == PT_INTEGRATOR_VOLPATH && is_zero(f))return;
    V3 c;
    if(light_is_delta(L))c = ls.L * f / pmf;
    else{
        pmf *= light_pdf_sampled(S, L, ls.p, ls.n, ip, sd, time);
        if(pmf <= 0.0f)return;
        const float w2 = pmf * pmf, w1 = spdf * spdf;
        c = light_L_sampled(S, L, ls, sd) * f * (w2 / (w1 + w2)) / pmf;
    }
    if(INTEG == PT_INTEGRATOR_VOLPATH && cur_medium >= 0 && !tr_walk)c = c * medium_Tr(S.media[cur_medium], t);   // Scene::IntersectTr, miss leg (Scene.cpp:13-17)
    out.want = true; out.o = ip; out.d = sd; out.tmax = t; out.contrib = c;
}

// 96 registers: forcing 6 or 8 resident CTAs (80 / 64 registers, spills) measured slower
template <int INTEG>
__global__ void __launch_bounds__(128, PTB_SHADE_MINB) k_shade(Arena A, ShadeScene S, BatchInfo B, uint32_t sort_min){
    WaveCtl& W = *A.ctl;
    const uint32_t n = W.n_cur;
    const int cur = (int)W.cur;
    const uint32_t* __restrict__ perm = n >= sort_min ? A.perm : nullptr;     // shade order (k_sort_*) or queue order
    const float4* __restrict__ q_rays = cur ? A.rays[1] : A.rays[0];
    const uint32_t* __restrict__ q_slot = cur ? A.qslot[1] : A.qslot[0];
    float4* __restrict__ nq_rays = cur ? A.rays[0] : A.rays[1];
    uint32_t* __restrict__ nq_slot = cur ? A.qslot[0] : A.qslot[1];
    // persistent: the block walks the queue in strides of the grid (the loop bound is block-uniform: every lane of a
    // warp takes part in the ballots of queue_slot)
    for(uint32_t tile = blockIdx.x * blockDim.x; tile < n; tile += gridDim.x * blockDim.x){
    const uint32_t tid = tile + threadIdx.x;
    const uint32_t j = (perm && tid < n) ? perm[tid] : tid;
    bool cont = false, nee = false;
    V3 no = v3(0.0f), nd = v3(0.0f); float ntime = 0.0f;
    NeeOut ne; ne.want = false;
    uint32_t slot = 0;
    if(j < n){
        slot = q_slot[j];
        const float4 ra = q_rays[2 * (size_t)j], rb = q_rays[2 * (size_t)j + 1];
        const V3 ro = v3(ra.x, ra.y, ra.z), rd = v3(rb.x, rb.y, rb.z);
        ntime = rb.w;
        const float4 h = A.hits[j];
        const uint32_t prim = __float_as_uint(h.w);
        float4 thr4 = *ps_at(A, slot, 0);
        V3 att = v3(thr4.x, thr4.y, thr4.z);
        float prev_pdf = thr4.w;
        float4 rad4 = *ps_at(A, slot, 1);
        V3 outp = v3(rad4.x, rad4.y, rad4.z);
        const uint4 st4 = *reinterpret_cast<const uint4*>(ps_at(A, slot, 2));     // rng, depth, mflag
        const float4 q3 = *ps_at(A, slot, 3);                                    // gid, spos
        uint32_t dw = st4.z, mf = st4.w;
        uint32_t depth = dw & 0xFFFFu, rr_depth = dw >> 16;
        bool spec = (mf & 1u) != 0;
        int medium = (int)(mf >> 1) - 1;
        uint32_t pixel, sample;
        Pcg g; g.inc = pcg_inc(B.seed, stream_of(B, (uint64_t)__float_as_uint(q3.x) | ((uint64_t)__float_as_uint(q3.y) << 32), pixel, sample)); g.state = (uint64_t)st4.x | ((uint64_t)st4.y << 32);

        if(prim == 0xFFFFFFFFu){
            // miss: infinite lights (Integrators.cpp:136-141, 197-207, 308-330)
            for(int k = 0; k < S.n_inf_lights; k++){
                const DevLight& L = S.lights[S.inf_lights[k]];
                const V3 le = inf_Le(S, L, rd);
                if(INTEG == PT_INTEGRATOR_SIMPLE || spec)outp += att * le;
                else if(prev_pdf > 0.0f){
                    const float lp = L.pmf_sample * light_pdf_sampled(S, L, v3(0.0f), v3(0.0f), ro, rd, ntime);
                    outp += att * le * (prev_pdf * prev_pdf / (prev_pdf * prev_pdf + lp * lp));
                }
            }
        } else{
            SurfInt si;
            build_interaction(S, ro, rd, ntime, h.x, h.y, h.z, prim, A.hinst[j], si);
            bool med_event = false; float med_dist = 0.0f;
            int ray_medium = -1;        // ray.medium while this bounce is evaluated (the shadow ray inherits it)
            if(INTEG == PT_INTEGRATOR_VOLPATH){
                if(medium < 0)medium = S.scene_medium;                           // Integrators.cpp:340-341
                ray_medium = medium;
                if(medium >= 0)att *= medium_sample(S.media[medium], si.t, g, med_event, med_dist);
            }
            float r[8]; float rr;
            if(INTEG == PT_INTEGRATOR_SIMPLE){
                r[0] = (float)rnd_double(g); r[1] = (float)rnd_double(g); r[4] = (float)rnd_double(g); rr = (float)rnd_double(g);
                r[2] = r[3] = r[5] = r[6] = r[7] = 0.0f;
            } else{
#pragma unroll
                for(int k = 0; k < 8; k++)r[k] = rnd_x8(g);                       // Sampler::get2Dx4f
                rr = INTEG == PT_INTEGRATOR_VOLPATH ? (float)rnd_double(g) : r[6];
            }
            bool alive = true, skip_rr = false;
            MatCtx mc; mc.kind = PT_MAT_CONDUCTOR; mc.ri = 1.0f; mc.color = v3(0.0f); mc.rough = 0.0f; mc.metallic = 0.0f;
            if(INTEG == PT_INTEGRATOR_VOLPATH && med_event){
                // medium event (Integrators.cpp:356-362)
                const DevMedium& M = S.media[medium];
                const V3 mp = ro + med_dist * rd;
                sample_ld<INTEG>(S, si, mc, true, mp, rd, ntime, M.g, medium, r[5], r[2], r[3], g, B.tr_walk != 0, ne);
                if(ne.want)ne.contrib = att * ne.contrib;
                outp += att * ld3(M.Le);
                no = mp; nd = hg_sample(M.g, rd, r[6], r[7]);
                medium = dot(nd, si.n) < 0.0f ? si.medium : -1;                  // interaction.getMedium(scattered)
                spec = false;
            } else{
                // emitted radiance with MIS against NEE (Integrators.cpp:153-156, 217-226, 364-373)
                if(si.light >= 0){
                    const DevLight& L = S.lights[si.light];
                    const V3 le = area_L(S, L, si.n, si.ns, si.u, si.v, rd);      // AreaLight::L
                    if(!is_zero(le)){
                        if(INTEG == PT_INTEGRATOR_SIMPLE || spec)outp += att * le;
                        else if(prev_pdf > 0.0f){
                            // PMF and PDF of interaction.AreaLight: the inner, untransformed light, world-space p / n
                            const float lp = L.pmf_hit * area_pdf(L, si.p, si.n, ro, rd);
                            outp += att * le * (prev_pdf * prev_pdf / (prev_pdf * prev_pdf + lp * lp));
                        }
                    }
                }
                if(INTEG == PT_INTEGRATOR_VOLPATH)spec = false;                   // Integrators.cpp:374
                if(si.mat < 0){
                    // medium boundary (Integrators.cpp:158-161, 228-232, 378-382): continue without RR
                    if(INTEG == PT_INTEGRATOR_PATH)spec = true;
                    no = ro + si.t * rd; nd = rd;
                    if(INTEG == PT_INTEGRATOR_VOLPATH)medium = dot(rd, si.n) < 0.0f ? si.medium : -1;
                    skip_rr = true;
                } else{
                    Scatter sc;
                    mat_fetch(S, si, mc);
                    if(!mat_scatter(mc, si, ro, rd, r[4], r[0], r[1], sc))alive = false;   // absorbed
                    else{
                        if(INTEG == PT_INTEGRATOR_VOLPATH){
                            int nm = dot(sc.d, si.n) < 0.0f ? si.medium : -1;
                            if(!(sc.flags & BX_TRANSMISSIVE) && dot(rd, si.ns) <= 0.0f)nm = medium;   // Integrators.cpp:389-394
                            medium = nm;
                        }
                        if(INTEG != PT_INTEGRATOR_SIMPLE){
                            spec = (sc.flags & BX_SPECULAR) != 0;
                            if(!spec){
                                sample_ld<INTEG>(S, si, mc, false, si.p, rd, ntime, 0.0f, ray_medium, r[5], r[2], r[3], g, B.tr_walk != 0, ne);
                                if(ne.want)ne.contrib = att * ne.contrib;
                                prev_pdf = mat_pdf(mc, si, rd, sc.d);
                            }
                        }
                        att *= sc.f * (fabsf(dot(si.ns, sc.d)) / sc.pdf);
                        no = sc.o; nd = sc.d;
                    }
                }
            }
            if(alive && !skip_rr){
                // Russian roulette (Integrators.cpp:171-175, 249-253, 411-416)
                if(rr_depth++ > 3){
                    const float q = fminf(0.95f, fmaxf(fmaxf(att.x, att.y), att.z));
                    if(rr >= q)alive = false;
                    else att = att / q;
                }
            }
            // loop condition of the next iteration: depth++ < maxDepth && sum(attenuation) > 0
            if(alive && (int)depth < B.max_depth && (att.x + att.y + att.z) > 0.0f){ cont = true; depth++; }
        }
        nee = ne.want;
        *ps_at(A, slot, 0) = make_float4(att.x, att.y, att.z, prev_pdf);
        *ps_at(A, slot, 1) = make_float4(outp.x, outp.y, outp.z, 0.0f);
        *ps_at(A, slot, 2) = make_float4(__uint_as_float((uint32_t)g.state), __uint_as_float((uint32_t)(g.state >> 32)), __uint_as_float(depth | (rr_depth << 16)),
                            __uint_as_float((spec ? 1u : 0u) | ((uint32_t)(medium + 1) << 1)));
    }
    const uint32_t qn = queue_slot(&W.n_next, cont);
    if(cont){
        nq_rays[2 * (size_t)qn] = make_float4(no.x, no.y, no.z, CUDART_INF_F);
        nq_rays[2 * (size_t)qn + 1] = make_float4(nd.x, nd.y, nd.z, ntime);
        nq_slot[qn] = slot;
    }
    const bool ended = j < n && !cont;
    const uint32_t qd = queue_slot(&W.n_done, ended);
    if(ended)A.done[qd] = slot;
    const uint32_t qs = queue_slot(&W.n_sh, nee);
    if(nee){
        A.srays[2 * (size_t)qs] = make_float4(ne.o.x, ne.o.y, ne.o.z, ne.tmax);
        A.srays[2 * (size_t)qs + 1] = make_float4(ne.d.x, ne.d.y, ne.d.z, ntime);
        A.sslot[qs] = slot;
        A.scontrib[qs] = make_float4(ne.contrib.x, ne.contrib.y, ne.contrib.z, __int_as_float(ne.medium));
    }
    }
}

// ---- host driver --------------------------------------------------------------------------------

static int make_filter(const pt_render_params* rp, FilterParams& F){
    F.kind = rp->filter; F.radx = rp->filter_radius[0]; F.rady = rp->filter_radius[1];
    F.p0 = rp->filter_param[0]; F.p1 = rp->filter_param[1]; F.gx = F.gy = 0.0;
    if(!(F.radx > 0.0f) || !(F.rady > 0.0f)){ set_error("filter radius must be positive"); return PT_ERR_INVALID; }
    F.rx = (int)std::ceil(F.radx - 0.5f); F.ry = (int)std::ceil(F.rady - 0.5f);     // FilmTile ctor, Film.hpp:58
    double integral = 1.0;
    switch(rp->filter){
    case PT_FILTER_BOX: integral = 4.0 * F.radx * F.rady; break;
    case PT_FILTER_GAUSSIAN: {
        const double s = F.p0;
        if(!(s > 0)){ set_error("gaussian sigma must be positive"); return PT_ERR_INVALID; }
        auto G = [&](double x){ return 0.5641895835477563 / (s * 1.4142135623730951) * std::exp(-(x * x) / (2 * s * s)); };
        auto GI = [&](double x0, double x1){ const double r2 = s * 1.4142135623730951; return 0.5 * (std::erf(-x0 / r2) - std::erf(-x1 / r2)); };
        F.gx = G(F.radx); F.gy = G(F.rady);
        integral = (GI(-F.radx, F.radx) - 2 * F.radx * F.gx) * (GI(-F.rady, F.rady) - 2 * F.rady * F.gy);
        break;
    }
    case PT_FILTER_MITCHELL: integral = (double)F.radx * F.rady / 4.0; break;
    case PT_FILTER_LANCZOS: {
        // LanczosFilter::Integral (Filter.hpp:126-139): 256x256 strata over [-r, r]^2 with "area" 2*rx*ry;
        // strata centres here instead of an unseeded jitter
        if(!(F.p0 > 0)){ set_error("lanczos tau must be positive"); return PT_ERR_INVALID; }
        auto sinc = [](double x){ const double pi = 3.141592653589793; return (1.0 - x * x == 1.0) ? 1.0 : std::sin(pi * x) / (pi * x); };
        auto ws = [&](double x, double r){ return std::fabs(x) > r ? 0.0 : sinc(x) * sinc(x / F.p0); };
        double acc = 0;
        for(int y = 0; y < 256; y++)for(int x = 0; x < 256; x++){
            const float ux = (float)((x + 0.5) / 256.0), uy = (float)((y + 0.5) / 256.0);
            const float px = -F.radx + 2 * F.radx * ux, py = -F.rady + 2 * F.rady * uy;
            acc += ws(px, F.radx) * ws(py, F.rady);
        }
        integral = 2.0 * F.radx * F.rady * acc / 65536.0;
        break;
    }
    default: set_error("unknown filter"); return PT_ERR_INVALID;
    }
    F.inv_integral = 1.0 / integral;
    return PT_OK;
}

static bool carve(char*& p, void** out, size_t bytes){ *out = p; p += (bytes + 255) & ~(size_t)255; return true; }

static size_t arena_layout(char* base, Arena& A, size_t cap){
    char* p = base;
    carve(p, (void**)&A.rays[0], cap * 32); carve(p, (void**)&A.rays[1], cap * 32);
    carve(p, (void**)&A.qslot[0], cap * 4); carve(p, (void**)&A.qslot[1], cap * 4);
    carve(p, (void**)&A.hits, cap * 16); carve(p, (void**)&A.hinst, cap * 4);
    carve(p, (void**)&A.srays, cap * 32); carve(p, (void**)&A.sslot, cap * 4); carve(p, (void**)&A.scontrib, cap * 16); carve(p, (void**)&A.socc, cap);
    carve(p, (void**)&A.ps, cap * 64); carve(p, (void**)&A.done, cap * 4);
    carve(p, (void**)&A.skey, cap); carve(p, (void**)&A.perm, cap * 4); carve(p, (void**)&A.shist, (size_t)kSortGrid * kSortKeys * 4 + 4096);
    carve(p, (void**)&A.ctl, 256);
    static_assert(sizeof(WaveCtl) <= 256, "WaveCtl outgrew its arena slot");
    return (size_t)(p - base);
}

// PTB_PROFILE=1: an event after every launch group, so the stream time of a frame can be split by kernel
// without a profiler (the intervals include the idle time before the group starts).
struct StageLog {
    bool on = false;
    cudaStream_t st = nullptr;
    std::vector<std::pair<int, cudaEvent_t>> marks;
    void mark(int tag){
        if(!on)return;
        cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, st);
        marks.push_back({ tag, e });
    }
    void report(){
        if(!on || marks.size() < 2)return;
        static const char* names[] = { "start", "generate", "trace_closest", "shade", "counters+sync", "trace_any+connect", "film", "adapt" };
        double acc[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
        for(size_t i = 1; i < marks.size(); i++){
            float ms = 0; cudaEventElapsedTime(&ms, marks[i - 1].second, marks[i].second);
            acc[marks[i].first] += ms;
        }
        fprintf(stderr, "[ptb] stages:");
        for(int k = 1; k < 8; k++)if(acc[k] > 0)fprintf(stderr, " %s %.3f ms;", names[k], acc[k]);
        fprintf(stderr, "\n");
        for(auto& m : marks)cudaEventDestroy(m.second);
        marks.clear();
    }
};

int render_passes(pt_scene* s, const pt_camera_desc* cam, const pt_render_params* rp, int clear, pt_render_stats* stats){
    const int W = rp->width, H = rp->height;
    if(W <= 0 || H <= 0 || W > 65535 || H > 65535){ set_error("bad resolution"); return PT_ERR_INVALID; }   // Film stores uint16_t (Film.hpp:266-267)
    if(rp->spp <= 0 || rp->max_depth < 0 || rp->pass_stride <= 0 || rp->pass_begin < 0){ set_error("bad spp / depth / pass range"); return PT_ERR_INVALID; }
    if(rp->max_depth > 65535){ set_error("max_depth above 65535: the path state keeps depth in 16 bits"); return PT_ERR_INVALID; }
    if(rp->integrator < PT_INTEGRATOR_SIMPLE || rp->integrator > PT_INTEGRATOR_VOLPATH){ set_error("unknown integrator"); return PT_ERR_INVALID; }
    if(rp->sampler != 0 && rp->sampler != 1){ set_error("unknown sampler"); return PT_ERR_INVALID; }
    if(rp->sampler == 1 && (rp->strata_x <= 0 || rp->strata_y <= 0 || rp->strata_x * rp->strata_y != rp->spp)){
        set_error("StratifiedSampler: spp must equal strata_x * strata_y (Sampler.hpp:84)"); return PT_ERR_INVALID; }
    const bool sharded_adaptive = rp->adaptive && (rp->pass_begin != 0 || rp->pass_stride != 1);
    if(sharded_adaptive && !s->allreduce){
        set_error("adaptive sampling decides per pixel from all of its samples: with pass sharding the ranks must sum their per-pixel statistics every round (pt_scene_set_allreduce)"); return PT_ERR_INVALID; }
    if(cam->medium >= (int)s->media.size()){ set_error("camera medium id out of range"); return PT_ERR_INVALID; }
    FilterParams F;
    int rc = make_filter(rp, F);
    if(rc != PT_OK)return rc;
    CamParams C;
    make_camera(cam, W, H, C);
    cudaStream_t st = (cudaStream_t)s->stream;
    DeviceBuffers& b = s->buf;
    const size_t film_bytes = sizeof(double) * 4 * (size_t)W * H;
    if(b.film && (s->film_w != W || s->film_h != H)){ cudaFree(b.film); b.film = nullptr; }
    if(!b.film){
        if(b.film_sq){ cudaFree(b.film_sq); b.film_sq = nullptr; }
        if(!cuda_ok(cudaMalloc(&b.film, film_bytes), "cudaMalloc film") || !cuda_ok(cudaMalloc(&b.film_sq, film_bytes / 4 * 3), "cudaMalloc film_sq"))return PT_ERR_CUDA;
        clear = 1;
    }
    s->film_w = W; s->film_h = H;
    if(clear && (!cuda_ok(cudaMemsetAsync(b.film, 0, film_bytes, st), "film clear") ||
                 !cuda_ok(cudaMemsetAsync(b.film_sq, 0, film_bytes / 4 * 3, st), "film clear")))return PT_ERR_CUDA;

    // passes of this call: pass_begin, pass_begin + stride, ... < spp
    const uint64_t n_pass = rp->pass_begin >= rp->spp ? 0 : (uint64_t)(rp->spp - rp->pass_begin + rp->pass_stride - 1) / rp->pass_stride;
    const uint64_t total = n_pass * (uint64_t)W * H;
    // Slot pool: the larger, the fewer wavefront iterations and the longer each kernel runs before its tail
    // (C3 / C4 on one B200: 4 Mi slots 376 / 329 Msamples/s, 16 Mi 416 / 347, 64 Mi 432 / 364).  64 Mi slots are 16 GB
    // of a 180 GB device; the pool never takes more than a quarter of what is free.  PTB_BATCH sets it by hand.
    size_t cap = 64u << 20;
    Arena A;
    if(const char* e = getenv("PTB_BATCH")){ const long long v = atoll(e); if(v >= 1024)cap = (size_t)v; }
    else{
        size_t free_b = 0, total_b = 0;
        if(cudaMemGetInfo(&free_b, &total_b) != cudaSuccess){ cudaGetLastError(); free_b = 0; }
        free_b += b.work_bytes;                  // what this scene already holds can be reused
        while(cap > (4u << 20) && arena_layout(nullptr, A, cap) > free_b / 4)cap >>= 1;
    }
    if(cap > total)cap = (size_t)std::max<uint64_t>(total, 1);
    const size_t need = arena_layout(nullptr, A, cap);
    if(need > b.work_bytes){
        if(b.work){ cudaFree(b.work); b.work = nullptr; b.work_bytes = 0; }
        if(!cuda_ok(cudaMalloc(&b.work, need), "cudaMalloc wavefront arena"))return PT_ERR_NOMEM;
        b.work_bytes = need;
    }
    arena_layout((char*)b.work, A, cap);
    // adaptive loop state: estimator sums (6 doubles / pixel), two pixel lists, a counter
    const size_t npix = (size_t)W * H;
    double* est = nullptr; double* est_round = nullptr; uint32_t* lists[2] = { nullptr, nullptr }; uint32_t* list_n = nullptr;
    if(rp->adaptive){
        const size_t bytes = npix * (48 + 8 + (sharded_adaptive ? 48 : 0)) + 256;
        if(bytes > b.adapt_bytes){
            if(b.adapt){ cudaFree(b.adapt); b.adapt = nullptr; b.adapt_bytes = 0; }
            if(!cuda_ok(cudaMalloc(&b.adapt, bytes), "cudaMalloc adaptive state"))return PT_ERR_NOMEM;
            b.adapt_bytes = bytes;
        }
        est = (double*)b.adapt;
        lists[0] = (uint32_t*)(est + 6 * npix); lists[1] = lists[0] + npix; list_n = lists[1] + npix;
        if(!cuda_ok(cudaMemsetAsync(est, 0, 48 * npix, st), "estimator clear"))return PT_ERR_CUDA;
        if(sharded_adaptive){
            // this rank's sums of the current round; after the all-reduce they are added to `est`, which
            // therefore holds the same global totals on every rank
            est_round = (double*)((char*)b.adapt + npix * (48 + 8) + 256);
            if(!cuda_ok(cudaMemsetAsync(est_round, 0, 48 * npix, st), "estimator clear"))return PT_ERR_CUDA;
        }
    }
    // pinned once per scene: cudaMallocHost / cudaFreeHost per frame cost milliseconds (more with other
    // large pinned buffers mapped in the process) and synchronise the device
    constexpr int kRing = 8;                 // watch slots: copies of WaveCtl the host reads a few bounces late
    if(!b.h_counters && !cuda_ok(cudaMallocHost(&b.h_counters, 256 * (kRing + 2)), "pinned counters"))return PT_ERR_CUDA;
    WaveCtl* h_ring = (WaveCtl*)b.h_counters;                            // slot k at byte 256 k
    auto h_slot = [&](int k){ return (WaveCtl*)((char*)b.h_counters + 256 * k); };
    uint32_t* h_words = (uint32_t*)((char*)b.h_counters + 256 * kRing);  // adaptive list size
    (void)h_ring;
    for(int k = 0; k < kRing; k++)if(!b.wave_events[k] && !cuda_ok(cudaEventCreateWithFlags((cudaEvent_t*)&b.wave_events[k], cudaEventDisableTiming), "cudaEventCreate"))return PT_ERR_CUDA;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if(!cuda_ok(cudaEventCreate(&e0), "cudaEventCreate"))return PT_ERR_CUDA;
    if(!cuda_ok(cudaEventCreate(&e1), "cudaEventCreate")){ cudaEventDestroy(e0); return PT_ERR_CUDA; }
    auto fail = [&](int code){ cudaEventDestroy(e0); cudaEventDestroy(e1); return code; };
    cudaEventRecord(e0, st);
    pt_render_stats R {};
    const unsigned long long overflow0 = s->overflow_seen;   // traversal-stack overflows are counted per scene; this call fails if the count moves

    // persistent grids: every kernel of the loop reads its batch size from WaveCtl
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->device);
    const unsigned g_shade = (unsigned)sms * PTB_SHADE_MINB, g_256 = (unsigned)sms * 8, g_tr = (unsigned)sms * 8;

    // one set of `count` camera samples through the wavefront.  The slot pool stays full: the slots whose
    // paths ended in a shade pass are splatted and handed to the next camera samples straight away, so the
    // long tail of deep paths is paid once per call and not once per batch.
    double host_wait_s = 0.0; unsigned iterations = 0;
    const bool tr_walk = rp->integrator == PT_INTEGRATOR_VOLPATH && s->has_boundary;
    // shade-order sort (k_sort_*).  Measured on one B200 (PTB_PROFILE, shade stage incl. the sort, round 2 build): C4 (VolPath: medium
    // events, textured materials) 325 -> 170 ms, C3 (PathIntegrator, four material kinds) 67.9 -> 38.5 ms (frame 291 -> 260 ms); C1 (one
    // material kind) 1.89 -> 2.82 ms (round 1).  So: on for VolPathIntegrator and for every scene whose materials are of more than one
    // kind; PTB_SORT=0 / 1 forces it, PTB_SORT_MIN sets the smallest queue worth sorting.
    static const int sort_env = []{ const char* e = getenv("PTB_SORT"); return e ? (atoi(e) != 0 ? 1 : 0) : -1; }();
    unsigned kinds = 0;
    for(const pt_material_desc& md : s->materials)kinds |= 1u << (md.kind & 31);
    const bool sort_shade = sort_env >= 0 ? sort_env == 1 : (rp->integrator == PT_INTEGRATOR_VOLPATH || (kinds & (kinds - 1)) != 0);
    static const uint32_t sort_min_env = []{ const char* e = getenv("PTB_SORT_MIN"); return e ? (uint32_t)atoll(e) : 65536u; }();
    const uint32_t sort_min = sort_shade ? sort_min_env : 0xFFFFFFFFu;      // the kernels compare the device-side queue size with it
    // how many bounces the host may enqueue beyond the last queue size it has seen (PTB_LOOKAHEAD)
    static const int lookahead = []{ const char* e = getenv("PTB_LOOKAHEAD"); const int v = e ? atoi(e) : 2; return v < 1 ? 1 : (v > kRing - 2 ? kRing - 2 : v); }();
    StageLog log; log.on = getenv("PTB_PROFILE") != nullptr; log.st = st;
    log.mark(0);
    const auto t_host0 = std::chrono::steady_clock::now();
    bool first_run = true;
    auto launch_ok = [&](const char* what){ return cuda_ok(cudaGetLastError(), what); };
    auto run = [&](uint64_t count, const uint32_t* list, uint32_t n_list, uint32_t sample_base) -> int {
        BatchInfo B;
        B.W = W; B.H = H; B.pass_begin = rp->pass_begin; B.pass_stride = rp->pass_stride; B.max_depth = rp->max_depth; B.seed = rp->seed;
        B.sampler = rp->sampler; B.sx = rp->strata_x; B.sy = rp->strata_y;
        B.list = list; B.n_list = n_list; B.sample_base = sample_base;
        B.tr_walk = tr_walk ? 1 : 0;
        if(count == 0)return PT_OK;                   // more ranks than samples per round
        // a run that fits the pool never regenerates a slot: queue order stays slot order -> planes; else records
        if(count <= cap){ A.ps_slot = 1; A.ps_quarter = (uint32_t)cap; } else{ A.ps_slot = 4; A.ps_quarter = 1; }
        double* est_now = est_round ? est_round : est;
        if(rp->max_depth <= 0){
            // `depth++ < maxDepth` fails at once: black samples that still carry their filter weight
            for(uint64_t gid = 0; gid < count; gid += cap){
                k_wave_init<<<1, 1, 0, st>>>(A, gid, count, cap, first_run ? 1 : 0);
                first_run = false;
                k_generate<<<g_256, 256, 0, st>>>(A, B, C);
                k_finish<<<g_256, 256, 0, st>>>(A, B, F, C, (double*)b.film, (double*)b.film_sq, est_now, 1);
                R.kernel_launches += 3;
            }
            return launch_ok("black frame") ? PT_OK : PT_ERR_CUDA;
        }
        k_wave_init<<<1, 1, 0, st>>>(A, 0, count, cap, first_run ? 1 : 0);
        first_run = false;
        k_generate<<<g_256, 256, 0, st>>>(A, B, C);
        R.kernel_launches += 2;
        log.mark(1);
        DynBatch dq; dq.n_dev = &A.ctl->n_cur; dq.sel_dev = &A.ctl->cur; dq.sel_stride = (size_t)(A.rays[1] - A.rays[0]);
        DynBatch ds; ds.n_dev = &A.ctl->n_sh;
        ConnectAcc acc; acc.slot = A.sslot; acc.contrib = A.scontrib; acc.ps = A.ps; acc.ps_slot = A.ps_slot; acc.ps_off = A.ps_quarter;
        // A batch that fits the pool ends after max_depth bounces at the latest and that is all the host needs to know
        // when the depth is small; otherwise the host watches n_cur `lookahead` bounces late and stops once it reads 0.
        const bool blind = count <= cap && rp->max_depth <= 12;
        for(unsigned it = 0;; it++){
            if(blind && it >= (unsigned)rp->max_depth)break;
            int rc = launch_trace_closest_dyn(s, A.rays[0], dq, A.hits, A.hinst, &A.ctl->cursor[0], st);
            if(rc != PT_OK)return rc;
            R.kernel_launches++;
            log.mark(2);
            if(sort_shade){
                k_sort_count<<<kSortGrid, kSortBlock, 0, st>>>(A, s->shade, sort_min);
                k_sort_scan<<<1, 1024, 0, st>>>(A, sort_min);
                k_sort_scatter<<<kSortGrid, kSortBlock, 0, st>>>(A, sort_min);
                R.kernel_launches += 3;
            }
            switch(rp->integrator){
            case PT_INTEGRATOR_SIMPLE: k_shade<PT_INTEGRATOR_SIMPLE><<<g_shade, 128, 0, st>>>(A, s->shade, B, sort_min); break;
            case PT_INTEGRATOR_PATH: k_shade<PT_INTEGRATOR_PATH><<<g_shade, 128, 0, st>>>(A, s->shade, B, sort_min); break;
            default: k_shade<PT_INTEGRATOR_VOLPATH><<<g_shade, 128, 0, st>>>(A, s->shade, B, sort_min); break;
            }
            R.kernel_launches++;
            log.mark(3);
            if(rp->integrator != PT_INTEGRATOR_SIMPLE){
                if(tr_walk){
                    if(s->has_alpha_tex || s->has_wrappers)launch_shadow_tr(A, s, g_tr, st, AlphaTex { s->shade, s->alpha_salt += 0x9E3779B97F4A7C15ull });
                    else launch_shadow_tr(A, s, g_tr, st, NoAlpha());
                    R.kernel_launches++;
                } else{
                    rc = launch_trace_any_dyn(s, A.srays, ds, nullptr, acc, &A.ctl->cursor[1], st);
                    if(rc != PT_OK)return rc;
                    R.kernel_launches++;
                }
                log.mark(5);
            }
            // after the shadow rays: a path may end in the very pass that queued its last one
            k_finish<<<g_256, 256, 0, st>>>(A, B, F, C, (double*)b.film, (double*)b.film_sq, est_now, 0);
            R.kernel_launches++;
            log.mark(6);
            if(!launch_ok("wavefront bounce"))return PT_ERR_CUDA;
            iterations++;
            if(blind)continue;
            // watch slot of this bounce: WaveCtl as k_finish left it, and an event behind the copy
            const int k = (int)(it % kRing);
            if(!cuda_ok(cudaMemcpyAsync(h_slot(k), A.ctl, sizeof(WaveCtl), cudaMemcpyDeviceToHost, st), "WaveCtl D2H") ||
               !cuda_ok(cudaEventRecord((cudaEvent_t)b.wave_events[k], st), "event"))return PT_ERR_CUDA;
            if(it >= (unsigned)lookahead){
                const int w = (int)((it - lookahead) % kRing);
                const auto tw0 = std::chrono::steady_clock::now();
                if(!cuda_ok(cudaEventSynchronize((cudaEvent_t)b.wave_events[w]), "wavefront watch"))return PT_ERR_CUDA;
                host_wait_s += std::chrono::duration<double>(std::chrono::steady_clock::now() - tw0).count();
                if(h_slot(w)->n_cur == 0)break;       // the queue ran dry `lookahead` bounces ago: what was enqueued since found nothing to do
            }
        }
        return PT_OK;
    };

    if(s->tile_world > 1 && !rp->adaptive){
        // Interleaved tile sharding (the split for frames with fewer passes than GPUs): this rank renders ALL the
        // passes of the 32 x 32 tiles whose index is congruent to its rank; the film reduce adds the disjoint tile sets.
        // The pixel list rides the mechanism of the adaptive rounds (BatchInfo::list).
        const int tw = (W + 31) / 32;
        std::vector<uint32_t> mine;
        for(int y = 0; y < H; y++)for(int x = 0; x < W; x++)if(((y / 32) * tw + x / 32) % s->tile_world == s->tile_rank)mine.push_back((uint32_t)(y * W + x));
        if(mine.size() * 4 > b.tile_list_bytes){
            if(b.tile_list){ cudaFree(b.tile_list); b.tile_list = nullptr; b.tile_list_bytes = 0; }
            if(!mine.empty() && !cuda_ok(cudaMalloc(&b.tile_list, mine.size() * 4), "cudaMalloc tile list"))return fail(PT_ERR_NOMEM);
            b.tile_list_bytes = mine.size() * 4;
        }
        if(!mine.empty()){
            if(!cuda_ok(cudaMemcpyAsync(b.tile_list, mine.data(), mine.size() * 4, cudaMemcpyHostToDevice, st), "tile list") || !cuda_ok(cudaStreamSynchronize(st), "tile list"))return fail(PT_ERR_CUDA);
            rc = run((uint64_t)mine.size() * n_pass, (const uint32_t*)b.tile_list, (uint32_t)mine.size(), 0);
        } else rc = PT_OK;
    } else rc = run(total, nullptr, 0, 0);
    if(rc != PT_OK)return fail(rc);
    // sharded adaptive rounds: sum this round's statistics over the ranks, fold them into the totals
    auto merge_round = [&]() -> int{
        if(!est_round)return PT_OK;
        if(!cuda_ok(cudaStreamSynchronize(st), "adaptive round"))return PT_ERR_CUDA;
        if(s->allreduce(s->allreduce_user, est_round, 6 * npix, st) != 0){ set_error("all-reduce callback failed"); return PT_ERR_STATE; }
        k_est_merge<<<(unsigned)((6 * npix + 255) / 256), 256, 0, st>>>(est, est_round, 6 * npix);
        R.kernel_launches++;
        return PT_OK;
    };
    if(rp->adaptive){
        // rounds 2..128 over the pixels the stopping rule keeps (Integrators.cpp:59,82-85)
        const uint32_t* in = nullptr;
        uint32_t n_in = (uint32_t)npix;
        for(int round = 1; round < 128 && n_in > 0; round++){
            uint32_t* out = lists[round & 1];
            if(!cuda_ok(cudaMemsetAsync(list_n, 0, 4, st), "list reset"))return fail(PT_ERR_CUDA);
            rc = merge_round();
            if(rc != PT_OK)return fail(rc);
            k_adapt<<<(n_in + 255) / 256, 256, 0, st>>>(est, in, n_in, out, list_n, (double)round * rp->spp);
            R.kernel_launches++;
            if(!cuda_ok(cudaMemcpyAsync(h_words, list_n, 4, cudaMemcpyDeviceToHost, st), "list size D2H") ||
               !cuda_ok(cudaStreamSynchronize(st), "adapt"))return fail(PT_ERR_CUDA);
            n_in = h_words[0]; in = out;
            if(n_in == 0)break;
            rc = run((uint64_t)n_in * n_pass, in, n_in, (uint32_t)round * (uint32_t)rp->spp);
            if(rc != PT_OK)return fail(rc);
        }
    }
    cudaEventRecord(e1, st);
    if(!cuda_ok(cudaMemcpyAsync(h_slot(kRing - 1), A.ctl, sizeof(WaveCtl), cudaMemcpyDeviceToHost, st), "WaveCtl D2H") ||
       !cuda_ok(cudaMemcpyAsync(h_words + 8, (unsigned long long*)b.stats + 8, 8, cudaMemcpyDeviceToHost, st), "overflow counter"))return fail(PT_ERR_CUDA);
    if(!cuda_ok(cudaStreamSynchronize(st), "render"))return fail(PT_ERR_CUDA);
    const WaveCtl& end = *h_slot(kRing - 1);
    R.nonfinite_samples = end.nonfinite;
    R.samples = end.samples; R.closest_rays = end.closest_rays; R.shadow_rays = end.shadow_rays;
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    R.gpu_ms = ms;
    if(getenv("PTB_PROFILE")){
        const double host_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_host0).count();
        log.report();
        fprintf(stderr, "[ptb] render: %.3f ms on the stream, host %.3f ms of which %.3f ms waiting for the watch slots, %u bounces enqueued (%llu ran), %llu launches\n",
                ms, host_s * 1e3, host_wait_s * 1e3, iterations, (unsigned long long)end.bounces, (unsigned long long)R.kernel_launches);
    }
    if(stats)*stats = R;
    s->overflow_seen = *(unsigned long long*)(h_words + 8);
    if(s->overflow_seen != overflow0){
        set_error("traversal stack overflow: " + std::to_string(*(unsigned long long*)(h_words + 8) - overflow0) + " rays needed more than " + std::to_string(kStack) +
                  " stack entries; their hits may be wrong");
        return fail(PT_ERR_OVERFLOW);
    }
    return fail(PT_OK);
}

} // namespace ptb
