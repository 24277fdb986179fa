// bvh_build_gpu.cu — the binned-SAH BVH2 build of BVHBase::BuildBaseThreaded (BVH.hpp:290-390) on the
// GPU.  Not LBVH: every decision of the reference builder is taken with the same
// float operations in the same order (8/16/32 bins by span, `cost < best` in axis-then-bin order, leaf when
// span <= 2 or bestCost >= area*span or the partition is degenerate), and the in-place partition
// reproduces the visiting order of a bidirectional std::partition, so the tree AND the item order inside
// leaves are those of the host builder (bvh_build.cpp) — flatten_bvh() turns either into byte-identical
// BVH2 / BVH4 / BVH8 arrays.  min / max / integer counts are order-independent, which is what makes a
// parallel build exact here.
//
// Two phases:
//  * large nodes (span > kSmall): level-synchronous, every node cut into 2048-item chunks, one CTA per
//    chunk: bounds -> bins (shared-memory atomics, flushed to the node's global bins) -> split decision ->
//    count -> partition.  ~20 levels x 7 small launches for 1 M items.
//  * small subtrees (span <= kSmall): one CTA builds the whole subtree with its items in shared memory;
//    nodes wider than a warp are processed by the CTA, nodes of <= 32 items by single warps (ballot /
//    __fns partition, per-warp bins).
#include <cuda_runtime.h>
#include <cfloat>
#include <cstdio>
#include <cstring>

#include "bvh_build.hpp"
#include "scene_host.hpp"

namespace ptb {

namespace {

struct GNode {            // same layout as BuildNode
    float lo[3], hi[3];
    uint32_t left, right, count, axis;
};
static_assert(sizeof(GNode) == sizeof(BuildNode), "GNode must match BuildNode");
static_assert(sizeof(GNode) == 40, "BuildNode layout");

constexpr int kSmall = 2048;        // largest subtree one CTA builds in shared memory
constexpr int kChunk = 2048;        // items per CTA in the large phase
constexpr int kThreads = 256;
constexpr int kPer = kChunk / kThreads;
constexpr int kWarps = kThreads / 32;
constexpr int kWarpList = 256;      // warp-task list of the subtree kernel

// ---- float <-> ordered int (atomicMin / atomicMax on floats) -------------------------
__device__ __forceinline__ int f2o(float f){ const int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7FFFFFFF; }
__device__ __forceinline__ float o2f(int i){ return __int_as_float(i >= 0 ? i : i ^ 0x7FFFFFFF); }

__device__ __forceinline__ float half_area(const float lo[3], const float hi[3]){
    const float ex = __fsub_rn(hi[0], lo[0]), ey = __fsub_rn(hi[1], lo[1]), ez = __fsub_rn(hi[2], lo[2]);
    return __fadd_rn(__fadd_rn(__fmul_rn(ex, ey), __fmul_rn(ey, ez)), __fmul_rn(ez, ex));   // AABB::Area, AABB.hpp:126-129
}
__device__ __forceinline__ float centroid(float lo, float hi){ return __fmul_rn(0.5f, __fadd_rn(hi, lo)); }   // BVH.hpp:89
__device__ __forceinline__ uint32_t bins_for(uint32_t span){ return span >= 1024u ? 32u : span >= 64u ? 16u : 8u; }   // BVH.hpp:316
__device__ __forceinline__ int bin_of(float c, float mn, float scale, int nbins){
    int id = (int)__fmul_rn(__fsub_rn(c, mn), scale);
    return id > nbins - 1 ? nbins - 1 : id;
}

struct Bin { int lo[3], hi[3]; uint32_t n, pad; };      // ordered-int boxes, 32 B

struct LTask {
    uint32_t first, last, node;
    uint32_t chunk0, nchunks;
    int blo[3], bhi[3], cmin[3], cmax[3];               // ordered ints
    int axis; float pos; int leaf;
    uint32_t good, m;
};

struct Chunk { uint32_t task, begin, end, local; };
struct SmallTask { uint32_t first, last, node, node_base; };

struct Counters {
    uint32_t ntasks[2];      // ping-pong level task counts
    uint32_t nchunks;
    uint32_t nsmall;
    uint32_t nnodes;
    uint32_t pad[3];
};

// candidate split (axis, bin) evaluated by one thread; bins as decoded floats in shared memory
struct SplitEval { float cost; int key; float pos; };

// cost of splitting after bin b on one axis (BVH.hpp:333-356).  binbox(i, lo, hi, n) reads bin i.
template <class BinRead>
__device__ __forceinline__ SplitEval eval_split(BinRead rd, int b, int nbins, uint32_t span, float mn, float mx, int key){
    const float inf = __int_as_float(0x7f800000);
    float llo[3] = { inf, inf, inf }, lhi[3] = { -inf, -inf, -inf }, rlo[3] = { inf, inf, inf }, rhi[3] = { -inf, -inf, -inf };
    uint32_t lsum = 0;
    for(int i = 0; i < nbins; i++){
        float lo[3], hi[3]; uint32_t n;
        rd(i, lo, hi, n);
        if(i <= b){
            lsum += n;
            for(int k = 0; k < 3; k++){ llo[k] = fminf(llo[k], lo[k]); lhi[k] = fmaxf(lhi[k], hi[k]); }
        } else{
            for(int k = 0; k < 3; k++){ rlo[k] = fminf(rlo[k], lo[k]); rhi[k] = fmaxf(rhi[k], hi[k]); }
        }
    }
    SplitEval e;
    e.cost = __fadd_rn(__fmul_rn(__uint2float_rn(lsum), half_area(llo, lhi)), __fmul_rn(__uint2float_rn(span - lsum), half_area(rlo, rhi)));
    const float scale = __fdiv_rn(__fsub_rn(mx, mn), __uint2float_rn((uint32_t)nbins));
    e.pos = __fadd_rn(mn, __fmul_rn(__uint2float_rn((uint32_t)(b + 1)), scale));
    e.key = key;
    if(!(e.cost < inf)){ e.cost = inf; e.key = 0x7fffffff; }     // NaN / inf never beats `best = inf` (BVH.hpp:350)
    return e;
}
__device__ __forceinline__ bool better(const SplitEval& a, const SplitEval& b){ return a.cost < b.cost || (a.cost == b.cost && a.key < b.key); }
__device__ __forceinline__ SplitEval warp_best(SplitEval e){
    for(int d = 16; d >= 1; d >>= 1){
        SplitEval o;
        o.cost = __shfl_xor_sync(0xffffffffu, e.cost, d); o.key = __shfl_xor_sync(0xffffffffu, e.key, d); o.pos = __shfl_xor_sync(0xffffffffu, e.pos, d);
        if(better(o, e))e = o;
    }
    return e;
}

// =====================================================================================
// Large phase
// =====================================================================================

__global__ void k_plan(LTask* tasks, Bin* bins, Chunk* chunks, Counters* cnt, int cur){
    __shared__ uint32_t s_scan[1024];
    __shared__ uint32_t s_base;
    const uint32_t nt = cnt->ntasks[cur];
    if(threadIdx.x == 0)s_base = 0;
    __syncthreads();
    for(uint32_t t0 = 0; t0 < nt; t0 += 1024){
        const uint32_t t = t0 + threadIdx.x;
        uint32_t nc = 0;
        if(t < nt)nc = (tasks[t].last - tasks[t].first + kChunk - 1) / kChunk;
        s_scan[threadIdx.x] = nc;
        __syncthreads();
        for(int d = 1; d < 1024; d <<= 1){
            uint32_t v = threadIdx.x >= d ? s_scan[threadIdx.x - d] : 0;
            __syncthreads();
            s_scan[threadIdx.x] += v;
            __syncthreads();
        }
        const uint32_t excl = s_base + s_scan[threadIdx.x] - nc;
        if(t < nt){
            LTask& T = tasks[t];
            T.chunk0 = excl; T.nchunks = nc;
            const int pinf = f2o(__int_as_float(0x7f800000)), ninf = f2o(__int_as_float(0xff800000));
            for(int k = 0; k < 3; k++){ T.blo[k] = pinf; T.bhi[k] = ninf; T.cmin[k] = f2o(FLT_MAX); T.cmax[k] = f2o(-FLT_MAX); }
            T.axis = 0; T.pos = 0.0f; T.leaf = 0; T.good = 0; T.m = 0;
            for(uint32_t c = 0; c < nc; c++){
                Chunk ch; ch.task = t; ch.begin = T.first + c * kChunk; ch.end = min(T.last, ch.begin + kChunk); ch.local = c;
                chunks[excl + c] = ch;
            }
        }
        __syncthreads();
        if(threadIdx.x == 1023)s_base += s_scan[1023];
        __syncthreads();
    }
    // reset the bins of this level's tasks
    const int pinf = f2o(__int_as_float(0x7f800000)), ninf = f2o(__int_as_float(0xff800000));
    for(uint32_t i = threadIdx.x; i < nt * 96u; i += 1024){
        Bin b;
        for(int k = 0; k < 3; k++){ b.lo[k] = pinf; b.hi[k] = ninf; }
        b.n = 0; b.pad = 0;
        bins[i] = b;
    }
    if(threadIdx.x == 0){ cnt->nchunks = s_base; cnt->ntasks[cur ^ 1] = 0; }
}

__global__ void __launch_bounds__(kThreads) k_bounds(const float4* __restrict__ ilo, const float4* __restrict__ ihi, const Chunk* __restrict__ chunks, LTask* tasks){
    const Chunk ch = chunks[blockIdx.x];
    const float inf = __int_as_float(0x7f800000);
    float v[12] = { inf, inf, inf, -inf, -inf, -inf, FLT_MAX, FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX };
    for(uint32_t i = ch.begin + threadIdx.x; i < ch.end; i += kThreads){
        const float4 lo = ilo[i], hi = ihi[i];
        const float l[3] = { lo.x, lo.y, lo.z }, h[3] = { hi.x, hi.y, hi.z };
        for(int k = 0; k < 3; k++){
            const float c = centroid(l[k], h[k]);
            v[k] = fminf(v[k], l[k]); v[3 + k] = fmaxf(v[3 + k], h[k]);
            v[6 + k] = fminf(v[6 + k], c); v[9 + k] = fmaxf(v[9 + k], c);
        }
    }
    for(int d = 16; d >= 1; d >>= 1)
        for(int k = 0; k < 12; k++){
            const float o = __shfl_xor_sync(0xffffffffu, v[k], d);
            v[k] = ((k / 3) & 1) ? fmaxf(v[k], o) : fminf(v[k], o);
        }
    __shared__ float s[kWarps][12];
    if((threadIdx.x & 31) == 0)for(int k = 0; k < 12; k++)s[threadIdx.x >> 5][k] = v[k];
    __syncthreads();
    if(threadIdx.x < 12){
        const int k = threadIdx.x;
        const bool mx = (k / 3) & 1;
        float r = s[0][k];
        for(int w = 1; w < kWarps; w++)r = mx ? fmaxf(r, s[w][k]) : fminf(r, s[w][k]);
        LTask& T = tasks[ch.task];
        int* dst = k < 3 ? &T.blo[k] : k < 6 ? &T.bhi[k - 3] : k < 9 ? &T.cmin[k - 6] : &T.cmax[k - 9];
        if(mx)atomicMax(dst, f2o(r)); else atomicMin(dst, f2o(r));
    }
}

__global__ void __launch_bounds__(kThreads) k_bin(const float4* __restrict__ ilo, const float4* __restrict__ ihi, const Chunk* __restrict__ chunks, const LTask* __restrict__ tasks, Bin* bins){
    const Chunk ch = chunks[blockIdx.x];
    const LTask& T = tasks[ch.task];
    __shared__ Bin sb[96];
    const int pinf = f2o(__int_as_float(0x7f800000)), ninf = f2o(__int_as_float(0xff800000));
    if(threadIdx.x < 96){
        Bin b;
        for(int k = 0; k < 3; k++){ b.lo[k] = pinf; b.hi[k] = ninf; }
        b.n = 0; b.pad = 0;
        sb[threadIdx.x] = b;
    }
    const uint32_t span = T.last - T.first;
    const int nbins = (int)bins_for(span);
    float mn[3], scale[3]; bool valid[3];
    for(int a = 0; a < 3; a++){
        const float mx = o2f(T.cmax[a]);
        mn[a] = o2f(T.cmin[a]);
        valid[a] = !(fabsf(__fsub_rn(mx, mn[a])) <= FLT_EPSILON);
        scale[a] = __fdiv_rn(__uint2float_rn((uint32_t)nbins), __fsub_rn(mx, mn[a]));
    }
    __syncthreads();
    for(uint32_t i = ch.begin + threadIdx.x; i < ch.end; i += kThreads){
        const float4 lo = ilo[i], hi = ihi[i];
        const float l[3] = { lo.x, lo.y, lo.z }, h[3] = { hi.x, hi.y, hi.z };
        int ol[3], oh[3];
        for(int k = 0; k < 3; k++){ ol[k] = f2o(l[k]); oh[k] = f2o(h[k]); }
        for(int a = 0; a < 3; a++){
            if(!valid[a])continue;
            Bin& b = sb[a * 32 + bin_of(centroid(l[a], h[a]), mn[a], scale[a], nbins)];
            atomicAdd(&b.n, 1u);
            for(int k = 0; k < 3; k++){ atomicMin(&b.lo[k], ol[k]); atomicMax(&b.hi[k], oh[k]); }
        }
    }
    __syncthreads();
    if(threadIdx.x < 96){
        const Bin& b = sb[threadIdx.x];
        if(b.n){
            Bin& g = bins[(size_t)ch.task * 96 + threadIdx.x];
            atomicAdd(&g.n, b.n);
            for(int k = 0; k < 3; k++){ atomicMin(&g.lo[k], b.lo[k]); atomicMax(&g.hi[k], b.hi[k]); }
        }
    }
}

// one CTA of 96 threads per task: thread (axis, b) evaluates the split after bin b
__global__ void __launch_bounds__(96) k_split(LTask* tasks, const Bin* __restrict__ bins, GNode* nodes){
    LTask& T = tasks[blockIdx.x];
    __shared__ float s_lo[96][3], s_hi[96][3];
    __shared__ uint32_t s_n[96];
    __shared__ SplitEval s_best[3];
    {
        const Bin b = bins[(size_t)blockIdx.x * 96 + threadIdx.x];
        for(int k = 0; k < 3; k++){ s_lo[threadIdx.x][k] = o2f(b.lo[k]); s_hi[threadIdx.x][k] = o2f(b.hi[k]); }
        s_n[threadIdx.x] = b.n;
    }
    __syncthreads();
    const uint32_t span = T.last - T.first;
    const int nbins = (int)bins_for(span);
    const int axis = threadIdx.x >> 5, b = threadIdx.x & 31;
    const float mn = o2f(T.cmin[axis]), mx = o2f(T.cmax[axis]);
    const bool valid = !(fabsf(__fsub_rn(mx, mn)) <= FLT_EPSILON);
    SplitEval e; e.cost = __int_as_float(0x7f800000); e.key = 0x7fffffff; e.pos = 0.0f;
    if(valid && b < nbins - 1){
        auto rd = [&](int i, float lo[3], float hi[3], uint32_t& n){
            for(int k = 0; k < 3; k++){ lo[k] = s_lo[axis * 32 + i][k]; hi[k] = s_hi[axis * 32 + i][k]; }
            n = s_n[axis * 32 + i];
        };
        e = eval_split(rd, b, nbins, span, mn, mx, axis * 32 + b);
    }
    e = warp_best(e);
    if(b == 0)s_best[axis] = e;
    __syncthreads();
    if(threadIdx.x == 0){
        SplitEval best = s_best[0];
        if(better(s_best[1], best))best = s_best[1];
        if(better(s_best[2], best))best = s_best[2];
        float lo[3], hi[3];
        for(int k = 0; k < 3; k++){ lo[k] = o2f(T.blo[k]); hi[k] = o2f(T.bhi[k]); }
        GNode& N = nodes[T.node];
        for(int k = 0; k < 3; k++){ N.lo[k] = lo[k]; N.hi[k] = hi[k]; }
        const float parent = __fmul_rn(half_area(lo, hi), __uint2float_rn(span));
        const bool none = best.key == 0x7fffffff;
        T.axis = none ? 0 : best.key >> 5;
        T.pos = none ? 0.0f : best.pos;
        T.leaf = (best.cost >= parent) ? 1 : 0;          // BVH.hpp:359
    }
}

__global__ void __launch_bounds__(kThreads) k_count(const float4* __restrict__ ilo, const float4* __restrict__ ihi, const Chunk* __restrict__ chunks, const LTask* __restrict__ tasks, uint32_t* chunk_good){
    const Chunk ch = chunks[blockIdx.x];
    const LTask& T = tasks[ch.task];
    if(T.leaf){ if(threadIdx.x == 0)chunk_good[blockIdx.x] = 0; return; }
    const int axis = T.axis; const float pos = T.pos;
    uint32_t n = 0;
    for(uint32_t i = ch.begin + threadIdx.x; i < ch.end; i += kThreads){
        const float4 lo = ilo[i], hi = ihi[i];
        const float c = axis == 0 ? centroid(lo.x, hi.x) : axis == 1 ? centroid(lo.y, hi.y) : centroid(lo.z, hi.z);
        n += c <= pos ? 1u : 0u;
    }
    for(int d = 16; d >= 1; d >>= 1)n += __shfl_xor_sync(0xffffffffu, n, d);
    __shared__ uint32_t s[kWarps];
    if((threadIdx.x & 31) == 0)s[threadIdx.x >> 5] = n;
    __syncthreads();
    if(threadIdx.x == 0){
        uint32_t t = 0;
        for(int w = 0; w < kWarps; w++)t += s[w];
        chunk_good[blockIdx.x] = t;
    }
}

// one warp per task: prefix over its chunks, children, next-level tasks
__global__ void k_emit(LTask* tasks, LTask* next, SmallTask* small, uint32_t* chunk_good, GNode* nodes, Counters* cnt, int cur, uint32_t leaf_size){
    const uint32_t t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if(t >= cnt->ntasks[cur])return;
    LTask& T = tasks[t];
    uint32_t run = 0;
    for(uint32_t c0 = 0; c0 < T.nchunks; c0 += 32){
        const uint32_t c = c0 + lane;
        const uint32_t v = c < T.nchunks ? chunk_good[T.chunk0 + c] : 0;
        uint32_t inc = v;
        for(int d = 1; d < 32; d <<= 1){ const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d); if(lane >= d)inc += o; }
        if(c < T.nchunks)chunk_good[T.chunk0 + c] = run + inc - v;     // exclusive prefix
        run += __shfl_sync(0xffffffffu, inc, 31);
    }
    if(lane != 0)return;
    const uint32_t span = T.last - T.first;
    const uint32_t mid = T.first + run;
    GNode& N = nodes[T.node];
    if(T.leaf || span <= leaf_size || mid == T.first || mid == T.last){
        T.leaf = 1;
        N.left = 0; N.right = T.first; N.count = span; N.axis = 0;
        return;
    }
    T.good = run;
    const uint32_t l = atomicAdd(&cnt->nnodes, 2u), r = l + 1;
    N.left = l; N.right = r; N.count = 0; N.axis = (uint32_t)T.axis;
    const uint32_t b[3] = { T.first, mid, T.last };
    for(int k = 0; k < 2; k++){
        const uint32_t sp = b[k + 1] - b[k];
        if(sp > (uint32_t)kSmall){
            LTask& C = next[atomicAdd(&cnt->ntasks[cur ^ 1], 1u)];
            C.first = b[k]; C.last = b[k + 1]; C.node = k ? r : l;
        } else{
            SmallTask S; S.first = b[k]; S.last = b[k + 1]; S.node = k ? r : l;
            S.node_base = atomicAdd(&cnt->nnodes, 2u * sp);
            small[atomicAdd(&cnt->nsmall, 1u)] = S;
        }
    }
}

// Partition with the visiting order of a bidirectional std::partition (stl_algo.h __partition): the k-th
// item > pos of [first, mid), counted from the left, trades places with the k-th item <= pos of [mid, last),
// counted from the right; everything else stays where it is.
__global__ void __launch_bounds__(kThreads) k_mark(const float4* __restrict__ ilo, const float4* __restrict__ ihi, const Chunk* __restrict__ chunks, LTask* tasks,
                                                   const uint32_t* __restrict__ chunk_good, uint32_t* bad_pos, uint32_t* good_pos){
    const Chunk ch = chunks[blockIdx.x];
    LTask& T = tasks[ch.task];
    if(T.leaf)return;
    const int axis = T.axis; const float pos = T.pos;
    const uint32_t first = T.first, G = T.good, mid = first + G;
    const uint32_t base = ch.begin + threadIdx.x * kPer;
    bool good[kPer];
    uint32_t mine = 0;
    for(int j = 0; j < kPer; j++){
        const uint32_t i = base + j;
        good[j] = false;
        if(i < ch.end){
            const float4 lo = ilo[i], hi = ihi[i];
            const float c = axis == 0 ? centroid(lo.x, hi.x) : axis == 1 ? centroid(lo.y, hi.y) : centroid(lo.z, hi.z);
            good[j] = c <= pos;
            mine += good[j] ? 1u : 0u;
        }
    }
    // block exclusive scan of `mine`
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t inc = mine;
    for(int d = 1; d < 32; d <<= 1){ const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d); if(lane >= d)inc += o; }
    __shared__ uint32_t s[kWarps];
    __shared__ uint32_t s_m;
    if(threadIdx.x == 0)s_m = 0;
    if(lane == 31)s[w] = inc;
    __syncthreads();
    uint32_t before = chunk_good[blockIdx.x] + inc - mine;
    for(int k = 0; k < w; k++)before += s[k];
    uint32_t nbad = 0;
    for(int j = 0; j < kPer; j++){
        const uint32_t i = base + j;
        if(i >= ch.end)break;
        if(!good[j] && i < mid){ bad_pos[first + (i - first) - before] = i; nbad++; }
        else if(good[j] && i >= mid){ good_pos[first + (G - before - 1)] = i; }
        before += good[j] ? 1u : 0u;
    }
    if(nbad)atomicAdd(&s_m, nbad);
    __syncthreads();
    if(threadIdx.x == 0 && s_m)atomicAdd(&T.m, s_m);
}

__global__ void __launch_bounds__(kThreads) k_swap(float4* ilo, float4* ihi, const Chunk* __restrict__ chunks, const LTask* __restrict__ tasks,
                                                   const uint32_t* __restrict__ bad_pos, const uint32_t* __restrict__ good_pos){
    const Chunk ch = chunks[blockIdx.x];
    const LTask& T = tasks[ch.task];
    if(T.leaf)return;
    const uint32_t m = T.m;
    for(uint32_t j = ch.local * kChunk + threadIdx.x; j < min(m, (ch.local + 1) * kChunk); j += kThreads){
        const uint32_t a = bad_pos[T.first + j], b = good_pos[T.first + j];
        const float4 la = ilo[a], ha = ihi[a], lb = ilo[b], hb = ihi[b];
        ilo[a] = lb; ihi[a] = hb; ilo[b] = la; ihi[b] = ha;
    }
}

// =====================================================================================
// Small phase: one CTA per subtree, items in shared memory
// =====================================================================================

struct SubShared {
    float lo[3][kSmall], hi[3][kSmall];
    uint32_t idx[kSmall];
    uint16_t bad[kSmall / 2], good[kSmall / 2];
    int bins[96][7];                       // CTA bins: lo3 hi3 n (ordered ints)
    int wbins[kWarps][24][7];              // per-warp bins (8 bins x 3 axes)
    float red[kWarps][12];
    uint32_t scan[kWarps];
    SplitEval best[3];
    uint32_t stack[16][3];                 // CTA-mode stack: first, last, node
    uint32_t wlist[kWarpList][3];          // nodes of <= 32 items waiting for a warp
    uint32_t wstack[kWarps][8][3];
    uint32_t sp, nw, next_node, m, cur[3];
    int decision[4];                       // leaf, axis, G ; pos in fpos
    float fpos;
};

__device__ __forceinline__ void write_leaf(GNode* nodes, uint32_t node, const float* v, uint32_t gfirst, uint32_t span){
    GNode N;
    for(int k = 0; k < 3; k++){ N.lo[k] = v[k]; N.hi[k] = v[3 + k]; }
    N.left = 0; N.right = gfirst; N.count = span; N.axis = 0;
    nodes[node] = N;
}
__device__ __forceinline__ void write_inner(GNode* nodes, uint32_t node, const float* v, uint32_t l, uint32_t r, int axis){
    GNode N;
    for(int k = 0; k < 3; k++){ N.lo[k] = v[k]; N.hi[k] = v[3 + k]; }
    N.left = l; N.right = r; N.count = 0; N.axis = (uint32_t)axis;
    nodes[node] = N;
}

// a warp builds the subtree over items [first, last) (<= 32) of the CTA's shared arrays
__device__ void warp_subtree(SubShared& S, GNode* nodes, uint32_t gbase, uint32_t first0, uint32_t last0, uint32_t node0, uint32_t leaf_size){
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const float inf = __int_as_float(0x7f800000);
    uint32_t (*st)[3] = S.wstack[w];
    int sp = 0;
    uint32_t first = first0, last = last0, node = node0;
    while(true){
        const uint32_t span = last - first;
        const bool act = (uint32_t)lane < span;
        const uint32_t p = first + (act ? lane : 0);
        float l[3], h[3], c[3];
        for(int k = 0; k < 3; k++){ l[k] = S.lo[k][p]; h[k] = S.hi[k][p]; c[k] = centroid(l[k], h[k]); }
        float v[12];
        for(int k = 0; k < 3; k++){
            v[k] = act ? l[k] : inf; v[3 + k] = act ? h[k] : -inf;
            v[6 + k] = act ? c[k] : FLT_MAX; v[9 + k] = act ? c[k] : -FLT_MAX;
        }
        for(int d = 16; d >= 1; d >>= 1)
            for(int k = 0; k < 12; k++){
                const float o = __shfl_xor_sync(0xffffffffu, v[k], d);
                v[k] = ((k / 3) & 1) ? fmaxf(v[k], o) : fminf(v[k], o);
            }
        bool leaf = span <= leaf_size;
        int axis = 0; float pos = 0.0f;
        if(!leaf){
            const int nbins = 8;
            int (*wb)[7] = S.wbins[w];
            const int pinf = f2o(inf), ninf = f2o(-inf);
            if(lane < 24){ for(int k = 0; k < 3; k++){ wb[lane][k] = pinf; wb[lane][3 + k] = ninf; } wb[lane][6] = 0; }
            __syncwarp();
            bool valid[3];
            for(int a = 0; a < 3; a++){
                const float mn = v[6 + a], mx = v[9 + a];
                valid[a] = !(fabsf(__fsub_rn(mx, mn)) <= FLT_EPSILON);
                if(valid[a] && act){
                    const float scale = __fdiv_rn(8.0f, __fsub_rn(mx, mn));
                    int* b = wb[a * 8 + bin_of(c[a], mn, scale, nbins)];
                    atomicAdd((uint32_t*)&b[6], 1u);
                    for(int k = 0; k < 3; k++){ atomicMin(&b[k], f2o(l[k])); atomicMax(&b[3 + k], f2o(h[k])); }
                }
            }
            __syncwarp();
            SplitEval e; e.cost = inf; e.key = 0x7fffffff; e.pos = 0.0f;
            {
                const int a = lane >> 3, b = lane & 7;
                if(lane < 24 && b < nbins - 1){
                    // valid[a] with a runtime index: select
                    const bool ok = a == 0 ? valid[0] : a == 1 ? valid[1] : valid[2];
                    const float mn = a == 0 ? v[6] : a == 1 ? v[7] : v[8], mx = a == 0 ? v[9] : a == 1 ? v[10] : v[11];
                    if(ok){
                        auto rd = [&](int i, float lo[3], float hi[3], uint32_t& n){
                            const int* bb = wb[a * 8 + i];
                            for(int k = 0; k < 3; k++){ lo[k] = o2f(bb[k]); hi[k] = o2f(bb[3 + k]); }
                            n = (uint32_t)bb[6];
                        };
                        e = eval_split(rd, b, nbins, span, mn, mx, a * 32 + b);
                    }
                }
            }
            e = warp_best(e);
            __syncwarp();
            const float parent = __fmul_rn(half_area(v, v + 3), __uint2float_rn(span));
            const bool none = e.key == 0x7fffffff;
            axis = none ? 0 : e.key >> 5; pos = none ? 0.0f : e.pos;
            leaf = e.cost >= parent;
        }
        uint32_t mid = first;
        if(!leaf){
            const float cc = axis == 0 ? c[0] : axis == 1 ? c[1] : c[2];
            const uint32_t all = span >= 32 ? 0xffffffffu : ((1u << span) - 1u);
            const uint32_t gm = __ballot_sync(0xffffffffu, act && cc <= pos);
            const uint32_t G = __popc(gm);
            if(G == 0 || G == span)leaf = true;
            else{
                const uint32_t left = (1u << G) - 1u;          // G < span <= 32, so G <= 31
                const uint32_t bad_left = ~gm & left, good_right = gm & all & ~left;
                int src = lane;
                if(act){
                    if((bad_left >> lane) & 1u){ const int k = __popc(bad_left & ((2u << lane) - 1u)); src = (int)__fns(good_right, 31, -k); }
                    else if((good_right >> lane) & 1u){ const int k = __popc(good_right >> lane); src = (int)__fns(bad_left, 0, k); }
                }
                const uint32_t q = first + (act ? src : 0);
                float nl[3], nh[3];
                for(int k = 0; k < 3; k++){ nl[k] = S.lo[k][q]; nh[k] = S.hi[k][q]; }
                const uint32_t ni = S.idx[q];
                __syncwarp();
                if(act){
                    for(int k = 0; k < 3; k++){ S.lo[k][p] = nl[k]; S.hi[k][p] = nh[k]; }
                    S.idx[p] = ni;
                }
                __syncwarp();
                mid = first + G;
            }
        }
        if(leaf){
            if(lane == 0)write_leaf(nodes, node, v, gbase + first, span);
            if(sp == 0)return;
            --sp;
            first = st[sp][0]; last = st[sp][1]; node = st[sp][2];
            __syncwarp();
            continue;
        }
        uint32_t lnode = 0;
        if(lane == 0)lnode = atomicAdd(&S.next_node, 2u);
        lnode = __shfl_sync(0xffffffffu, lnode, 0);
        if(lane == 0)write_inner(nodes, node, v, lnode, lnode + 1, axis);
        // the smaller child next, the larger one on the stack (depth <= log2 32)
        const bool left_small = (mid - first) <= (last - mid);
        if(lane == 0){
            st[sp][0] = left_small ? mid : first; st[sp][1] = left_small ? last : mid; st[sp][2] = left_small ? lnode + 1 : lnode;
        }
        ++sp;
        __syncwarp();
        if(left_small){ last = mid; node = lnode; } else{ first = mid; node = lnode + 1; }
    }
}

__global__ void __launch_bounds__(kThreads) k_subtree(const float4* __restrict__ ilo, const float4* __restrict__ ihi, const SmallTask* __restrict__ tasks,
                                                      GNode* nodes, uint32_t* order, uint32_t leaf_size){
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SubShared& S = *reinterpret_cast<SubShared*>(smem_raw);
    const SmallTask task = tasks[blockIdx.x];
    const uint32_t gbase = task.first, n = task.last - task.first;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const float inf = __int_as_float(0x7f800000);
    for(uint32_t i = tid; i < n; i += kThreads){
        const float4 lo = ilo[gbase + i], hi = ihi[gbase + i];
        S.lo[0][i] = lo.x; S.lo[1][i] = lo.y; S.lo[2][i] = lo.z;
        S.hi[0][i] = hi.x; S.hi[1][i] = hi.y; S.hi[2][i] = hi.z;
        S.idx[i] = __float_as_uint(lo.w);
    }
    if(tid == 0){
        S.sp = 0; S.nw = 0; S.next_node = task.node_base;
        if(n > 32){ S.stack[0][0] = 0; S.stack[0][1] = n; S.stack[0][2] = task.node; S.sp = 1; }
        else{ S.wlist[0][0] = 0; S.wlist[0][1] = n; S.wlist[0][2] = task.node; S.nw = 1; }
    }
    __syncthreads();

    while(true){
        // ---- drain the warp list when it is nearly full or the CTA has nothing else to do ----
        const uint32_t sp = S.sp, nw = S.nw;
        __syncthreads();                                  // everyone has read sp / nw before thread 0 moves them
        if(nw > 0 && (sp == 0 || nw > kWarpList - 2)){
            for(uint32_t k = w; k < nw; k += kWarps)warp_subtree(S, nodes, gbase, S.wlist[k][0], S.wlist[k][1], S.wlist[k][2], leaf_size);
            __syncthreads();
            if(tid == 0)S.nw = 0;
            __syncthreads();
            continue;
        }
        if(sp == 0)break;
        if(tid == 0){ S.cur[0] = S.stack[sp - 1][0]; S.cur[1] = S.stack[sp - 1][1]; S.cur[2] = S.stack[sp - 1][2]; S.sp = sp - 1; }
        __syncthreads();
        const uint32_t first = S.cur[0], last = S.cur[1], node = S.cur[2];
        const uint32_t span = last - first;            // > 32 here

        // ---- bounds ----
        float v[12] = { inf, inf, inf, -inf, -inf, -inf, FLT_MAX, FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX };
        for(uint32_t i = first + tid; i < last; i += kThreads)
            for(int k = 0; k < 3; k++){
                const float l = S.lo[k][i], h = S.hi[k][i], c = centroid(l, h);
                v[k] = fminf(v[k], l); v[3 + k] = fmaxf(v[3 + k], h);
                v[6 + k] = fminf(v[6 + k], c); v[9 + k] = fmaxf(v[9 + k], c);
            }
        for(int d = 16; d >= 1; d >>= 1)
            for(int k = 0; k < 12; k++){
                const float o = __shfl_xor_sync(0xffffffffu, v[k], d);
                v[k] = ((k / 3) & 1) ? fmaxf(v[k], o) : fminf(v[k], o);
            }
        if(lane == 0)for(int k = 0; k < 12; k++)S.red[w][k] = v[k];
        if(tid < 96){
            const int pinf = f2o(inf), ninf = f2o(-inf);
            for(int k = 0; k < 3; k++){ S.bins[tid][k] = pinf; S.bins[tid][3 + k] = ninf; }
            S.bins[tid][6] = 0;
        }
        __syncthreads();
        for(int k = 0; k < 12; k++){
            float r = S.red[0][k];
            for(int ww = 1; ww < kWarps; ww++)r = ((k / 3) & 1) ? fmaxf(r, S.red[ww][k]) : fminf(r, S.red[ww][k]);
            v[k] = r;
        }

        // ---- bins ----
        const int nbins = (int)bins_for(span);
        bool valid[3]; float scale[3];
        for(int a = 0; a < 3; a++){
            valid[a] = !(fabsf(__fsub_rn(v[9 + a], v[6 + a])) <= FLT_EPSILON);
            scale[a] = __fdiv_rn(__uint2float_rn((uint32_t)nbins), __fsub_rn(v[9 + a], v[6 + a]));
        }
        for(uint32_t i = first + tid; i < last; i += kThreads){
            int ol[3], oh[3]; float c[3];
            for(int k = 0; k < 3; k++){ const float l = S.lo[k][i], h = S.hi[k][i]; ol[k] = f2o(l); oh[k] = f2o(h); c[k] = centroid(l, h); }
            for(int a = 0; a < 3; a++){
                if(!valid[a])continue;
                int* b = S.bins[a * 32 + bin_of(c[a], v[6 + a], scale[a], nbins)];
                atomicAdd((uint32_t*)&b[6], 1u);
                for(int k = 0; k < 3; k++){ atomicMin(&b[k], ol[k]); atomicMax(&b[3 + k], oh[k]); }
            }
        }
        __syncthreads();

        // ---- split ----
        if(tid < 96){
            const int a = tid >> 5, b = tid & 31;
            SplitEval e; e.cost = inf; e.key = 0x7fffffff; e.pos = 0.0f;
            const bool ok = a == 0 ? valid[0] : a == 1 ? valid[1] : valid[2];
            const float mn = a == 0 ? v[6] : a == 1 ? v[7] : v[8], mx = a == 0 ? v[9] : a == 1 ? v[10] : v[11];
            if(ok && b < nbins - 1){
                auto rd = [&](int i, float lo[3], float hi[3], uint32_t& nn){
                    const int* bb = S.bins[a * 32 + i];
                    for(int k = 0; k < 3; k++){ lo[k] = o2f(bb[k]); hi[k] = o2f(bb[3 + k]); }
                    nn = (uint32_t)bb[6];
                };
                e = eval_split(rd, b, nbins, span, mn, mx, a * 32 + b);
            }
            e = warp_best(e);
            if(b == 0)S.best[a] = e;
        }
        __syncthreads();
        SplitEval best = S.best[0];
        if(better(S.best[1], best))best = S.best[1];
        if(better(S.best[2], best))best = S.best[2];
        const float parent = __fmul_rn(half_area(v, v + 3), __uint2float_rn(span));
        const bool none = best.key == 0x7fffffff;
        const int axis = none ? 0 : best.key >> 5;
        const float pos = none ? 0.0f : best.pos;
        bool leaf = best.cost >= parent;                 // uniform across the CTA

        // ---- partition ----
        uint32_t mid = first;
        if(!leaf){
            const uint32_t per = (span + kThreads - 1) / kThreads;
            const uint32_t b0 = min(last, first + tid * per), b1 = min(last, b0 + per);
            uint32_t mine = 0;
            for(uint32_t i = b0; i < b1; i++)mine += centroid(S.lo[axis][i], S.hi[axis][i]) <= pos ? 1u : 0u;
            uint32_t inc = mine;
            for(int d = 1; d < 32; d <<= 1){ const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d); if(lane >= d)inc += o; }
            if(lane == 31)S.scan[w] = inc;
            if(tid == 0)S.m = 0;
            __syncthreads();
            uint32_t before = inc - mine, G = 0;
            for(int k = 0; k < kWarps; k++){ if(k < w)before += S.scan[k]; G += S.scan[k]; }
            mid = first + G;
            if(G == 0 || G == span)leaf = true;
            else{
                uint32_t nbad = 0, firstbad = 0;
                for(uint32_t i = b0; i < b1; i++){
                    const bool g = centroid(S.lo[axis][i], S.hi[axis][i]) <= pos;
                    if(!g && i < mid){ const uint32_t k = (i - first) - before; S.bad[k] = (uint16_t)i; if(!nbad)firstbad = k; nbad++; }
                    else if(g && i >= mid)S.good[G - before - 1] = (uint16_t)i;
                    before += g ? 1u : 0u;
                }
                (void)firstbad;
                if(nbad)atomicAdd(&S.m, nbad);
                __syncthreads();
                const uint32_t m = S.m;
                for(uint32_t j = tid; j < m; j += kThreads){
                    const uint32_t a = S.bad[j], b = S.good[j];
                    for(int k = 0; k < 3; k++){
                        const float t0 = S.lo[k][a]; S.lo[k][a] = S.lo[k][b]; S.lo[k][b] = t0;
                        const float t1 = S.hi[k][a]; S.hi[k][a] = S.hi[k][b]; S.hi[k][b] = t1;
                    }
                    const uint32_t t2 = S.idx[a]; S.idx[a] = S.idx[b]; S.idx[b] = t2;
                }
            }
        }
        if(tid == 0){
            if(leaf)write_leaf(nodes, node, v, gbase + first, span);
            else{
                const uint32_t l = atomicAdd(&S.next_node, 2u);
                write_inner(nodes, node, v, l, l + 1, axis);
                const uint32_t b[3] = { first, mid, last };
                // smaller child on top of the stack: it is popped next, so the stack stays <= log2(kSmall / 32) + 1 deep
                const int order2[2] = { (mid - first) <= (last - mid) ? 1 : 0, (mid - first) <= (last - mid) ? 0 : 1 };
                for(int q = 0; q < 2; q++){
                    const int k = order2[q];
                    const uint32_t sp2 = b[k + 1] - b[k];
                    if(sp2 > 32){ uint32_t* e = S.stack[S.sp++]; e[0] = b[k]; e[1] = b[k + 1]; e[2] = l + k; }
                    else{ uint32_t* e = S.wlist[S.nw++]; e[0] = b[k]; e[1] = b[k + 1]; e[2] = l + k; }
                }
            }
        }
        __syncthreads();
    }
    for(uint32_t i = tid; i < n; i += kThreads)order[gbase + i] = S.idx[i];
}

__global__ void k_pack_items(const Box* __restrict__ boxes, uint32_t n, float4* ilo, float4* ihi){
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if(i >= n)return;
    const Box b = boxes[i];
    ilo[i] = make_float4(b.lo[0], b.lo[1], b.lo[2], __uint_as_float(i));
    ihi[i] = make_float4(b.hi[0], b.hi[1], b.hi[2], 0.0f);
}
__global__ void k_order_from_items(const float4* __restrict__ ilo, uint32_t n, uint32_t* order){
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if(i < n)order[i] = __float_as_uint(ilo[i].w);
}

struct DevMem {
    void* p = nullptr;
    ~DevMem(){ if(p)cudaFree(p); }
    bool alloc(size_t bytes, const char* what){ return cuda_ok(cudaMalloc(&p, bytes ? bytes : 16), what); }
    template <class T> T* as(){ return (T*)p; }
};

} // namespace

// Build on the current device.  On success `out` is what build_binary_bvh() returns for the same boxes
// (node numbering aside: nodes are reached through left / right, and unused slots stay zero).
int build_binary_bvh_gpu(const Box* boxes, uint32_t n, uint32_t leaf_size, BinaryBvh& out, GpuBuildTimes* times){
    return build_binary_bvh_gpu_ex(boxes, nullptr, n, leaf_size, &out, nullptr, times);
}

DeviceBvh2::~DeviceBvh2(){ if(nodes)cudaFree(nodes); if(order)cudaFree(order); }

// The general form: the item boxes come from the host (boxes) or are on the device already (d_boxes_in); the tree goes
// back to the host (out) and / or stays on the device for the device flatten (keep: scene_pack_gpu.cu).
int build_binary_bvh_gpu_ex(const Box* boxes, const Box* d_boxes_in, uint32_t n, uint32_t leaf_size, BinaryBvh* out, DeviceBvh2* keep, GpuBuildTimes* times){
    if(out){ out->nodes.clear(); out->order.clear(); }
    if(n == 0)return PT_OK;
    cudaEvent_t ev[4];
    for(auto& e : ev)if(!cuda_ok(cudaEventCreate(&e), "cudaEventCreate"))return PT_ERR_CUDA;
    struct EvGuard { cudaEvent_t* e; ~EvGuard(){ for(int i = 0; i < 4; i++)cudaEventDestroy(e[i]); } } guard { ev };

    const uint32_t max_tasks = n / kSmall + 2;
    const uint32_t max_chunks = n / kChunk + max_tasks + 2;
    const size_t node_cap = 4 * (size_t)n + 16;     // <= 2 per large inner node (< n of them) + 2 * span per subtree
    DevMem d_boxes, d_ilo, d_ihi, d_nodes, d_order, d_tasks[2], d_bins, d_chunks, d_small, d_good, d_bad_pos, d_good_pos, d_cnt;
    if((!d_boxes_in && !d_boxes.alloc(sizeof(Box) * (size_t)n, "cudaMalloc boxes")) || !d_ilo.alloc(16 * (size_t)n, "cudaMalloc items") || !d_ihi.alloc(16 * (size_t)n, "cudaMalloc items") ||
       !d_nodes.alloc(sizeof(GNode) * node_cap, "cudaMalloc nodes") || !d_order.alloc(4 * (size_t)n, "cudaMalloc order") ||
       !d_tasks[0].alloc(sizeof(LTask) * max_tasks, "cudaMalloc tasks") || !d_tasks[1].alloc(sizeof(LTask) * max_tasks, "cudaMalloc tasks") ||
       !d_bins.alloc(sizeof(Bin) * 96 * (size_t)max_tasks, "cudaMalloc bins") || !d_chunks.alloc(sizeof(Chunk) * max_chunks, "cudaMalloc chunks") ||
       !d_small.alloc(sizeof(SmallTask) * ((size_t)n + 2), "cudaMalloc small tasks") || !d_good.alloc(4 * (size_t)max_chunks, "cudaMalloc chunk counts") ||
       !d_bad_pos.alloc(4 * (size_t)n, "cudaMalloc pairs") || !d_good_pos.alloc(4 * (size_t)n, "cudaMalloc pairs") || !d_cnt.alloc(sizeof(Counters), "cudaMalloc counters"))
        return PT_ERR_CUDA;
    Counters* h_cnt = nullptr;
    if(!cuda_ok(cudaMallocHost(&h_cnt, sizeof(Counters)), "cudaMallocHost"))return PT_ERR_CUDA;
    struct HostGuard { Counters* p; ~HostGuard(){ cudaFreeHost(p); } } hguard { h_cnt };

    cudaStream_t st = 0;
    cudaEventRecord(ev[0], st);
    if(!d_boxes_in && !cuda_ok(cudaMemcpyAsync(d_boxes.p, boxes, sizeof(Box) * (size_t)n, cudaMemcpyHostToDevice, st), "H2D boxes"))return PT_ERR_CUDA;
    cudaEventRecord(ev[1], st);
    k_pack_items<<<(n + 255) / 256, 256, 0, st>>>(d_boxes_in ? d_boxes_in : d_boxes.as<Box>(), n, d_ilo.as<float4>(), d_ihi.as<float4>());
    if(!cuda_ok(cudaMemsetAsync(d_nodes.p, 0, sizeof(GNode) * node_cap, st), "node pool clear"))return PT_ERR_CUDA;   // unused slots stay zero

    Counters init; memset(&init, 0, sizeof init);
    init.nnodes = 1;
    LTask root; memset(&root, 0, sizeof root);
    root.first = 0; root.last = n; root.node = 0;
    SmallTask sroot { 0, n, 0, 1 };
    if(n > (uint32_t)kSmall){
        init.ntasks[0] = 1;
        cudaMemcpyAsync(d_tasks[0].p, &root, sizeof root, cudaMemcpyHostToDevice, st);
    } else{
        init.nsmall = 1; init.nnodes = 1 + 2 * n;
        cudaMemcpyAsync(d_small.p, &sroot, sizeof sroot, cudaMemcpyHostToDevice, st);
    }
    cudaMemcpyAsync(d_cnt.p, &init, sizeof init, cudaMemcpyHostToDevice, st);

    int cur = 0, levels = 0;
    uint32_t ntasks = init.ntasks[0], nchunks = 0;
    auto plan = [&]() -> bool{      // chunk list of level `cur` + the two counts the host needs for its grids
        k_plan<<<1, 1024, 0, st>>>(d_tasks[cur].as<LTask>(), d_bins.as<Bin>(), d_chunks.as<Chunk>(), d_cnt.as<Counters>(), cur);
        if(!cuda_ok(cudaMemcpyAsync(h_cnt, d_cnt.p, sizeof(Counters), cudaMemcpyDeviceToHost, st), "D2H counters") ||
           !cuda_ok(cudaStreamSynchronize(st), "bvh build level"))return false;
        ntasks = h_cnt->ntasks[cur]; nchunks = h_cnt->nchunks;
        if(ntasks > max_tasks || nchunks > max_chunks){ set_error("gpu bvh build: task list overflow"); return false; }
        return true;
    };
    if(ntasks > 0 && !plan())return PT_ERR_CUDA;
    while(ntasks > 0){
        k_bounds<<<nchunks, kThreads, 0, st>>>(d_ilo.as<float4>(), d_ihi.as<float4>(), d_chunks.as<Chunk>(), d_tasks[cur].as<LTask>());
        k_bin<<<nchunks, kThreads, 0, st>>>(d_ilo.as<float4>(), d_ihi.as<float4>(), d_chunks.as<Chunk>(), d_tasks[cur].as<LTask>(), d_bins.as<Bin>());
        k_split<<<ntasks, 96, 0, st>>>(d_tasks[cur].as<LTask>(), d_bins.as<Bin>(), d_nodes.as<GNode>());
        k_count<<<nchunks, kThreads, 0, st>>>(d_ilo.as<float4>(), d_ihi.as<float4>(), d_chunks.as<Chunk>(), d_tasks[cur].as<LTask>(), d_good.as<uint32_t>());
        k_emit<<<(ntasks * 32 + 127) / 128, 128, 0, st>>>(d_tasks[cur].as<LTask>(), d_tasks[cur ^ 1].as<LTask>(), d_small.as<SmallTask>(), d_good.as<uint32_t>(),
                                                         d_nodes.as<GNode>(), d_cnt.as<Counters>(), cur, leaf_size);
        k_mark<<<nchunks, kThreads, 0, st>>>(d_ilo.as<float4>(), d_ihi.as<float4>(), d_chunks.as<Chunk>(), d_tasks[cur].as<LTask>(), d_good.as<uint32_t>(),
                                            d_bad_pos.as<uint32_t>(), d_good_pos.as<uint32_t>());
        k_swap<<<nchunks, kThreads, 0, st>>>(d_ilo.as<float4>(), d_ihi.as<float4>(), d_chunks.as<Chunk>(), d_tasks[cur].as<LTask>(), d_bad_pos.as<uint32_t>(), d_good_pos.as<uint32_t>());
        if(!cuda_ok(cudaGetLastError(), "bvh build level launch"))return PT_ERR_CUDA;
        cur ^= 1;
        levels++;
        if(!plan())return PT_ERR_CUDA;
    }
    // items of large-phase leaves are final where they are; subtrees overwrite their own ranges
    k_order_from_items<<<(n + 255) / 256, 256, 0, st>>>(d_ilo.as<float4>(), n, d_order.as<uint32_t>());
    const uint32_t nsmall = n > (uint32_t)kSmall ? h_cnt->nsmall : 1u;
    if(nsmall){
        // function attributes are per device: set on every build (a few microseconds), not once per process
        if(!cuda_ok(cudaFuncSetAttribute(k_subtree, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SubShared)), "smem attribute"))return PT_ERR_CUDA;
        k_subtree<<<nsmall, kThreads, sizeof(SubShared), st>>>(d_ilo.as<float4>(), d_ihi.as<float4>(), d_small.as<SmallTask>(), d_nodes.as<GNode>(), d_order.as<uint32_t>(), leaf_size);
        if(!cuda_ok(cudaGetLastError(), "k_subtree launch"))return PT_ERR_CUDA;
    }
    cudaEventRecord(ev[2], st);
    if(!cuda_ok(cudaMemcpyAsync(h_cnt, d_cnt.p, sizeof(Counters), cudaMemcpyDeviceToHost, st), "D2H counters") ||
       !cuda_ok(cudaStreamSynchronize(st), "bvh subtree build"))return PT_ERR_CUDA;
    const uint32_t nnodes = h_cnt->nnodes;
    if(nnodes > node_cap){ set_error("gpu bvh build: node pool overflow"); return PT_ERR_CUDA; }
    if(out){
        out->nodes.resize(nnodes);
        out->order.resize(n);
        if(!cuda_ok(cudaMemcpyAsync(out->nodes.data(), d_nodes.p, sizeof(GNode) * (size_t)nnodes, cudaMemcpyDeviceToHost, st), "D2H nodes") ||
           !cuda_ok(cudaMemcpyAsync(out->order.data(), d_order.p, 4 * (size_t)n, cudaMemcpyDeviceToHost, st), "D2H order"))return PT_ERR_CUDA;
    }
    cudaEventRecord(ev[3], st);
    if(!cuda_ok(cudaStreamSynchronize(st), "bvh download"))return PT_ERR_CUDA;
    if(times){
        float a = 0, b = 0, c = 0;
        cudaEventElapsedTime(&a, ev[0], ev[1]); cudaEventElapsedTime(&b, ev[1], ev[2]); cudaEventElapsedTime(&c, ev[2], ev[3]);
        times->upload_ms += a; times->build_ms += b; times->download_ms += c; times->levels += levels; times->subtrees += nsmall; times->items += n;
    }
    if(keep){
        if(keep->nodes)cudaFree(keep->nodes);
        if(keep->order)cudaFree(keep->order);
        keep->nodes = d_nodes.p; d_nodes.p = nullptr;
        keep->order = d_order.p; d_order.p = nullptr;
        keep->n_nodes = nnodes; keep->n = n;
    }
    return PT_OK;
}

} // namespace ptb
