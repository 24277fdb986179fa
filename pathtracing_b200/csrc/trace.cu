// trace.cu — ray-cast kernels: batched Scene::Intersect / Scene::IntersectPred
// (Scene.cpp:4-6,31-33) for BVH2 and BVH4, one thread per ray.
//
// Grid: ceil(n / 128) CTAs of 128 threads (4 warps); with <= 64 registers this
// gives 16 CTAs = 64 warps per SM, i.e. full occupancy to hide the dependent
// node-fetch latency that bounds incoherent traversal.  Rays are read as two
// 128-bit loads, hits written as one 128-bit store.
#include <cstdio>
#include "scene_host.hpp"
#include "trace.cuh"

namespace ptb {

constexpr int kTraceBlock = 128;

template <int WIDTH>
__global__ void __launch_bounds__(kTraceBlock) k_trace_closest(DevScene S, const float4* __restrict__ rays, size_t n, float4* __restrict__ hits, uint32_t* __restrict__ inst){
    const size_t i = (size_t)blockIdx.x * kTraceBlock + threadIdx.x;
    if(i >= n)return;
    const float4 a = __ldg(rays + 2 * i), b = __ldg(rays + 2 * i + 1);
    RayCtx r;
    make_ray(r, a.x, a.y, a.z, b.x, b.y, b.z);
    HitRec h;
    trace_scene<WIDTH, false, false>(S, r, a.w, h, nullptr, NoAlpha());
    hits[i] = make_float4(h.t, h.u, h.v, __uint_as_float(h.prim));
    if(inst)inst[i] = h.inst;
}

template <int WIDTH>
__global__ void __launch_bounds__(kTraceBlock) k_trace_any(DevScene S, const float4* __restrict__ rays, size_t n, uint8_t* __restrict__ occ){
    const size_t i = (size_t)blockIdx.x * kTraceBlock + threadIdx.x;
    if(i >= n)return;
    const float4 a = __ldg(rays + 2 * i), b = __ldg(rays + 2 * i + 1);
    RayCtx r;
    make_ray(r, a.x, a.y, a.z, b.x, b.y, b.z);
    HitRec h;
    const bool o = trace_scene<WIDTH, true, false>(S, r, a.w, h, nullptr, NoAlpha());
    occ[i] = o ? 1 : 0;
}

// counting build of the same closest-hit traversal
template <int WIDTH>
__global__ void __launch_bounds__(kTraceBlock) k_trace_stats(DevScene S, const float4* __restrict__ rays, size_t n, unsigned long long* __restrict__ out){
    const size_t i = (size_t)blockIdx.x * kTraceBlock + threadIdx.x;
    TraceCounters c = { 0, 0, 0, 0 };
    unsigned long long hitc = 0;
    if(i < n){
        const float4 a = __ldg(rays + 2 * i), b = __ldg(rays + 2 * i + 1);
        RayCtx r;
        make_ray(r, a.x, a.y, a.z, b.x, b.y, b.z);
        HitRec h;
        if(trace_scene<WIDTH, false, true>(S, r, a.w, h, &c, NoAlpha()))hitc = 1;
    }
    // warp reduce, one atomic per warp per counter
    unsigned long long v[5] = { c.tlas_nodes, c.blas_nodes, c.instances, c.triangles, hitc };
#pragma unroll
    for(int k = 0; k < 5; k++){
        unsigned long long x = v[k];
        for(int o = 16; o > 0; o >>= 1)x += __shfl_down_sync(0xffffffffu, x, o);
        if((threadIdx.x & 31) == 0 && x)atomicAdd(out + k, x);
    }
}

static int check_launch(const char* what){
    cudaError_t e = cudaGetLastError();
    if(e != cudaSuccess){ set_error(std::string(what) + ": " + cudaGetErrorString(e)); return PT_ERR_CUDA; }
    return PT_OK;
}

int launch_trace_closest(const DevScene& S, const void* d_rays, size_t n, void* d_hits, void* d_inst, void* stream){
    if(n == 0)return PT_OK;
    const unsigned grid = (unsigned)((n + kTraceBlock - 1) / kTraceBlock);
    cudaStream_t st = (cudaStream_t)stream;
    if(S.width == 2)k_trace_closest<2><<<grid, kTraceBlock, 0, st>>>(S, (const float4*)d_rays, n, (float4*)d_hits, (uint32_t*)d_inst);
    else k_trace_closest<4><<<grid, kTraceBlock, 0, st>>>(S, (const float4*)d_rays, n, (float4*)d_hits, (uint32_t*)d_inst);
    return check_launch("k_trace_closest");
}

int launch_trace_any(const DevScene& S, const void* d_rays, size_t n, void* d_occ, void* stream){
    if(n == 0)return PT_OK;
    const unsigned grid = (unsigned)((n + kTraceBlock - 1) / kTraceBlock);
    cudaStream_t st = (cudaStream_t)stream;
    if(S.width == 2)k_trace_any<2><<<grid, kTraceBlock, 0, st>>>(S, (const float4*)d_rays, n, (uint8_t*)d_occ);
    else k_trace_any<4><<<grid, kTraceBlock, 0, st>>>(S, (const float4*)d_rays, n, (uint8_t*)d_occ);
    return check_launch("k_trace_any");
}

int launch_trace_stats(const DevScene& S, const void* d_rays, size_t n, void* d_counters, void* stream){
    if(n == 0)return PT_OK;
    const unsigned grid = (unsigned)((n + kTraceBlock - 1) / kTraceBlock);
    cudaStream_t st = (cudaStream_t)stream;
    if(S.width == 2)k_trace_stats<2><<<grid, kTraceBlock, 0, st>>>(S, (const float4*)d_rays, n, (unsigned long long*)d_counters);
    else k_trace_stats<4><<<grid, kTraceBlock, 0, st>>>(S, (const float4*)d_rays, n, (unsigned long long*)d_counters);
    return check_launch("k_trace_stats");
}

} // namespace ptb
