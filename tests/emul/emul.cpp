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
#endif

#include "../../pathtracing_b200/csrc/scene_host.hpp"
#include "../../pathtracing_b200/csrc/trace.cuh"

using namespace ptb;

struct EmulScene {
    pt_scene s;
    PackedTraversal P;
    DevScene D;
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
PT_API int emul_scene_add_light(void* p, const pt_light_desc* d){ return scene_add_light(&((EmulScene*)p)->s, d); }

PT_API int emul_scene_commit(void* p, int width){
    EmulScene* e = (EmulScene*)p;
    int rc = pack_traversal(&e->s, width, e->P);
    if(rc != PT_OK)return rc;
    e->D.n2 = e->P.n2.data(); e->D.n4 = e->P.n4.data(); e->D.tris = e->P.tris.data(); e->D.inst = e->P.insts.data();
    e->D.tlas_root = e->P.tlas_root; e->D.width = width;
    e->D.single_root = e->P.single_root; e->D.single_id = e->P.single_id;
    for(int a = 0; a < 3; a++){ e->D.tlo[a] = e->P.bounds.lo[a]; e->D.thi[a] = e->P.bounds.hi[a]; }
    return PT_OK;
}

PT_API int emul_scene_bounds(void* p, float lo[3], float hi[3]){
    EmulScene* e = (EmulScene*)p;
    for(int a = 0; a < 3; a++){ lo[a] = e->P.bounds.lo[a]; hi[a] = e->P.bounds.hi[a]; }
    return PT_OK;
}

PT_API int emul_node_count(void* p){ EmulScene* e = (EmulScene*)p; return (int)(e->D.width == 2 ? e->P.n2.size() : e->P.n4.size()); }

PT_API int emul_trace_closest(void* p, const pt_ray* rays, size_t n, pt_hit* hits, uint32_t* inst, unsigned long long* stats){
    EmulScene* e = (EmulScene*)p;
    const DevScene& D = e->D;
    int nt = (int)std::thread::hardware_concurrency(); if(nt < 1)nt = 1;
    std::vector<TraceCounters> cs(nt, TraceCounters{0, 0, 0, 0});
    auto work = [&](int tid){
        for(size_t i = tid; i < n; i += nt){
            HitRec h;
            const float4* rb = reinterpret_cast<const float4*>(rays);
            if(D.width == 2)trace_scene<2, false, true>(D, rb, i, h, &cs[tid], NoAlpha());
            else trace_scene<4, false, true>(D, rb, i, h, &cs[tid], NoAlpha());
            hits[i].t = h.t; hits[i].u = h.u; hits[i].v = h.v; hits[i].prim = h.prim;
            if(inst)inst[i] = h.inst;
        }
    };
    std::vector<std::thread> pool;
    for(int t = 0; t < nt; t++)pool.emplace_back(work, t);
    for(auto& t : pool)t.join();
    if(stats){
        stats[0] = stats[1] = stats[2] = stats[3] = 0;
        for(auto& c : cs){ stats[0] += c.tlas_nodes; stats[1] += c.blas_nodes; stats[2] += c.instances; stats[3] += c.triangles; }
    }
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
            bool o = D.width == 2 ? trace_scene<2, true, false>(D, rb, i, h, nullptr, NoAlpha())
                                  : trace_scene<4, true, false>(D, rb, i, h, nullptr, NoAlpha());
            occ[i] = o ? 1 : 0;
        }
    };
    std::vector<std::thread> pool;
    for(int t = 0; t < nt; t++)pool.emplace_back(work, t);
    for(auto& t : pool)t.join();
    return PT_OK;
}

} // extern "C"
