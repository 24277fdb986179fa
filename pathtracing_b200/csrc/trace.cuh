// trace.cuh — device-side BVH2/BVH4 TLAS+BLAS traversal for sm_100a.
//
// One lane owns one ray and drives a small per-ray state machine (TravState +
// one unified TLAS/BLAS stack: shared memory for the first rows, local memory beyond).  One call of
// trav_step() does at most: one pop, one inner-node test, one leaf item (one
// triangle, one instance entry or one instance exit).  Kernels call it from a
// persistent loop and refill finished lanes with new rays (trace.cu, render
// kernels), so a warp never waits for its slowest ray and a leaf costs the
// warp one triangle test per round instead of a whole nested loop
// (profiles/r01_trace4_v1.md: the nested-loop version ran at 6.2 of 32 lanes).
//
// Node records are fetched with 256-bit loads, triangles and instances with
// 128-bit loads (pt_types.h).  The arithmetic of every box and triangle test is written with
// explicit round-to-nearest intrinsics in the operation order of the
// reference, so that no FMA contraction can change which candidate wins
// (SURVEY.md §7 "hard part 1"); pushes and pops happen in the reference's
// order, so exact ties resolve to the same triangle too:
//   box     : AABB::HitDistance (AABB.hpp:72-111) / BVH4 SSE slab test (BVH.hpp:1142-1189)
//   closest : glm::intersectRayTriangle (glm/gtx/intersect.inl:29-94) + Shape.cpp:204
//   any-hit : TriangleShape::IntersectPred (Shape.cpp:246-268)
//   instance: TransformedPrimitive::Intersect/IntersectPred (Primitive.cpp:41-66)
//   BVH2    : BVH2::Intersect / IntersectPred (BVH.hpp:503-548, 476-500)
//   BVH4    : BVH4::Intersect / IntersectPred (BVH.hpp:1111-1211, 1019-1109)
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include "pt_types.h"

namespace ptb {

constexpr float kShadowEps = 0.00001f;          // shadowEpsilon, AABB.hpp:6
constexpr float kFltEps = 1.1920928955078125e-7f;
#ifndef PTB_STACK
#define PTB_STACK 96                            // tests build a second library with a small stack to exercise the overflow error (Makefile)
#endif
constexpr int kStack = PTB_STACK;               // unified TLAS+BLAS stack (the reference: 32 + 32, unchecked)
constexpr uint32_t REF_LEAVE = 0xFFFFFFFEu;     // stack sentinel: end of the current instance

struct RayCtx {
    float ox, oy, oz, dx, dy, dz, ix, iy, iz;
};

struct HitRec {
    float t, u, v;
    uint32_t prim, inst;
    uint32_t overflow;   // the traversal stack overflowed on this ray (TF_OVERFLOW)
};

struct TraceCounters {
    unsigned long long tlas_nodes, blas_nodes, instances, triangles;
};

__device__ __forceinline__ float f_sub(float a, float b){ return __fsub_rn(a, b); }
__device__ __forceinline__ float f_add(float a, float b){ return __fadd_rn(a, b); }
__device__ __forceinline__ float f_mul(float a, float b){ return __fmul_rn(a, b); }
__device__ __forceinline__ float f_div(float a, float b){ return __fdiv_rn(a, b); }

// Ray ctor, Ray.hpp:32-37
#if defined(__CUDA_ARCH__)
__device__ __forceinline__ float f_rcp(float a){ return __frcp_rn(a); }   // correctly rounded: == 1.0f / a
__device__ __forceinline__ float f_sqrt(float a){ return __fsqrt_rn(a); }
#else
__device__ __forceinline__ float f_sqrt(float a){ return sqrtf(a); }
__device__ __forceinline__ float f_rcp(float a){ return 1.0f / a; }
#endif
__device__ __forceinline__ float inv_dir(float d){ return fabsf(d) < 1e-32f ? 1e32f : f_rcp(d); }

__device__ __forceinline__ void make_ray(RayCtx& r, float ox, float oy, float oz, float dx, float dy, float dz){
    r.ox = ox; r.oy = oy; r.oz = oz; r.dx = dx; r.dy = dy; r.dz = dz;
    r.ix = inv_dir(dx); r.iy = inv_dir(dy); r.iz = inv_dir(dz);
}

// slab test; returns entry/exit
__device__ __forceinline__ void slab(const RayCtx& r, float lx, float ly, float lz, float hx, float hy, float hz, float& tEntry, float& tExit){
    const float ax = f_mul(f_sub(lx, r.ox), r.ix), bx = f_mul(f_sub(hx, r.ox), r.ix);
    const float ay = f_mul(f_sub(ly, r.oy), r.iy), by = f_mul(f_sub(hy, r.oy), r.iy);
    const float az = f_mul(f_sub(lz, r.oz), r.iz), bz = f_mul(f_sub(hz, r.oz), r.iz);
    tEntry = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
    tExit = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
}

// AABB::HitDistance(ray, max) != inf   (AABB.hpp:110)
__device__ __forceinline__ bool box_hit2(const RayCtx& r, const float* lo, const float* hi, float tmax){
    float a, b;
    slab(r, lo[0], lo[1], lo[2], hi[0], hi[1], hi[2], a, b);
    return a <= b && b >= kShadowEps && a <= tmax;
}

struct TriData { float4 a, b, c; };

// 256-bit read-only global load (LDG.E.256, sm_100+): a 128 B cluster arrives in
// 4 L1 wavefronts per lane instead of 8 (profiles/r01_trace4_v2.md: the kernel is
// bound by l1tex data-pipe wavefronts, not by bytes).  p must be 32 B aligned.
struct F8 { float4 lo, hi; };
__device__ __forceinline__ F8 ldg256(const void* p){
    F8 r;
#if defined(__CUDA_ARCH__)
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(r.lo.x), "=f"(r.lo.y), "=f"(r.lo.z), "=f"(r.lo.w), "=f"(r.hi.x), "=f"(r.hi.y), "=f"(r.hi.z), "=f"(r.hi.w) : "l"(p));
#else
    r.lo = reinterpret_cast<const float4*>(p)[0]; r.hi = reinterpret_cast<const float4*>(p)[1];
#endif
    return r;
}

__device__ __forceinline__ TriData load_tri64(const Tri64* tris, uint32_t i){
    const float4* p = reinterpret_cast<const float4*>(tris + i);
    const F8 w0 = ldg256(p), w1 = ldg256(p + 2);
    TriData t;
    t.a = w0.lo; t.b = w0.hi; t.c = w1.lo;
    return t;
}

__device__ __forceinline__ TriData load_tri(const Tri* tris, uint32_t i){
    const float4* p = reinterpret_cast<const float4*>(tris + i);
    TriData t;
    t.a = __ldg(p); t.b = __ldg(p + 1); t.c = __ldg(p + 2);
    return t;
}

// glm::intersectRayTriangle + the t-range reject of TriangleShape::Intersect.
// Accepts t == tmax (the reference rejects only t > max, Shape.cpp:204).
__device__ __forceinline__ bool tri_closest(const RayCtx& r, const TriData& T, float tmax, float& t, float& u, float& v){
    const float e1x = T.b.x, e1y = T.b.y, e1z = T.b.z;
    const float e2x = T.c.x, e2y = T.c.y, e2z = T.c.z;
    // p = cross(dir, edge2)
    const float px = f_sub(f_mul(r.dy, e2z), f_mul(e2y, r.dz));
    const float py = f_sub(f_mul(r.dz, e2x), f_mul(e2z, r.dx));
    const float pz = f_sub(f_mul(r.dx, e2y), f_mul(e2x, r.dy));
    const float det = f_add(f_add(f_mul(e1x, px), f_mul(e1y, py)), f_mul(e1z, pz));
    const float sx = f_sub(r.ox, T.a.x), sy = f_sub(r.oy, T.a.y), sz = f_sub(r.oz, T.a.z);
    const float bx = f_add(f_add(f_mul(sx, px), f_mul(sy, py)), f_mul(sz, pz));
    // Perpendicular = cross(dist, edge1)
    const float qx = f_sub(f_mul(sy, e1z), f_mul(e1y, sz));
    const float qy = f_sub(f_mul(sz, e1x), f_mul(e1z, sx));
    const float qz = f_sub(f_mul(sx, e1y), f_mul(e1x, sy));
    const float by = f_add(f_add(f_mul(r.dx, qx), f_mul(r.dy, qy)), f_mul(r.dz, qz));
    if(det > 0.0f){
        if(bx < 0.0f || bx > det)return false;
        if(by < 0.0f || f_add(bx, by) > det)return false;
    } else if(det < 0.0f){
        if(bx > 0.0f || bx < det)return false;
        if(by > 0.0f || f_add(bx, by) < det)return false;
    } else return false;
    const float inv_det = f_div(1.0f, det);
    const float tt = f_mul(f_add(f_add(f_mul(e2x, qx), f_mul(e2y, qy)), f_mul(e2z, qz)), inv_det);
    if(tt > tmax || tt < kShadowEps)return false;
    t = tt; u = f_mul(bx, inv_det); v = f_mul(by, inv_det);
    return true;
}

// TriangleShape::IntersectPred
__device__ __forceinline__ bool tri_any(const RayCtx& r, const TriData& T, float tmax){
    const float e1x = T.b.x, e1y = T.b.y, e1z = T.b.z;
    const float e2x = T.c.x, e2y = T.c.y, e2z = T.c.z;
    const float hx = f_sub(f_mul(r.dy, e2z), f_mul(e2y, r.dz));
    const float hy = f_sub(f_mul(r.dz, e2x), f_mul(e2z, r.dx));
    const float hz = f_sub(f_mul(r.dx, e2y), f_mul(e2x, r.dy));
    const float det = f_add(f_add(f_mul(e1x, hx), f_mul(e1y, hy)), f_mul(e1z, hz));
    if(det > -kFltEps && det < kFltEps)return false;
    const float inv_det = f_div(1.0f, det);
    const float sx = f_sub(r.ox, T.a.x), sy = f_sub(r.oy, T.a.y), sz = f_sub(r.oz, T.a.z);
    const float u = f_mul(f_add(f_add(f_mul(sx, hx), f_mul(sy, hy)), f_mul(sz, hz)), inv_det);
    if(u < 0.0f || u > 1.0f)return false;
    const float qx = f_sub(f_mul(sy, e1z), f_mul(e1y, sz));
    const float qy = f_sub(f_mul(sz, e1x), f_mul(e1z, sx));
    const float qz = f_sub(f_mul(sx, e1y), f_mul(e1x, sy));
    const float v = f_mul(f_add(f_add(f_mul(r.dx, qx), f_mul(r.dy, qy)), f_mul(r.dz, qz)), inv_det);
    if(v < 0.0f || f_add(u, v) > 1.0f)return false;
    const float t = f_mul(f_add(f_add(f_mul(e2x, qx), f_mul(e2y, qy)), f_mul(e2z, qz)), inv_det);
    return t <= tmax && t >= kShadowEps;
}

// Hook for alpha-tested materials (GeometricPrimitive::Intersect/IntersectPred,
// Primitive.cpp:6-26).  The ray-cast entry points pass NoAlpha; the integrator
// passes a functor that evaluates the material's alpha texture.
// operator()(mesh, prim, inst_record, u, v, ray index) -> accept the hit?
// kWrap: does this kernel variant understand analytic shapes and wrapper chains (sphere / quad GeometricPrimitives,
// AnimatedPrimitive, TransformedPrimitive over a shape; pt_types.h)?  The plain variant -- TLAS of BLAS instances,
// the C2-C5 case -- does not carry that code: with the wrappers compiled in it needs 78 registers instead of 64
// (6 instead of 8 CTAs per SM), and the shape branch alone costs the BVH2 kernel 7 % (63 registers instead of 55,
// 940 instead of 1008 Mrays/s on the C2 soup, scratch/bvh2_bisect.py).  A scene that holds a shape or a wrapper
// runs the general (AlphaTex) variant.
// kSingle: the scene is ONE untransformed BLAS (DevScene::single_root: traversal starts inside it and never meets a TLAS leaf or an
// instance exit), so that code is not compiled in -- the C2 case; the instance code costs the plain kernel registers and schedule
// even when no lane ever runs it (1631 -> 1573 Mrays/s when the world-box test joined it, profiles/r02_trace.md).
struct NoAlpha {
    static constexpr bool kWrap = false;
    static constexpr bool kSingle = false;
    __device__ __forceinline__ bool operator()(uint32_t, uint32_t, uint32_t, float, float, size_t) const{ return true; }
    __device__ __forceinline__ bool shape(uint32_t, bool, float, float, float, uint32_t, size_t) const{ return true; }
};

struct NoAlphaWrap : NoAlpha { static constexpr bool kWrap = true; };   // whole-ray walks: host emulation, tiny scenes
struct NoAlphaSingle : NoAlpha { static constexpr bool kSingle = true; };

struct InstData { float4 m0, m1, m2; uint4 meta; float4 b0; float4 b1; };

__device__ __forceinline__ InstData load_inst(const Instance* inst, uint32_t i){
    const float4* p = reinterpret_cast<const float4*>(inst + i);
    InstData d;
    d.m0 = __ldg(p); d.m1 = __ldg(p + 1); d.m2 = __ldg(p + 2);
    d.meta = __ldg(reinterpret_cast<const uint4*>(p + 3));
    d.b0 = __ldg(p + 4); d.b1 = __ldg(p + 5);
    return d;
}

// Material::Alpha on a candidate hit of a sphere / quad record flagged INST_ALPHA
template <class Alpha>
__device__ __forceinline__ bool shape_alpha(const Alpha& alpha, const RayCtx& r, const InstData& c, float t, float u, float v, uint32_t inst_rec, size_t ray_idx){
    if(c.meta.y & INST_SPHERE){
        const float ax = f_sub(f_add(r.ox, f_mul(r.dx, t)), c.m0.x), ay = f_sub(f_add(r.oy, f_mul(r.dy, t)), c.m0.y), az = f_sub(f_add(r.oz, f_mul(r.dz, t)), c.m0.z);
        return alpha.shape(c.meta.x, true, ax, ay, az, inst_rec, ray_idx);
    }
    return alpha.shape(c.meta.x, false, u, v, 0.0f, inst_rec, ray_idx);
}

// SphereShape::Intersect / IntersectPred (Shape.cpp:3-56): both roots, strict `eps < t < max`.
// Record layout: scene_build.cpp pack_shape.
__device__ __forceinline__ bool sphere_hit(const RayCtx& r, const InstData& d, float tmax, float& t){
    const float ocx = f_sub(r.ox, d.m0.x), ocy = f_sub(r.oy, d.m0.y), ocz = f_sub(r.oz, d.m0.z);
    const float a = f_add(f_add(f_mul(r.dx, r.dx), f_mul(r.dy, r.dy)), f_mul(r.dz, r.dz));
    const float b = f_add(f_add(f_mul(ocx, r.dx), f_mul(ocy, r.dy)), f_mul(ocz, r.dz));
    const float c = f_sub(f_add(f_add(f_mul(ocx, ocx), f_mul(ocy, ocy)), f_mul(ocz, ocz)), f_mul(d.m0.w, d.m0.w));
    const float disc = f_sub(f_mul(b, b), f_mul(a, c));
    if(!(disc > 0.0f))return false;
    const float sq = f_sqrt(disc);
    t = f_div(f_sub(-b, sq), a);
    if(t < tmax && t > kShadowEps)return true;
    t = f_div(f_add(-b, sq), a);
    return t < tmax && t > kShadowEps;
}

// QuadShape::Intersect / IntersectPred (Shape.cpp:316-359): plane hit with the normal turned against the
// ray, `eps <= t <= max`, then the (alpha, beta) interior test.
__device__ __forceinline__ bool quad_hit(const RayCtx& r, const InstData& d, float tmax, float& t, float& alpha, float& beta){
    const float Qx = d.m0.x, Qy = d.m0.y, Qz = d.m0.z, ux = d.m0.w, uy = d.m1.x, uz = d.m1.y, vx = d.m1.z, vy = d.m1.w, vz = d.m2.x;
    float nx = d.m2.y, ny = d.m2.z, nz = d.m2.w, D = d.b0.w;
    const float wx = d.b0.x, wy = d.b0.y, wz = d.b0.z;
    if(f_add(f_add(f_mul(r.dx, nx), f_mul(r.dy, ny)), f_mul(r.dz, nz)) > 0.0f){ nx = -nx; ny = -ny; nz = -nz; D = -D; }
    const float denom = f_add(f_add(f_mul(nx, r.dx), f_mul(ny, r.dy)), f_mul(nz, r.dz));
    if(fabsf(denom) < 1e-8f)return false;
    t = f_div(f_sub(D, f_add(f_add(f_mul(nx, r.ox), f_mul(ny, r.oy)), f_mul(nz, r.oz))), denom);
    if(t < kShadowEps || t > tmax)return false;
    const float hx = f_sub(f_add(r.ox, f_mul(t, r.dx)), Qx), hy = f_sub(f_add(r.oy, f_mul(t, r.dy)), Qy), hz = f_sub(f_add(r.oz, f_mul(t, r.dz)), Qz);
    // alpha = dot(w, cross(planar_hit, v)), beta = dot(w, cross(u, planar_hit))
    const float ax = f_sub(f_mul(hy, vz), f_mul(vy, hz)), ay = f_sub(f_mul(hz, vx), f_mul(vz, hx)), az = f_sub(f_mul(hx, vy), f_mul(vx, hy));
    const float bx = f_sub(f_mul(uy, hz), f_mul(hy, uz)), by = f_sub(f_mul(uz, hx), f_mul(hz, ux)), bz = f_sub(f_mul(ux, hy), f_mul(hx, uy));
    alpha = f_add(f_add(f_mul(wx, ax), f_mul(wy, ay)), f_mul(wz, az));
    beta = f_add(f_add(f_mul(wx, bx), f_mul(wy, by)), f_mul(wz, bz));
    return alpha >= 0.0f && alpha <= 1.0f && beta >= 0.0f && beta <= 1.0f;
}

// glm mat4 * vec4 restricted to xyz: (m0*x + m1*y) + (m2*z + m3*w)
__device__ __forceinline__ void xf_apply(const InstData& d, float x, float y, float z, float w, float& ox, float& oy, float& oz){
    // inv[c*3+r]: m0 = (c0r0,c0r1,c0r2,c1r0) m1 = (c1r1,c1r2,c2r0,c2r1) m2 = (c2r2,c3r0,c3r1,c3r2)
    ox = f_add(f_add(f_mul(d.m0.x, x), f_mul(d.m0.w, y)), f_add(f_mul(d.m1.z, z), f_mul(d.m2.y, w)));
    oy = f_add(f_add(f_mul(d.m0.y, x), f_mul(d.m1.x, y)), f_add(f_mul(d.m1.w, z), f_mul(d.m2.z, w)));
    oz = f_add(f_add(f_mul(d.m0.z, x), f_mul(d.m1.y, y)), f_add(f_mul(d.m2.x, z), f_mul(d.m2.w, w)));
}

// TransformedPrimitive::Intersect[Pred] head (Primitive.cpp:41-52): ray into the primitive's space,
// direction renormalised, max scaled by the length.
__device__ __forceinline__ void xf_ray(const InstData& d, RayCtx& r, float& tmax, float& len){
    float ddx, ddy, ddz, oox, ooy, ooz;
    xf_apply(d, r.dx, r.dy, r.dz, 0.0f, ddx, ddy, ddz);
    len = f_sqrt(f_add(f_add(f_mul(ddx, ddx), f_mul(ddy, ddy)), f_mul(ddz, ddz)));
    xf_apply(d, r.ox, r.oy, r.oz, 1.0f, oox, ooy, ooz);
    make_ray(r, oox, ooy, ooz, f_div(ddx, len), f_div(ddy, len), f_div(ddz, len));
    tmax = f_mul(tmax, len);
}

// AnimatedPrimitive::Intersect[Pred] (Primitive.cpp:82-89): a TransformedPrimitive with
// glm::translate(mat4(1), dir * t), t = clamp(time - tb.x, tb.x, tb.y) / (tb.y - tb.x).  glm::inverse of
// that matrix is the identity with column 3 = -dir*t (tests/test_raycast_emul.py checks it bit for bit);
// it goes through the same mat4 * vec4 arithmetic as every other transform.
__device__ __forceinline__ void anim_ray(const InstData& a, float time, RayCtx& r, float& tmax, float& len){
    const float lo = a.m0.w, hi = a.m1.x;
    const float tt = f_div(fminf(fmaxf(f_sub(time, lo), lo), hi), f_sub(hi, lo));
    InstData T;
    T.m0 = make_float4(1.0f, 0.0f, 0.0f, 0.0f);
    T.m1 = make_float4(1.0f, 0.0f, 0.0f, 0.0f);
    T.m2 = make_float4(1.0f, -f_mul(a.m0.x, tt), -f_mul(a.m0.y, tt), -f_mul(a.m0.z, tt));
    xf_ray(T, r, tmax, len);
}

// ---- per-ray traversal state ------------------------------------------------------------

enum : uint32_t {
    TF_IN_INST = 1u,     // below a TLAS leaf: leaf refs now mean triangles
    TF_XF = 2u,          // the current instance is a TransformedPrimitive: r/tmax are object space
    TF_FOUND_LOCAL = 4u, // a hit was accepted inside the current instance
    TF_DONE = 8u,
    TF_OCCLUDED = 16u,   // any-hit result
    TF_ANIM2 = 32u,      // AnimatedPrimitive over a TransformedPrimitive: the outer |dir| sits on the stack under REF_LEAVE
    TF_OVERFLOW = 64u    // a push found the traversal stack full and was dropped: the result may be wrong, the caller counts it and fails the call
};

struct TravState {
    RayCtx r;            // ray in the current space (world, or object space of the instance)
    float tmax;          // closest accepted t so far, current space
    float wtmax;         // world-space tmax saved while inside a transformed instance
    float len;           // |M^-1 dir| of the current transformed instance
    uint32_t cur;        // what to process next: inner ref | leaf ref | REF_LEAVE | REF_EMPTY (= pop)
    int sp;
    uint32_t flags;
    uint32_t inst_rec;   // record index (leaf order) of the current instance
    uint32_t inst_id;    // its id (pt_scene_add_instance order)
    uint32_t octant;     // BVH4 visit-order selector: sign bits of r.d
    float hu, hv;
    uint32_t hprim, hinst;
};

__device__ __forceinline__ uint32_t ray_octant(const RayCtx& r){
    return (r.dx < 0.0f ? 1u : 0u) | (r.dy < 0.0f ? 2u : 0u) | (r.dz < 0.0f ? 4u : 0u);
}

// Traversal stack: the first KS entries of every lane live in shared memory (row = level, column =
// thread: a warp's 64-bit accesses are conflict-free whatever level each lane is at), deeper entries
// spill to the lane's local-memory window.  profiles/r01_trace4_v4.md: with the whole stack in local
// memory 58 % of the pops missed L1 and the stack lines evicted the upper tree levels.
template <int KS, int STRIDE>
struct TravStack {
    uint2* sm;      // &smem[threadIdx.x]
    uint2* lm;      // kStack - KS local entries
    float* wr;      // this lane's column of the nine world-ray rows behind the stack rows (stack_save_world_ray)
    __device__ __forceinline__ void st(int i, uint2 v) const{ if(i < KS)sm[i * STRIDE] = v; else lm[i - KS] = v; }
    __device__ __forceinline__ uint2 ld(int i) const{ return i < KS ? sm[i * STRIDE] : lm[i - KS]; }
};
template <>
struct TravStack<0, 1> {    // all local (host emulation, tiny launches)
    uint2* lm;
    __device__ __forceinline__ void st(int i, uint2 v) const{ lm[i] = v; }
    __device__ __forceinline__ uint2 ld(int i) const{ return lm[i]; }
};
using LocalStack = TravStack<0, 1>;

template <class STK>
__device__ __forceinline__ void trav_push(const STK& stack, int& sp, uint32_t& flags, uint32_t ref, float entry){
    if(sp < kStack){ stack.st(sp, make_uint2(ref, __float_as_uint(entry))); ++sp; }
    else flags |= TF_OVERFLOW;
}

// Scene::Intersect / IntersectPred entry (Scene.cpp:4-6,31-33).
template <int WIDTH>
__device__ __forceinline__ void trav_init(TravState& s, const DevScene& S, const float4 a, const float4 b){
    make_ray(s.r, a.x, a.y, a.z, b.x, b.y, b.z);
    s.tmax = a.w; s.wtmax = a.w; s.len = 1.0f;
    s.sp = 0; s.flags = 0; s.inst_rec = 0; s.inst_id = 0xFFFFFFFFu;
    s.hu = 0.0f; s.hv = 0.0f; s.hprim = 0xFFFFFFFFu; s.hinst = 0xFFFFFFFFu;
    s.octant = ray_octant(s.r);
    s.cur = S.tlas_root;
    if(S.single_root != REF_EMPTY){
        // one untransformed instance: TLAS root box == BLAS root box, the leaf holds just this BLAS
        s.cur = S.single_root; s.flags = TF_IN_INST; s.inst_id = S.single_id;
    }
    if(WIDTH == 2){
        // BVH2::Intersect tests the root box first (BVH.hpp:504,477)
        if(!box_hit2(s.r, S.tlo, S.thi, s.tmax)){ s.cur = REF_EMPTY; s.flags = TF_DONE; }
    }
}

// ---- inner nodes ----------------------------------------------------------------------

// BVH2 child-pair node.  Closest: push order by the sign of dir[node.axis]
// (BVH.hpp:520-545); any-hit: child 0 then child 1 (BVH.hpp:489-490).
template <bool ANY, class STK>
__device__ __forceinline__ void node_step2(TravState& s, const DevScene& S, const STK& stack){
    const float4* p = reinterpret_cast<const float4*>(S.n2 + s.cur);
    const F8 w0 = ldg256(p), w1 = ldg256(p + 2);
    const float4 q0 = w0.lo, q1 = w0.hi, q2 = w1.lo;
    const uint4 q3 = make_uint4(__float_as_uint(w1.hi.x), __float_as_uint(w1.hi.y), __float_as_uint(w1.hi.z), __float_as_uint(w1.hi.w));
    float a0, b0, a1, b1;
    slab(s.r, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, a0, b0);
    slab(s.r, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, a1, b1);
    const bool h0 = a0 <= b0 && b0 >= kShadowEps && a0 <= s.tmax;
    const bool h1 = a1 <= b1 && b1 >= kShadowEps && a1 <= s.tmax;
    bool near0 = false;   // any-hit: child 1 is pushed last, so it is visited first
    if(!ANY){
        const float da = q3.z == 0 ? s.r.dx : (q3.z == 1 ? s.r.dy : s.r.dz);
        near0 = da > 0.0f;
    }
    const uint32_t nref = near0 ? q3.x : q3.y, fref = near0 ? q3.y : q3.x;
    const bool hn = near0 ? h0 : h1, hf = near0 ? h1 : h0;
    if(hn){
        if(hf)trav_push(stack, s.sp, s.flags, fref, 0.0f);
        s.cur = nref;
    } else s.cur = hf ? fref : REF_EMPTY;
}

// Up to three pushes of one cluster visit.  Common case (all of them land in the shared-memory rows): predicated
// stores, no branch per push; the generic path (rows spilling to local memory, overflow) only near the row limit.
// profiles/r02_trace44_v1.md: with one smem-or-local branch inside each of the four unrolled pushes the push code was
// 12 % of all instructions of the kernel.
template <int KS, int STRIDE>
__device__ __forceinline__ bool stack_fast(const TravStack<KS, STRIDE>&, int sp){ return KS >= 4 && sp + 3 <= KS; }
__device__ __forceinline__ bool stack_fast(const LocalStack&, int){ return false; }
template <int KS, int STRIDE>
__device__ __forceinline__ void stack_st_sm(const TravStack<KS, STRIDE>& k, int i, uint2 v){ k.sm[i * STRIDE] = v; }
__device__ __forceinline__ void stack_st_sm(const LocalStack& k, int i, uint2 v){ k.lm[i] = v; }
// the same store under a predicate, as ONE predicated instruction: written as `if(p)store`, the compiler turns each of a
// cluster's three pushes into a divergent branch with its own convergence barrier (7 instructions around one STS)
template <int KS, int STRIDE>
__device__ __forceinline__ void stack_st_sm_if(const TravStack<KS, STRIDE>& k, int i, uint2 v, bool p){
#if defined(__CUDA_ARCH__)
    const unsigned addr = (unsigned)__cvta_generic_to_shared(k.sm + i * STRIDE);
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.u32 q, %3, 0;\n\t@q st.shared.v2.u32 [%0], {%1, %2};\n\t}" :: "r"(addr), "r"(v.x), "r"(v.y), "r"((unsigned)p) : "memory");
#else
    if(p)k.sm[i * STRIDE] = v;
#endif
}
__device__ __forceinline__ void stack_st_sm_if(const LocalStack& k, int i, uint2 v, bool p){ if(p)k.lm[i] = v; }

// The world-space ray of a lane that is inside a transformed instance.  Leaving the instance used to re-read the ray from global
// memory and rebuild its three exact reciprocals (about 40 instructions, run for 4 of 32 lanes in most rounds of an instanced scene:
// 13 % of the instructions of a C3 frame's closest-hit kernel, profiles/r02_trace.md); nine shared-memory words per lane hold it instead.
// The restored values are the ones make_ray would compute again, bit for bit.  Stacks without shared memory keep the re-read.
constexpr int kWorldRayRows = 9;
template <int KS, int STRIDE>
__device__ __forceinline__ void stack_save_world_ray(const TravStack<KS, STRIDE>& k, const RayCtx& r){
    k.wr[0] = r.ox; k.wr[STRIDE] = r.oy; k.wr[2 * STRIDE] = r.oz; k.wr[3 * STRIDE] = r.dx; k.wr[4 * STRIDE] = r.dy; k.wr[5 * STRIDE] = r.dz;
    k.wr[6 * STRIDE] = r.ix; k.wr[7 * STRIDE] = r.iy; k.wr[8 * STRIDE] = r.iz;
}
template <int KS, int STRIDE>
__device__ __forceinline__ bool stack_load_world_ray(const TravStack<KS, STRIDE>& k, RayCtx& r){
    r.ox = k.wr[0]; r.oy = k.wr[STRIDE]; r.oz = k.wr[2 * STRIDE]; r.dx = k.wr[3 * STRIDE]; r.dy = k.wr[4 * STRIDE]; r.dz = k.wr[5 * STRIDE];
    r.ix = k.wr[6 * STRIDE]; r.iy = k.wr[7 * STRIDE]; r.iz = k.wr[8 * STRIDE];
    return true;
}
__device__ __forceinline__ void stack_save_world_ray(const LocalStack&, const RayCtx&){}
__device__ __forceinline__ bool stack_load_world_ray(const LocalStack&, RayCtx&){ return false; }

// pop: rows below KS come straight from shared memory (no branch in the common case of a shallow stack)
template <int KS, int STRIDE>
__device__ __forceinline__ uint2 stack_ld(const TravStack<KS, STRIDE>& k, int i){ return k.ld(i); }
__device__ __forceinline__ uint2 stack_ld(const LocalStack& k, int i){ return k.lm[i]; }

__device__ __forceinline__ uint32_t pick4(const uint4& v, unsigned k){
    return k == 0 ? v.x : (k == 1 ? v.y : (k == 2 ? v.z : v.w));
}

// BVH4 cluster: 4-wide slab test (BVH.hpp:1142-1189; hit iff tExit >= eps &&
// tEntry < max && tEntry <= tExit, note the strict '<'), then the children
// that were hit are pushed far -> near in the cluster's fixed visit order for
// this ray octant (the folded LUT[signs][perm] / maskLUT of BVH.hpp:1193-1204).
// The nearest one stays in s.cur instead of going through the stack: it was
// just tested against the current tmax, so the pop-time cull cannot reject it.
template <bool ANY, class STK>
__device__ __forceinline__ void node_step4(TravState& s, const DevScene& S, const STK& stack){
    const float4* p = reinterpret_cast<const float4*>(S.n4 + s.cur);
    const F8 wx = ldg256(p), wy = ldg256(p + 2), wz = ldg256(p + 4), wr = ldg256(p + 6);
    const float4 xmin = wx.lo, xmax = wx.hi, ymin = wy.lo, ymax = wy.hi, zmin = wz.lo, zmax = wz.hi;
    const uint4 refs = make_uint4(__float_as_uint(wr.lo.x), __float_as_uint(wr.lo.y), __float_as_uint(wr.lo.z), __float_as_uint(wr.lo.w));
    const uint2 ord = make_uint2(__float_as_uint(wr.hi.x), __float_as_uint(wr.hi.y));
    float e0, e1, e2, e3, x;
    unsigned mask = 0;
    const float tmax = s.tmax;
    slab(s.r, xmin.x, ymin.x, zmin.x, xmax.x, ymax.x, zmax.x, e0, x); if(x >= kShadowEps && e0 < tmax && e0 <= x)mask |= 1u;
    slab(s.r, xmin.y, ymin.y, zmin.y, xmax.y, ymax.y, zmax.y, e1, x); if(x >= kShadowEps && e1 < tmax && e1 <= x)mask |= 2u;
    slab(s.r, xmin.z, ymin.z, zmin.z, xmax.z, ymax.z, zmax.z, e2, x); if(x >= kShadowEps && e2 < tmax && e2 <= x)mask |= 4u;
    slab(s.r, xmin.w, ymin.w, zmin.w, xmax.w, ymax.w, zmax.w, e3, x); if(x >= kShadowEps && e3 < tmax && e3 <= x)mask |= 8u;
    // branch-free: the selects stay SEL/predicated stores, the warp does not split per child
    uint32_t cur = REF_EMPTY;
    float cure = 0.0f;
    int sp = s.sp;
    bool ovf = false;
    const unsigned word = s.octant < 4 ? ord.x : ord.y;
    const unsigned code = (word >> ((s.octant & 3u) * 8u)) & 0xFFu;
    auto child = [&](int k, uint32_t& r, float& e, bool& hit){
        if(ANY){
            // any-hit: the answer does not depend on the visit order; BVH4::IntersectPred pushes unordered too (BVH.hpp:1019-1109)
            hit = ((mask >> k) & 1u) != 0;
            r = k == 0 ? refs.x : k == 1 ? refs.y : k == 2 ? refs.z : refs.w;
            e = k == 0 ? e0 : k == 1 ? e1 : k == 2 ? e2 : e3;
        } else{
            const unsigned lane = (code >> (2 * k)) & 3u;   // k = 0 is the LAST visited lane
            hit = ((mask >> lane) & 1u) != 0;
            const bool hi = (lane & 2u) != 0, lo = (lane & 1u) != 0;
            const uint32_t ra = lo ? refs.y : refs.x, rb = lo ? refs.w : refs.z;
            const float ea = lo ? e1 : e0, eb = lo ? e3 : e2;
            r = hi ? rb : ra;
            e = hi ? eb : ea;
        }
    };
    if(stack_fast(stack, sp)){
        // the three possible pushes land in shared-memory rows: straight-line code, one predicated store each
#pragma unroll
        for(int k = 0; k < 4; k++){
            uint32_t r; float e; bool hit;
            child(k, r, e, hit);
            const bool want = hit && cur != REF_EMPTY;
            if(k > 0)stack_st_sm_if(stack, sp, make_uint2(cur, __float_as_uint(cure)), want);
            sp += want ? 1 : 0;
            cur = hit ? r : cur;
            cure = hit ? e : cure;
        }
    } else{
#pragma unroll
        for(int k = 0; k < 4; k++){
            uint32_t r; float e; bool hit;
            child(k, r, e, hit);
            const bool want = hit && cur != REF_EMPTY;
            const bool push = want && sp < kStack;
            ovf |= want && !push;
            if(push)stack.st(sp, make_uint2(cur, __float_as_uint(cure)));
            sp += push ? 1 : 0;
            cur = hit ? r : cur;
            cure = hit ? e : cure;
        }
    }
    s.sp = sp;
    s.cur = cur;
    if(ovf)s.flags |= TF_OVERFLOW;
}

// Fast BVH4 cluster (pt_types.h Bvh4QNode): two 256-bit loads.  The slab planes of child c along x are
// org.x + q * step.x; with a = step.x * inv.x and b = (org.x - o.x) * inv.x every plane distance is one FMA
// q * a + b.  The boxes were rounded outwards by at least a quarter step at build time, far more than the rounding
// of these FMAs, so no exact box is ever missed; `e <= tmax` (not the reference's strict `<`) keeps the superset
// property for candidates at exactly tmax.  Visit order and pop-time cull as in node_step4.
__device__ __forceinline__ float qbyte(uint32_t w, int c){ return __uint2float_rn((w >> (8 * c)) & 0xFFu); }

#ifndef PTB_SORTED_PUSH
#define PTB_SORTED_PUSH 1
#endif
constexpr bool kSortedPush4q = PTB_SORTED_PUSH != 0;
template <bool ANY, class STK>
__device__ __forceinline__ void node_step4q(TravState& s, const DevScene& S, const STK& stack){
    const float4* p = reinterpret_cast<const float4*>(S.n4q + s.cur);
    const F8 w0 = ldg256(p), w1 = ldg256(p + 2);
    const uint32_t exw = __float_as_uint(w0.lo.w);
    const uint4 refs = make_uint4(__float_as_uint(w0.hi.x), __float_as_uint(w0.hi.y), __float_as_uint(w0.hi.z), __float_as_uint(w0.hi.w));
    const uint2 ord = make_uint2(__float_as_uint(w1.hi.z), __float_as_uint(w1.hi.w));
    const float ax = __uint_as_float((exw & 0xFFu) << 23) * s.r.ix, ay = __uint_as_float(((exw >> 8) & 0xFFu) << 23) * s.r.iy,
                az = __uint_as_float(((exw >> 16) & 0xFFu) << 23) * s.r.iz;
    const float bx = (w0.lo.x - s.r.ox) * s.r.ix, by = (w0.lo.y - s.r.oy) * s.r.iy, bz = (w0.lo.z - s.r.oz) * s.r.iz;
    // near / far plane words per axis by the sign of the ray's reciprocal direction: the slab interval needs no min / max
    const bool nx = s.r.ix < 0.0f, ny = s.r.iy < 0.0f, nz = s.r.iz < 0.0f;
    const uint32_t qnx = nx ? __float_as_uint(w1.lo.w) : __float_as_uint(w1.lo.x), qfx = nx ? __float_as_uint(w1.lo.x) : __float_as_uint(w1.lo.w);
    const uint32_t qny = ny ? __float_as_uint(w1.hi.x) : __float_as_uint(w1.lo.y), qfy = ny ? __float_as_uint(w1.lo.y) : __float_as_uint(w1.hi.x);
    const uint32_t qnz = nz ? __float_as_uint(w1.hi.y) : __float_as_uint(w1.lo.z), qfz = nz ? __float_as_uint(w1.lo.z) : __float_as_uint(w1.hi.y);
    const float tmax = s.tmax;
    float e[4];
    unsigned mask = 0;
#pragma unroll
    for(int c = 0; c < 4; c++){
        const float en = fmaxf(fmaxf(fmaf(qbyte(qnx, c), ax, bx), fmaf(qbyte(qny, c), ay, by)), fmaf(qbyte(qnz, c), az, bz));
        const float ex = fminf(fminf(fmaf(qbyte(qfx, c), ax, bx), fmaf(qbyte(qfy, c), ay, by)), fmaf(qbyte(qfz, c), az, bz));
        e[c] = en;
        if(ex >= kShadowEps && en <= fminf(tmax, ex))mask |= 1u << c;
    }
    // (an absent child is stored with its near planes at 255 and its far planes at 0 steps: it never passes, whatever the ray)
    uint32_t cur = REF_EMPTY;
    float cure = 0.0f;
    int sp = s.sp;
    const unsigned word = s.octant < 4 ? ord.x : ord.y;
    const unsigned code = (word >> ((s.octant & 3u) * 8u)) & 0xFFu;
    bool ovf = false;
    // child k of the visit sequence (k = 0 is the LAST visited lane): its ref, entry distance, and whether the ray enters it
    auto child = [&](int k, uint32_t& r, float& en, bool& hit){
        if(ANY){
            // any-hit: the answer does not depend on the visit order (the reference pushes unordered too, BVH.hpp:1019-1109)
            hit = ((mask >> k) & 1u) != 0;
            r = k == 0 ? refs.x : k == 1 ? refs.y : k == 2 ? refs.z : refs.w;
            en = e[k];
        } else{
            const unsigned lane = (code >> (2 * k)) & 3u;
            hit = ((mask >> lane) & 1u) != 0;
            const bool hi = (lane & 2u) != 0, lo = (lane & 1u) != 0;
            const uint32_t ra = lo ? refs.y : refs.x, rb = lo ? refs.w : refs.z;
            const float ea = lo ? e[1] : e[0], eb = lo ? e[3] : e[2];
            r = hi ? rb : ra;
            en = hi ? eb : ea;
        }
    };
    if(!ANY && kSortedPush4q){
        // closest hit, fast layout: children ordered by their entry distance with a 5-exchange network instead of the
        // cluster's per-octant visit bytes (which the exact layout has to follow to reproduce the reference's tie-breaks):
        // 25 instructions instead of 48 for the decode, and the nearest child really comes first.  Hits differ from the
        // octant order only where two triangles are hit at bit-equal t (north_star: "ties excepted").
        float k0 = (mask & 1u) ? e[0] : CUDART_INF_F, k1 = (mask & 2u) ? e[1] : CUDART_INF_F, k2 = (mask & 4u) ? e[2] : CUDART_INF_F, k3 = (mask & 8u) ? e[3] : CUDART_INF_F;
        uint32_t r0 = refs.x, r1 = refs.y, r2 = refs.z, r3 = refs.w;
#define PTB_CAS(ka, kb, ra, rb) { const bool sw = kb < ka; const float ta = sw ? kb : ka, tb = sw ? ka : kb; const uint32_t ua = sw ? rb : ra, ub = sw ? ra : rb; ka = ta; kb = tb; ra = ua; rb = ub; }
        PTB_CAS(k0, k1, r0, r1) PTB_CAS(k2, k3, r2, r3) PTB_CAS(k0, k2, r0, r2) PTB_CAS(k1, k3, r1, r3) PTB_CAS(k1, k2, r1, r2)
#undef PTB_CAS
        const bool v0 = k0 < CUDART_INF_F, v1 = k1 < CUDART_INF_F, v2 = k2 < CUDART_INF_F, v3 = k3 < CUDART_INF_F;
        if(stack_fast(stack, sp)){
            // the three possible pushes land in shared-memory rows: one predicated store each
            stack_st_sm_if(stack, sp, make_uint2(r3, __float_as_uint(k3)), v3);
            sp += v3 ? 1 : 0;
            stack_st_sm_if(stack, sp, make_uint2(r2, __float_as_uint(k2)), v2);
            sp += v2 ? 1 : 0;
            stack_st_sm_if(stack, sp, make_uint2(r1, __float_as_uint(k1)), v1);
            sp += v1 ? 1 : 0;
        } else{
            const uint32_t rr[3] = { r3, r2, r1 };
            const float kk[3] = { k3, k2, k1 };
            const bool vv[3] = { v3, v2, v1 };
#pragma unroll
            for(int q = 0; q < 3; q++){
                const bool push = vv[q] && sp < kStack;
                ovf |= vv[q] && !push;
                if(push)stack.st(sp, make_uint2(rr[q], __float_as_uint(kk[q])));
                sp += push ? 1 : 0;
            }
        }
        cur = v0 ? r0 : REF_EMPTY;
    } else if(stack_fast(stack, sp)){
        // the three possible pushes land in shared-memory rows: straight-line code, one predicated store each
#pragma unroll
        for(int k = 0; k < 4; k++){
            uint32_t r; float en; bool hit;
            child(k, r, en, hit);
            const bool want = hit && cur != REF_EMPTY;
            if(k > 0)stack_st_sm_if(stack, sp, make_uint2(cur, __float_as_uint(cure)), want);
            sp += want ? 1 : 0;
            cur = hit ? r : cur;
            cure = hit ? en : cure;
        }
    } else{
#pragma unroll
        for(int k = 0; k < 4; k++){
            uint32_t r; float en; bool hit;
            child(k, r, en, hit);
            const bool want = hit && cur != REF_EMPTY;
            const bool push = want && sp < kStack;
            ovf |= want && !push;
            if(push)stack.st(sp, make_uint2(cur, __float_as_uint(cure)));
            sp += push ? 1 : 0;
            cur = hit ? r : cur;
            cure = hit ? en : cure;
        }
    }
    s.sp = sp;
    s.cur = cur;
    if(ovf)s.flags |= TF_OVERFLOW;
}

// BVH8 cluster (BVH8::Intersect / IntersectPred, BVH.hpp:1374-1540): the same slab test and strict
// `tEntry < max` as BVH4, children pushed in the per-octant order of buildPerm so that the nearest pops
// first, entry-distance cull at pop.  Two 4-lane halves of three 256-bit loads each, one for the refs, and
// the 4-byte visit code of this ray's octant: all eight sectors of the cluster.
template <class T>
__device__ __forceinline__ T sel8(const T (&a)[8], unsigned lane){
    const bool b0 = (lane & 1u) != 0, b1 = (lane & 2u) != 0, b2 = (lane & 4u) != 0;
    const T t0 = b0 ? a[1] : a[0], t1 = b0 ? a[3] : a[2], t2 = b0 ? a[5] : a[4], t3 = b0 ? a[7] : a[6];
    const T u0 = b1 ? t1 : t0, u1 = b1 ? t3 : t2;
    return b2 ? u1 : u0;
}

template <bool ANY, class STK>
__device__ __forceinline__ void node_step8(TravState& s, const DevScene& S, const STK& stack){
    const float4* p = reinterpret_cast<const float4*>(S.n8 + s.cur);
    float e[8], x;
    unsigned mask = 0;
    const float tmax = s.tmax;
#pragma unroll
    for(int h = 0; h < 2; h++){
        const F8 wx = ldg256(p + 6 * h), wy = ldg256(p + 6 * h + 2), wz = ldg256(p + 6 * h + 4);
        slab(s.r, wx.lo.x, wy.lo.x, wz.lo.x, wx.hi.x, wy.hi.x, wz.hi.x, e[4 * h + 0], x); if(x >= kShadowEps && e[4 * h + 0] < tmax && e[4 * h + 0] <= x)mask |= 1u << (4 * h);
        slab(s.r, wx.lo.y, wy.lo.y, wz.lo.y, wx.hi.y, wy.hi.y, wz.hi.y, e[4 * h + 1], x); if(x >= kShadowEps && e[4 * h + 1] < tmax && e[4 * h + 1] <= x)mask |= 2u << (4 * h);
        slab(s.r, wx.lo.z, wy.lo.z, wz.lo.z, wx.hi.z, wy.hi.z, wz.hi.z, e[4 * h + 2], x); if(x >= kShadowEps && e[4 * h + 2] < tmax && e[4 * h + 2] <= x)mask |= 4u << (4 * h);
        slab(s.r, wx.lo.w, wy.lo.w, wz.lo.w, wx.hi.w, wy.hi.w, wz.hi.w, e[4 * h + 3], x); if(x >= kShadowEps && e[4 * h + 3] < tmax && e[4 * h + 3] <= x)mask |= 8u << (4 * h);
    }
    const F8 wr = ldg256(p + 12);
    const uint32_t refs[8] = { __float_as_uint(wr.lo.x), __float_as_uint(wr.lo.y), __float_as_uint(wr.lo.z), __float_as_uint(wr.lo.w),
                               __float_as_uint(wr.hi.x), __float_as_uint(wr.hi.y), __float_as_uint(wr.hi.z), __float_as_uint(wr.hi.w) };
    const unsigned code = __ldg(reinterpret_cast<const unsigned*>(p + 14) + s.octant);
    uint32_t cur = REF_EMPTY;
    float cure = 0.0f;
    int sp = s.sp;
    bool ovf = false;
#pragma unroll
    for(int k = 7; k >= 0; k--){                 // far to near: the nearest hit lane ends up in `cur`
        const unsigned lane = (code >> (3 * k)) & 7u;
        const uint32_t r = sel8(refs, lane);
        const float en = sel8(e, lane);
        const bool hit = ((mask >> lane) & 1u) != 0 && r != REF_EMPTY;   // absent lanes have zero boxes, never children
        const bool want = hit && cur != REF_EMPTY;
        const bool push = want && sp < kStack;
        ovf |= want && !push;
        if(push)stack.st(sp, make_uint2(cur, __float_as_uint(cure)));
        sp += push ? 1 : 0;
        cur = hit ? r : cur;
        cure = hit ? en : cure;
    }
    s.sp = sp;
    s.cur = cur;
    if(ovf)s.flags |= TF_OVERFLOW;
}

// ---- one round of the state machine -------------------------------------------------------

// Where the world-space ray comes from: a ray buffer (read-only loads), or registers for a caller that
// changes the ray between traversals inside one kernel (the Scene::IntersectTr walk in render.cu).
struct RayRegs { float4 a, b; };
__device__ __forceinline__ void fetch_ray(const float4* rays, size_t i, float4& a, float4& b){ a = __ldg(rays + 2 * i); b = __ldg(rays + 2 * i + 1); }
__device__ __forceinline__ void fetch_ray(const RayRegs& r, size_t, float4& a, float4& b){ a = r.a; b = r.b; }
// a ray buffer that the copy engine is still filling while the kernel runs (streamed host-buffer ray cast): the
// non-coherent path (ld.global.nc) is specified only for data that stays read-only for the whole kernel, so
// these reads go through L2 (ld.global.cg)
struct CgRays { const float4* p; };
#if defined(__CUDA_ARCH__)
__device__ __forceinline__ void fetch_ray(const CgRays& r, size_t i, float4& a, float4& b){ a = __ldcg(r.p + 2 * i); b = __ldcg(r.p + 2 * i + 1); }
#else
__device__ __forceinline__ void fetch_ray(const CgRays& r, size_t i, float4& a, float4& b){ a = r.p[2 * i]; b = r.p[2 * i + 1]; }
#endif

// compact rays of the host-buffer pipeline (pt_trace_closest_compact): {o, d} in 24 B, tmax = inf, time = 0; 8-byte aligned
struct CgRays24 { const float2* p; };
#if defined(__CUDA_ARCH__)
__device__ __forceinline__ void fetch_ray(const CgRays24& r, size_t i, float4& a, float4& b){
    const float2 x = __ldcg(r.p + 3 * i), y = __ldcg(r.p + 3 * i + 1), z = __ldcg(r.p + 3 * i + 2);
    a = make_float4(x.x, x.y, y.x, CUDART_INF_F); b = make_float4(y.y, z.x, z.y, 0.0f);
}
#else
inline void fetch_ray(const CgRays24& r, size_t i, float4& a, float4& b){
    const float2 x = r.p[3 * i], y = r.p[3 * i + 1], z = r.p[3 * i + 2];
    a = make_float4(x.x, x.y, y.x, INFINITY); b = make_float4(y.y, z.x, z.y, 0.0f);
}
#endif

// One round; the ray is finished once s.flags & TF_DONE.
// The world-space ray is re-read from its source when a transformed instance is left (cheaper than
// holding 9 more registers).
template <int WIDTH, bool ANY, bool STATS, class STK, class Alpha, class RAYS>
__device__ __forceinline__ void trav_step(TravState& s, const DevScene& S, const STK& stack, const RAYS& rays, size_t ray_idx,
                                          TraceCounters* cnt, const Alpha& alpha, bool do_leaf = true){
    // do_leaf == false (decided per warp by the caller, k_trace): a lane that holds a leaf item keeps it for a later
    // round, so that the leaf code runs when enough lanes of the warp have one (postponed leaf processing)
    if(!do_leaf && (s.cur & REF_LEAF) && s.cur != REF_EMPTY)return;
    // -- pop (BVH4 closest: cull by entry distance, BVH.hpp:1135) --
    if(s.cur == REF_EMPTY){
        // (a tighter loop for the all-shared-memory case was tried: the compiler then drops the convergence point behind it and the
        // lanes that leave early run the cluster step on their own, 10.8 -> 17 ms; with the pop hoisted to k_trace's converged loop
        // plus __syncwarp it is 10.9 ms -- no gain, profiles/r02_trace.md)
        while(s.sp > 0){
            const uint2 e = stack_ld(stack, --s.sp);
            if(WIDTH >= 4 && !ANY && __uint_as_float(e.y) > s.tmax)continue;
            if(e.x == REF_EMPTY)continue;      // empty lane of a cluster (BVH.hpp:38-43)
            s.cur = e.x;
            break;
        }
        if(s.cur == REF_EMPTY){ s.flags |= TF_DONE; return; }
    }
    // -- inner node --
    if(!(s.cur & REF_LEAF)){
        if(STATS){ if(s.flags & TF_IN_INST)cnt->blas_nodes++; else cnt->tlas_nodes++; }
        if(WIDTH == 2)node_step2<ANY>(s, S, stack);
        else if(WIDTH == 4)node_step4<ANY>(s, S, stack);
        else if(WIDTH == kWidthFast4)node_step4q<ANY>(s, S, stack);
        else node_step8<ANY>(s, S, stack);
    }
    // -- leaf items --
    if(do_leaf && (s.cur & REF_LEAF) && s.cur != REF_EMPTY){
        if(!Alpha::kSingle && s.cur == REF_LEAVE){
            // end of TransformedPrimitive::Intersect (Primitive.cpp:54-63): t back to world space
            if(s.flags & TF_XF){
                if(!stack_load_world_ray(stack, s.r)){
                    float4 a, b;
                    fetch_ray(rays, ray_idx, a, b);
                    make_ray(s.r, a.x, a.y, a.z, b.x, b.y, b.z);
                }
                s.octant = ray_octant(s.r);
                float t = f_div(s.tmax, s.len);
                if((Alpha::kWrap || STATS) && (s.flags & TF_ANIM2)){
                    // the outer TransformedPrimitive(translate) divides once more by its own length
                    const uint2 e = stack.ld(--s.sp);
                    t = f_div(t, __uint_as_float(e.y));
                }
                s.tmax = (s.flags & TF_FOUND_LOCAL) ? t : s.wtmax;
            }
            s.flags &= ~(TF_IN_INST | TF_XF | TF_FOUND_LOCAL | TF_ANIM2);
            s.cur = REF_EMPTY;
        } else if(!Alpha::kSingle && !(s.flags & TF_IN_INST)){
            // TLAS leaf item: BVHBase::intersectPrimitives over shared_ptr<Primitive> (BVH.hpp:107-135),
            // each either the BLAS itself or a TransformedPrimitive (Primitive.cpp:41-66)
            const uint32_t i = s.cur & ~REF_LEAF;
            const InstData d = load_inst(S.inst, i);
            if(STATS)cnt->instances++;
            const uint32_t nxt = i + 1 + ((d.meta.y >> INST_EXTRA_SHIFT) & INST_EXTRA_MASK);
            if((Alpha::kWrap || STATS) && (d.meta.y & (INST_ANIM | INST_WRAP))){
                // wrapper chain (pt_types.h): [AnimatedPrimitive]? [TransformedPrimitive]? (BLAS | sphere | quad)
                RayCtx r = s.r;
                float tm = s.tmax, len_a = 1.0f, len_b = 1.0f;
                uint32_t j = i;
                InstData c = d;
                const bool anim = (d.meta.y & INST_ANIM) != 0;
                if(anim){
                    float4 a, b;
                    fetch_ray(rays, ray_idx, a, b);
                    anim_ray(d, b.w, r, tm, len_a);
                    c = load_inst(S.inst, ++j);
                }
                const bool xf = (c.meta.y & INST_XF) != 0;
                if(xf)xf_ray(c, r, tm, len_b);
                if(c.meta.y & INST_WRAP)c = load_inst(S.inst, ++j);
                if(c.meta.y & (INST_SPHERE | INST_QUAD)){
                    float t, u = 0.0f, v = 0.0f;
                    bool hit = (c.meta.y & INST_SPHERE) ? sphere_hit(r, c, tm, t) : quad_hit(r, c, tm, t, u, v);
                    if(hit && (c.meta.y & INST_ALPHA))hit = shape_alpha(alpha, r, c, t, u, v, i, ray_idx);
                    if(hit){
                        if(ANY){ s.flags |= TF_OCCLUDED | TF_DONE; return; }
                        if(xf)t = f_div(t, len_b);
                        if(anim)t = f_div(t, len_a);
                        s.tmax = t; s.hu = u; s.hv = v; s.hprim = 0u; s.hinst = d.meta.z;
                    }
                    s.cur = (d.meta.y & INST_LAST) ? REF_EMPTY : (REF_LEAF | nxt);
                    return;
                }
                // AnimatedPrimitive over a BLAS instance
                if(!(d.meta.y & INST_LAST))trav_push(stack, s.sp, s.flags, REF_LEAF | nxt, -CUDART_INF_F);
                if(xf){ trav_push(stack, s.sp, s.flags, REF_EMPTY, len_a); s.flags |= TF_ANIM2; }
                trav_push(stack, s.sp, s.flags, REF_LEAVE, -CUDART_INF_F);
                s.flags |= TF_IN_INST | TF_XF;
                s.inst_rec = i; s.inst_id = d.meta.z;
                s.wtmax = s.tmax; s.len = xf ? len_b : len_a;
                stack_save_world_ray(stack, s.r);
                s.r = r; s.tmax = tm;
                s.octant = ray_octant(s.r);
                s.cur = c.meta.x;
                if(WIDTH == 2){
                    const float rlo[3] = { c.b0.x, c.b0.y, c.b0.z }, rhi[3] = { c.b0.w, c.b1.x, c.b1.y };
                    if(!box_hit2(s.r, rlo, rhi, s.tmax))s.cur = REF_EMPTY;
                }
                return;
            }
            if((Alpha::kWrap || STATS) && (d.meta.y & (INST_SPHERE | INST_QUAD))){
                // GeometricPrimitive over an analytic shape sitting directly in the TLAS leaf
                float t, u = 0.0f, v = 0.0f;
                bool hit = (d.meta.y & INST_SPHERE) ? sphere_hit(s.r, d, s.tmax, t) : quad_hit(s.r, d, s.tmax, t, u, v);
                if(hit && (d.meta.y & INST_ALPHA))hit = shape_alpha(alpha, s.r, d, t, u, v, i, ray_idx);
                if(hit){
                    if(ANY){ s.flags |= TF_OCCLUDED | TF_DONE; return; }
                    s.tmax = t; s.hu = u; s.hv = v; s.hprim = 0u; s.hinst = d.meta.z;
                }
                s.cur = (d.meta.y & INST_LAST) ? REF_EMPTY : (REF_LEAF | nxt);
                return;
            }
            if(WIDTH != 2 && (d.meta.y & INST_WBOX)){
                // the instance's own world box first (pt_types.h INST_WBOX); a miss costs one slab test instead of the transform
                const float wlo[3] = { d.b0.x, d.b0.y, d.b0.z }, whi[3] = { d.b0.w, d.b1.x, d.b1.y };
                if(!box_hit2(s.r, wlo, whi, s.tmax)){
                    s.cur = (d.meta.y & INST_LAST) ? REF_EMPTY : (REF_LEAF | nxt);
                    return;
                }
            }
            if(!(d.meta.y & INST_LAST))trav_push(stack, s.sp, s.flags, REF_LEAF | nxt, -CUDART_INF_F);
            trav_push(stack, s.sp, s.flags, REF_LEAVE, -CUDART_INF_F);
            s.flags |= TF_IN_INST;
            s.inst_rec = i; s.inst_id = d.meta.z;
            if(d.meta.y & INST_XF){
                s.wtmax = s.tmax;
                stack_save_world_ray(stack, s.r);
                xf_ray(d, s.r, s.tmax, s.len);
                s.octant = ray_octant(s.r);
                s.flags |= TF_XF;
            }
            s.cur = d.meta.x;
            if(WIDTH == 2){
                const float rlo[3] = { d.b0.x, d.b0.y, d.b0.z }, rhi[3] = { d.b0.w, d.b1.x, d.b1.y };
                if(!box_hit2(s.r, rlo, rhi, s.tmax))s.cur = REF_EMPTY;
            }
        } else{
            // BLAS leaf item: one triangle of BVHBase::intersectPrimitives[Pred] (BVH.hpp:107-135)
            const uint32_t i = s.cur & ~REF_LEAF;
            const TriData T = WIDTH == kWidthFast4 ? load_tri64(S.tris64, i) : load_tri(S.tris, i);
            const uint32_t tf = __float_as_uint(T.c.w);
            if(STATS)cnt->triangles++;
            float t, u, v;
            if(ANY){
                bool occ;
                if(tf & TRI_ALPHA_PATH){
                    // material->HasAlpha(): shape->Intersect + alpha test (Primitive.cpp:6-14)
                    occ = tri_closest(s.r, T, s.tmax, t, u, v) &&
                          (!(tf & TRI_ALPHA_TEX) || alpha(__float_as_uint(T.b.w), __float_as_uint(T.a.w), s.inst_rec, u, v, ray_idx));
                } else occ = tri_any(s.r, T, s.tmax);
                if(occ){ s.flags |= TF_OCCLUDED | TF_DONE; return; }
            } else{
                if(tri_closest(s.r, T, s.tmax, t, u, v)){
                    if(!(tf & TRI_ALPHA_TEX) || alpha(__float_as_uint(T.b.w), __float_as_uint(T.a.w), s.inst_rec, u, v, ray_idx)){
                        s.tmax = t; s.hu = u; s.hv = v; s.hprim = __float_as_uint(T.a.w); s.hinst = s.inst_id;
                        s.flags |= TF_FOUND_LOCAL;
                    }
                }
            }
            s.cur = (tf & TRI_LAST) ? REF_EMPTY : (REF_LEAF | (i + 1));
        }
    }
}

// Whole-ray convenience wrapper (host emulation tests, counting kernel).
template <int WIDTH, bool ANY, bool STATS, class Alpha, class RAYS>
__device__ __forceinline__ bool trace_scene(const DevScene& S, const RAYS& rays, size_t ray_idx, HitRec& hit, TraceCounters* cnt, const Alpha& alpha){
    uint2 stack_mem[kStack];
    LocalStack stack; stack.lm = stack_mem;
    TravState s;
    float4 ra, rb;
    fetch_ray(rays, ray_idx, ra, rb);
    trav_init<WIDTH>(s, S, ra, rb);
    while(!(s.flags & TF_DONE))trav_step<WIDTH, ANY, STATS>(s, S, stack, rays, ray_idx, cnt, alpha);
    const bool found = s.hprim != 0xFFFFFFFFu;
    hit.t = found ? s.tmax : CUDART_INF_F; hit.u = s.hu; hit.v = s.hv; hit.prim = s.hprim; hit.inst = s.hinst;
    hit.overflow = (s.flags & TF_OVERFLOW) ? 1u : 0u;
    return ANY ? (s.flags & TF_OCCLUDED) != 0 : found;
}

} // namespace ptb
