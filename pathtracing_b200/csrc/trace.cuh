// trace.cuh — device-side BVH2/BVH4 TLAS+BLAS traversal for sm_100a.
//
// One thread owns one ray.  Every node/triangle/instance record is fetched
// with 128-bit loads (pt_types.h); the traversal stack lives in the thread's
// local memory window (L1-resident).  The arithmetic of every box and triangle
// test is written with explicit round-to-nearest intrinsics in the operation
// order of the reference, so that no FMA contraction can change which
// candidate wins (SURVEY.md §7 "hard part 1"):
//   box     : AABB::HitDistance (AABB.hpp:72-111) / BVH4 SSE slab test (BVH.hpp:1142-1189)
//   closest : glm::intersectRayTriangle (glm/gtx/intersect.inl:29-94) + Shape.cpp:204
//   any-hit : TriangleShape::IntersectPred (Shape.cpp:246-268)
//   instance: TransformedPrimitive::Intersect/IntersectPred (Primitive.cpp:41-66)
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include "pt_types.h"

namespace ptb {

constexpr float kShadowEps = 0.00001f;          // shadowEpsilon, AABB.hpp:6
constexpr float kFltEps = 1.1920928955078125e-7f;
constexpr int kStackTlas = 32;
constexpr int kStackBlas = 64;

struct RayCtx {
    float ox, oy, oz, dx, dy, dz, ix, iy, iz;
};

struct HitRec {
    float t, u, v;
    uint32_t prim, inst;
};

struct TraceCounters {
    unsigned long long tlas_nodes, blas_nodes, instances, triangles;
};

__device__ __forceinline__ float f_sub(float a, float b){ return __fsub_rn(a, b); }
__device__ __forceinline__ float f_add(float a, float b){ return __fadd_rn(a, b); }
__device__ __forceinline__ float f_mul(float a, float b){ return __fmul_rn(a, b); }
__device__ __forceinline__ float f_div(float a, float b){ return __fdiv_rn(a, b); }

// Ray ctor, Ray.hpp:32-37
__device__ __forceinline__ float inv_dir(float d){ return fabsf(d) < 1e-32f ? 1e32f : f_div(1.0f, d); }

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
struct NoAlpha {
    __device__ __forceinline__ bool operator()(uint32_t, uint32_t, float, float) const{ return true; }
};

// Leaf loop: BVHBase::intersectPrimitives / intersectPrimitivesPred (BVH.hpp:107-135)
template <bool ANY, bool STATS, class Alpha>
__device__ __forceinline__ bool leaf_tris(const DevScene& S, const RayCtx& r, uint32_t first, float& tmax, HitRec& hit, TraceCounters* cnt, const Alpha& alpha){
    bool found = false;
    for(uint32_t i = first;; i++){
        const TriData T = load_tri(S.tris, i);
        const uint32_t flags = __float_as_uint(T.c.w);
        if(STATS)cnt->triangles++;
        float t, u, v;
        if(ANY){
            if(flags & TRI_ALPHA_PATH){
                if(tri_closest(r, T, tmax, t, u, v)){
                    if(!(flags & TRI_ALPHA_TEX) || alpha(__float_as_uint(T.b.w), __float_as_uint(T.a.w), u, v))return true;
                }
            } else if(tri_any(r, T, tmax))return true;
        } else{
            if(tri_closest(r, T, tmax, t, u, v)){
                if(!(flags & TRI_ALPHA_TEX) || alpha(__float_as_uint(T.b.w), __float_as_uint(T.a.w), u, v)){
                    tmax = t; hit.t = t; hit.u = u; hit.v = v; hit.prim = __float_as_uint(T.a.w);
                    found = true;
                }
            }
        }
        if(flags & TRI_LAST)break;
    }
    return found;
}

// ---- BVH2 ---------------------------------------------------------------------------

// BVH2::Intersect (BVH.hpp:503-548) / IntersectPred (:476-500) over one BLAS.
template <bool ANY, bool STATS, class Alpha>
__device__ __forceinline__ bool blas2(const DevScene& S, const RayCtx& r, uint32_t root, const float* rlo, const float* rhi, float& tmax, HitRec& hit, TraceCounters* cnt, const Alpha& alpha){
    if(!box_hit2(r, rlo, rhi, tmax))return false;
    uint32_t stack[kStackBlas];
    int sp = 0;
    stack[sp++] = root;
    bool found = false;
    while(sp){
        const uint32_t ref = stack[--sp];
        if(ref & REF_LEAF){
            if(leaf_tris<ANY, STATS>(S, r, ref & ~REF_LEAF, tmax, hit, cnt, alpha)){
                if(ANY)return true;
                found = true;
            }
            continue;
        }
        const float4* p = reinterpret_cast<const float4*>(S.n2 + ref);
        const float4 q0 = __ldg(p), q1 = __ldg(p + 1), q2 = __ldg(p + 2);
        const uint4 q3 = __ldg(reinterpret_cast<const uint4*>(p + 3));
        if(STATS)cnt->blas_nodes++;
        float a0, b0, a1, b1;
        slab(r, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, a0, b0);
        slab(r, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, a1, b1);
        const bool h0 = a0 <= b0 && b0 >= kShadowEps && a0 <= tmax;
        const bool h1 = a1 <= b1 && b1 >= kShadowEps && a1 <= tmax;
        if(ANY){
            if(h0 && sp < kStackBlas)stack[sp++] = q3.x;
            if(h1 && sp < kStackBlas)stack[sp++] = q3.y;
        } else{
            const float da = q3.z == 0 ? r.dx : (q3.z == 1 ? r.dy : r.dz);
            if(da > 0.0f){      // child 0 is nearer: push 1 then 0
                if(h1 && sp < kStackBlas)stack[sp++] = q3.y;
                if(h0 && sp < kStackBlas)stack[sp++] = q3.x;
            } else{
                if(h0 && sp < kStackBlas)stack[sp++] = q3.x;
                if(h1 && sp < kStackBlas)stack[sp++] = q3.y;
            }
        }
    }
    return found;
}

// ---- BVH4 ---------------------------------------------------------------------------

struct Slab4 { float entry[4]; unsigned mask; };

// 4-wide slab test of BVH4::Intersect (BVH.hpp:1142-1189): hit iff
// tExit >= eps && tEntry < max && tEntry <= tExit  (note the strict '<').
__device__ __forceinline__ Slab4 slab4(const RayCtx& r, const Bvh4Node* n, float tmax, uint4& refs, uint2& order){
    const float4* p = reinterpret_cast<const float4*>(n);
    const float4 xmin = __ldg(p), xmax = __ldg(p + 1), ymin = __ldg(p + 2), ymax = __ldg(p + 3), zmin = __ldg(p + 4), zmax = __ldg(p + 5);
    refs = __ldg(reinterpret_cast<const uint4*>(p + 6));
    const uint4 tail = __ldg(reinterpret_cast<const uint4*>(p + 7));
    order = make_uint2(tail.x, tail.y);
    Slab4 s; s.mask = 0;
    float a, b;
    slab(r, xmin.x, ymin.x, zmin.x, xmax.x, ymax.x, zmax.x, a, b); s.entry[0] = a; if(b >= kShadowEps && a < tmax && a <= b)s.mask |= 1u;
    slab(r, xmin.y, ymin.y, zmin.y, xmax.y, ymax.y, zmax.y, a, b); s.entry[1] = a; if(b >= kShadowEps && a < tmax && a <= b)s.mask |= 2u;
    slab(r, xmin.z, ymin.z, zmin.z, xmax.z, ymax.z, zmax.z, a, b); s.entry[2] = a; if(b >= kShadowEps && a < tmax && a <= b)s.mask |= 4u;
    slab(r, xmin.w, ymin.w, zmin.w, xmax.w, ymax.w, zmax.w, a, b); s.entry[3] = a; if(b >= kShadowEps && a < tmax && a <= b)s.mask |= 8u;
    return s;
}

__device__ __forceinline__ uint32_t pick4(const uint4& v, unsigned k){
    return k == 0 ? v.x : (k == 1 ? v.y : (k == 2 ? v.z : v.w));
}

template <bool ANY, bool STATS, class Alpha>
__device__ __forceinline__ bool blas4(const DevScene& S, const RayCtx& r, uint32_t root, float& tmax, HitRec& hit, TraceCounters* cnt, const Alpha& alpha){
    uint32_t stack[kStackBlas];
    float entry[kStackBlas];
    int sp = 0;
    entry[sp] = 0.0f;
    stack[sp++] = root;
    const unsigned octant = (r.dx < 0.0f ? 1u : 0u) | (r.dy < 0.0f ? 2u : 0u) | (r.dz < 0.0f ? 4u : 0u);
    bool found = false;
    while(sp){
        --sp;
        if(!ANY && entry[sp] > tmax)continue;       // BVH.hpp:1135
        const uint32_t ref = stack[sp];
        if(ref == REF_EMPTY)continue;
        if(ref & REF_LEAF){
            if(leaf_tris<ANY, STATS>(S, r, ref & ~REF_LEAF, tmax, hit, cnt, alpha)){
                if(ANY)return true;
                found = true;
            }
            continue;
        }
        uint4 refs; uint2 ord;
        const Slab4 s = slab4(r, S.n4 + ref, tmax, refs, ord);
        if(STATS)cnt->blas_nodes++;
        if(ANY){
            // BVH4::IntersectPred pushes lanes 0..3 unordered (BVH.hpp:1099-1102)
#pragma unroll
            for(unsigned k = 0; k < 4; k++)
                if((s.mask >> k) & 1u){ if(sp < kStackBlas){ stack[sp++] = pick4(refs, k); } }
        } else{
            // visit order byte for this octant; push far -> near
            const unsigned word = octant < 4 ? ord.x : ord.y;
            const unsigned code = (word >> ((octant & 3u) * 8u)) & 0xFFu;
#pragma unroll
            for(int k = 0; k < 4; k++){
                const unsigned lane = (code >> (2 * k)) & 3u;   // k=0 is the LAST visited lane
                if((s.mask >> lane) & 1u){
                    if(sp < kStackBlas){
                        entry[sp] = lane == 0 ? s.entry[0] : (lane == 1 ? s.entry[1] : (lane == 2 ? s.entry[2] : s.entry[3]));
                        stack[sp++] = pick4(refs, lane);
                    }
                }
            }
        }
    }
    return found;
}

// ---- instances ------------------------------------------------------------------------

struct InstData { float4 m0, m1, m2; uint4 meta; float4 b0; float4 b1; };

__device__ __forceinline__ InstData load_inst(const Instance* inst, uint32_t i){
    const float4* p = reinterpret_cast<const float4*>(inst + i);
    InstData d;
    d.m0 = __ldg(p); d.m1 = __ldg(p + 1); d.m2 = __ldg(p + 2);
    d.meta = __ldg(reinterpret_cast<const uint4*>(p + 3));
    d.b0 = __ldg(p + 4); d.b1 = __ldg(p + 5);
    return d;
}

// glm mat4 * vec4 restricted to xyz: (m0*x + m1*y) + (m2*z + m3*w)
__device__ __forceinline__ void xf_apply(const InstData& d, float x, float y, float z, float w, float& ox, float& oy, float& oz){
    // inv[c*3+r]: m0 = (c0r0,c0r1,c0r2,c1r0) m1 = (c1r1,c1r2,c2r0,c2r1) m2 = (c2r2,c3r0,c3r1,c3r2)
    ox = f_add(f_add(f_mul(d.m0.x, x), f_mul(d.m0.w, y)), f_add(f_mul(d.m1.z, z), f_mul(d.m2.y, w)));
    oy = f_add(f_add(f_mul(d.m0.y, x), f_mul(d.m1.x, y)), f_add(f_mul(d.m1.w, z), f_mul(d.m2.z, w)));
    oz = f_add(f_add(f_mul(d.m0.z, x), f_mul(d.m1.y, y)), f_add(f_mul(d.m2.x, z), f_mul(d.m2.w, w)));
}

// One TLAS leaf: loop over its instances (BVHBase::intersectPrimitives over
// shared_ptr<Primitive>), each either the BLAS itself or a TransformedPrimitive.
template <int WIDTH, bool ANY, bool STATS, class Alpha>
__device__ __forceinline__ bool leaf_instances(const DevScene& S, const RayCtx& r, uint32_t first, float& tmax, HitRec& hit, TraceCounters* cnt, const Alpha& alpha){
    bool found = false;
    for(uint32_t i = first;; i++){
        const InstData d = load_inst(S.inst, i);
        if(STATS)cnt->instances++;
        const float rlo[3] = { d.b0.x, d.b0.y, d.b0.z }, rhi[3] = { d.b0.w, d.b1.x, d.b1.y };
        bool h;
        if(d.meta.y & INST_XF){
            float ddx, ddy, ddz, oox, ooy, ooz;
            xf_apply(d, r.dx, r.dy, r.dz, 0.0f, ddx, ddy, ddz);
            const float len = sqrtf(f_add(f_add(f_mul(ddx, ddx), f_mul(ddy, ddy)), f_mul(ddz, ddz)));
            xf_apply(d, r.ox, r.oy, r.oz, 1.0f, oox, ooy, ooz);
            RayCtx lr;
            make_ray(lr, oox, ooy, ooz, f_div(ddx, len), f_div(ddy, len), f_div(ddz, len));
            float lmax = f_mul(tmax, len);
            HitRec lh;
            if(WIDTH == 2)h = blas2<ANY, STATS>(S, lr, d.meta.x, rlo, rhi, lmax, lh, cnt, alpha);
            else h = blas4<ANY, STATS>(S, lr, d.meta.x, lmax, lh, cnt, alpha);
            if(h && !ANY){
                hit = lh;
                hit.t = f_div(lh.t, len);   // Primitive.cpp:60
                tmax = hit.t;
            }
        } else{
            if(WIDTH == 2)h = blas2<ANY, STATS>(S, r, d.meta.x, rlo, rhi, tmax, hit, cnt, alpha);
            else h = blas4<ANY, STATS>(S, r, d.meta.x, tmax, hit, cnt, alpha);
        }
        if(h){
            if(ANY)return true;
            hit.inst = d.meta.z;
            found = true;
        }
        if(d.meta.y & INST_LAST)break;
    }
    return found;
}

// ---- TLAS -----------------------------------------------------------------------------

// Scene::Intersect / IntersectPred (Scene.cpp:4-6,31-33) -> TLASx -> instances -> BLASx.
template <int WIDTH, bool ANY, bool STATS, class Alpha>
__device__ __forceinline__ bool trace_scene(const DevScene& S, const RayCtx& r, float tmax, HitRec& hit, TraceCounters* cnt, const Alpha& alpha){
    hit.t = CUDART_INF_F; hit.u = 0.0f; hit.v = 0.0f; hit.prim = 0xFFFFFFFFu; hit.inst = 0xFFFFFFFFu;
    bool found = false;
    if(WIDTH == 2){
        if(!box_hit2(r, S.tlo, S.thi, tmax))return false;
        uint32_t stack[kStackTlas];
        int sp = 0;
        stack[sp++] = S.tlas_root;
        while(sp){
            const uint32_t ref = stack[--sp];
            if(ref & REF_LEAF){
                if(leaf_instances<2, ANY, STATS>(S, r, ref & ~REF_LEAF, tmax, hit, cnt, alpha)){
                    if(ANY)return true;
                    found = true;
                }
                continue;
            }
            const float4* p = reinterpret_cast<const float4*>(S.n2 + ref);
            const float4 q0 = __ldg(p), q1 = __ldg(p + 1), q2 = __ldg(p + 2);
            const uint4 q3 = __ldg(reinterpret_cast<const uint4*>(p + 3));
            if(STATS)cnt->tlas_nodes++;
            float a0, b0, a1, b1;
            slab(r, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, a0, b0);
            slab(r, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, a1, b1);
            const bool h0 = a0 <= b0 && b0 >= kShadowEps && a0 <= tmax;
            const bool h1 = a1 <= b1 && b1 >= kShadowEps && a1 <= tmax;
            const float da = q3.z == 0 ? r.dx : (q3.z == 1 ? r.dy : r.dz);
            if(!ANY && da > 0.0f){
                if(h1 && sp < kStackTlas)stack[sp++] = q3.y;
                if(h0 && sp < kStackTlas)stack[sp++] = q3.x;
            } else{
                if(h0 && sp < kStackTlas)stack[sp++] = q3.x;
                if(h1 && sp < kStackTlas)stack[sp++] = q3.y;
            }
        }
    } else{
        uint32_t stack[kStackTlas];
        float entry[kStackTlas];
        int sp = 0;
        entry[sp] = 0.0f;
        stack[sp++] = S.tlas_root;
        const unsigned octant = (r.dx < 0.0f ? 1u : 0u) | (r.dy < 0.0f ? 2u : 0u) | (r.dz < 0.0f ? 4u : 0u);
        while(sp){
            --sp;
            if(!ANY && entry[sp] > tmax)continue;
            const uint32_t ref = stack[sp];
            if(ref == REF_EMPTY)continue;
            if(ref & REF_LEAF){
                if(leaf_instances<4, ANY, STATS>(S, r, ref & ~REF_LEAF, tmax, hit, cnt, alpha)){
                    if(ANY)return true;
                    found = true;
                }
                continue;
            }
            uint4 refs; uint2 ord;
            const Slab4 s = slab4(r, S.n4 + ref, tmax, refs, ord);
            if(STATS)cnt->tlas_nodes++;
            if(ANY){
#pragma unroll
                for(unsigned k = 0; k < 4; k++)
                    if((s.mask >> k) & 1u){ if(sp < kStackTlas){ stack[sp++] = pick4(refs, k); } }
            } else{
                const unsigned word = octant < 4 ? ord.x : ord.y;
                const unsigned code = (word >> ((octant & 3u) * 8u)) & 0xFFu;
#pragma unroll
                for(int k = 0; k < 4; k++){
                    const unsigned lane = (code >> (2 * k)) & 3u;
                    if((s.mask >> lane) & 1u){
                        if(sp < kStackTlas){
                            entry[sp] = lane == 0 ? s.entry[0] : (lane == 1 ? s.entry[1] : (lane == 2 ? s.entry[2] : s.entry[3]));
                            stack[sp++] = pick4(refs, lane);
                        }
                    }
                }
            }
        }
    }
    return found;
}

} // namespace ptb
