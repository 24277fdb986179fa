// alpha.cuh — the alpha test GeometricPrimitive::Intersect / IntersectPred run on every candidate hit
// (Primitive.cpp:6-26): Material::Alpha(uv) of MicrofacetDiffuse / MicrofacetDielectric
// (Material.hpp:336-343, 572-579) through AlphaTester (Material.hpp:181-198).
// Only triangles flagged TRI_ALPHA_TEX reach it (scene_build.cpp): the rest can never be rejected.
#pragma once
#include "shade_math.cuh"

namespace ptb {

// AlphaTester's Blend mode draws random_float() from the calling thread's generator; any independent
// uniform has the same distribution, so the draw is a hash of (launch salt, ray, triangle, instance).
PT_HDI float alpha_uniform(uint64_t salt, uint64_t ray, uint32_t prim, uint32_t inst_rec){
    uint64_t x = salt ^ (ray * 0x9E3779B97F4A7C15ull) ^ ((uint64_t)prim << 32 | inst_rec);
    x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
    x ^= x >> 27; x *= 0x94D049BB133111EBull;
    x ^= x >> 31;
    return (float)(x >> 40) * (1.0f / 16777216.0f);
}

struct AlphaTex {
    static constexpr bool kWrap = true;      // the general traversal variant (trace.cuh NoAlpha)
    static constexpr bool kSingle = false;
    ShadeScene S;
    uint64_t salt;
    // Material::Alpha(uv) (Material.hpp:336-342, 572-578) through AlphaTester (:181-198)
    PT_HDI bool test(const DevMaterial& M, float u, float v, uint32_t prim, uint32_t inst_rec, size_t ray) const{
        const float a = M.alpha >= 0 ? tex_eval(S, M.alpha, u, v, v3(0.0f)).x : tex_alpha(S, M.tex, u, v);
        switch(M.alpha_mode){
        case PT_ALPHA_BLEND: return alpha_uniform(salt, (uint64_t)ray, prim, inst_rec) < a;
        case PT_ALPHA_MASK: return a > M.alpha_cutoff;
        default: return true;
        }
    }
    // a sphere / quad GeometricPrimitive: (a, b, c) = hit point - center (SphereShape::Intersect: uv = GetSphereUV of the
    // normal, Shape.cpp:11-17) or (a, b) = the quad's (alpha, beta) (Shape.cpp:333-340)
    PT_HDI bool shape(uint32_t material, bool sphere, float a, float b, float c, uint32_t inst_rec, size_t ray) const{
        float u = a, v = b;
        if(sphere)sphere_uv(v3(a, b, c), u, v);
        return test(S.materials[material], u, v, 0u, inst_rec, ray);
    }
    PT_HDI bool operator()(uint32_t mesh, uint32_t prim, uint32_t inst_rec, float bu, float bv, size_t ray) const{
        const DevMesh m = S.meshes[mesh];
        if(m.material < 0)return true;
        const DevMaterial& M = S.materials[m.material];
        // uv of the candidate hit as TriangleShape::Intersect interpolates it (Shape.cpp:205-210)
        const uint32_t tri = prim - m.prim_base;
        const uint32_t i0 = m.vtx_base + S.idx[m.idx_base + 3 * tri], i1 = m.vtx_base + S.idx[m.idx_base + 3 * tri + 1], i2 = m.vtx_base + S.idx[m.idx_base + 3 * tri + 2];
        const float w = 1.0f - bu - bv;
        const float u = bu * S.uv[2 * (size_t)i1] + bv * S.uv[2 * (size_t)i2] + w * S.uv[2 * (size_t)i0];
        const float v = bu * S.uv[2 * (size_t)i1 + 1] + bv * S.uv[2 * (size_t)i2 + 1] + w * S.uv[2 * (size_t)i0 + 1];
        return test(M, u, v, prim, inst_rec, ray);
    }
};

} // namespace ptb
