// shade.cuh — device-side evaluation over SoA hit records: surface interaction
// rebuild, materials, light sampling, media.  Each function restates one
// reference routine (cited) in fp32 with the reference's formulas, including its
// quirks (SURVEY.md §7 "hard part 3"), because the converged image has to match
// the reference's, not a textbook's.
#pragma once
#include "shade_math.cuh"
#include "pt_types.h"

namespace ptb {

constexpr float kEps = 0.00001f;    // shadowEpsilon, AABB.hpp:6

// SurfaceInteraction (Interaction.hpp:36-51) rebuilt from a 16 B hit record
struct SurfInt {
    V3 p, n, ns, tangent;
    float u, v, t;
    int32_t mat;        // material id or -1 (medium boundary)
    int32_t medium;     // mesh medium id or -1
    int32_t light;      // DevLight index of the area light on this triangle, or -1
    int32_t emis_tex;
};

PT_HDI V3 xf_point(const float* m, V3 p){   // m[c*3+r], c = 0..3 : (m0*x + m1*y) + (m2*z + m3)
    return v3((m[0] * p.x + m[3] * p.y) + (m[6] * p.z + m[9]),
              (m[1] * p.x + m[4] * p.y) + (m[7] * p.z + m[10]),
              (m[2] * p.x + m[5] * p.y) + (m[8] * p.z + m[11]));
}
PT_HDI V3 xf_vec(const float* m, V3 d){
    return v3((m[0] * d.x + m[3] * d.y) + m[6] * d.z, (m[1] * d.x + m[4] * d.y) + m[7] * d.z, (m[2] * d.x + m[5] * d.y) + m[8] * d.z);
}
PT_HDI V3 m3_mul(const float* m, V3 d){      // m[c*3+r]
    return v3(m[0] * d.x + m[3] * d.y + m[6] * d.z, m[1] * d.x + m[4] * d.y + m[7] * d.z, m[2] * d.x + m[5] * d.y + m[8] * d.z);
}

// Material::sample_normalMap (Material.hpp:340-344, 584-588)
PT_HDI V3 sample_normal_map(const ShadeScene& S, int mat, V3 ns, V3 tangent, float u, float v){
    if(mat < 0)return ns;
    const DevMaterial& M = S.materials[mat];
    if((M.kind != PT_MAT_DIFFUSE && M.kind != PT_MAT_DIELECTRIC) || M.norm < 0)return ns;
    const V3 c = tex_eval(S, M.norm, u, v, ns);
    const V3 nn = normalize(2.0f * c - v3(1.0f));
    return to_world(onb_from_frame(ns, tangent), nn);
}

// AnimatedPrimitive / AnimatedLight offset at a ray time (Primitive.cpp:82-89, Light.cpp:341-356):
// dir * (clamp(time - tb.x, tb.x, tb.y) / (tb.y - tb.x))
PT_HDI V3 anim_offset(const DevInstShade& I, float time){
    const float tt = fminf(fmaxf(time - I.atb[0], I.atb[0]), I.atb[1]) / (I.atb[1] - I.atb[0]);
    return ld3(I.adir) * tt;
}

// TriangleShape::Intersect's shading frame (Shape.cpp:205-243) + GeometricPrimitive::Intersect's
// stamping (Primitive.cpp:20-24) + TransformedPrimitive::Intersect's world transform (Primitive.cpp:54-63),
// applied once more with translate(dir * t) for an AnimatedPrimitive (Primitive.cpp:86-89)
PT_HDN void build_interaction(const ShadeScene& S, V3 ro, V3 rd, float time, float t, float bu, float bv, uint32_t prim, uint32_t inst, SurfInt& si){
    const DevInstShade& I = S.inst[inst];
    const float t_world = t;
    V3 off = v3(0.0f);
    if(I.animated){
        // the outer TransformedPrimitive(translate): origin back by the offset, direction renormalised
        off = anim_offset(I, time);
        ro = ro - off;
        const float la = length(rd);
        rd = rd / la; t = t * la;
    }
    if(I.shape >= 0){
        // SphereShape::Intersect (Shape.cpp:11-20) / QuadShape::Intersect (Shape.cpp:333-340), in the shape's space
        V3 o = ro, d = rd;
        float tl = t;
        if(I.has_xf){
            const V3 dd = xf_vec(I.inv, rd);
            const float len = length(dd);
            o = xf_point(I.inv, ro); d = dd / len; tl = t * len;
        }
        const V3 at = o + tl * d;
        V3 ns, n;
        if(I.shape == PT_SHAPE_SPHERE){
            ns = normalize(at - ld3(I.shp)); n = ns;
            sphere_uv(n, si.u, si.v);
        } else{
            n = ld3(I.shp + 9);
            ns = dot(d, n) > 0.0f ? -n : n;
            si.u = bu; si.v = bv;
        }
        const V3 up = fabsf(ns.x) > 0.9999f ? v3(0, 1, 0) : v3(1, 0, 0);
        V3 tangent = normalize(cross(up, ns));
        V3 p = at + kEps * ns;
        if(I.has_xf){
            p = xf_point(I.fwd, p);
            n = normalize(m3_mul(I.nrm, n));
            ns = normalize(m3_mul(I.nrm, ns));
            tangent = normalize(xf_vec(I.fwd, tangent));
        }
        si.tangent = tangent;
        si.p = p + off; si.n = n; si.ns = ns; si.t = t_world;
        si.mat = I.material; si.medium = I.medium; si.emis_tex = -1;
        si.light = I.light_base;
        return;
    }
    const ShadeTri st = S.stri[I.tri_base + prim];      // vertices and mesh in one load (shade_types.h)
    const DevMesh m = S.meshes[st.mesh];
    const uint32_t i0 = st.i0, i1 = st.i1, i2 = st.i2;
    const V3 v0 = ld3(S.pos + 3 * (size_t)i0), v1 = ld3(S.pos + 3 * (size_t)i1), v2 = ld3(S.pos + 3 * (size_t)i2);
    V3 o = ro, d = rd;
    float tl = t;
    if(I.has_xf){
        const V3 dd = xf_vec(I.inv, rd);
        const float len = length(dd);
        o = xf_point(I.inv, ro); d = dd / len; tl = t * len;
    }
    const float w = 1.0f - bu - bv;
    si.u = bu * S.uv[2 * (size_t)i1] + bv * S.uv[2 * (size_t)i2] + w * S.uv[2 * (size_t)i0];
    si.v = bu * S.uv[2 * (size_t)i1 + 1] + bv * S.uv[2 * (size_t)i2 + 1] + w * S.uv[2 * (size_t)i0 + 1];
    V3 nn = normalize(bu * ld3(S.nrm + 3 * (size_t)i1) + bv * ld3(S.nrm + 3 * (size_t)i2) + w * ld3(S.nrm + 3 * (size_t)i0));
    const V3 N = normalize(cross(v1 - v0, v2 - v0));
    if(dot(N, nn) < 0.0f)nn = -nn;
    V3 p = (o + tl * d) + kEps * N * (dot(d, N) > 0.0f ? -1.0f : 1.0f);
    V3 tangent;
    if(m.has_tangents){
        const V3 tg = bu * ld3(S.tan + 3 * (size_t)i1) + bv * ld3(S.tan + 3 * (size_t)i2) + w * ld3(S.tan + 3 * (size_t)i0);
        tangent = normalize(tg - nn * dot(nn, tg));
    } else{
        const V3 up = fabsf(nn.x) > 0.9999f ? v3(0, 1, 0) : v3(1, 0, 0);
        tangent = normalize(cross(up, nn));
    }
    V3 ns = sample_normal_map(S, m.material, nn, tangent, si.u, si.v);
    V3 n = N;
    if(I.has_xf){
        p = xf_point(I.fwd, p);
        n = normalize(m3_mul(I.nrm, n));
        ns = normalize(m3_mul(I.nrm, ns));
        tangent = normalize(xf_vec(I.fwd, tangent));
    }
    si.p = p + off; si.n = n; si.ns = ns; si.tangent = tangent; si.t = t_world;
    si.mat = m.material; si.medium = m.medium; si.emis_tex = m.emissive_tex;
    si.light = -1;
    if(m.emissive_tex >= 0 && I.light_base >= 0){
        const int32_t k = S.tri_emis[I.tri_base + prim];
        if(k >= 0){
            const int32_t li = I.light_base + k;
            if(S.lights[li].valid)si.light = li;
        }
    }
}

// ---- microfacet distribution (Material.hpp:55-142), isotropic use only ---------------------
struct Ggx { float a; };   // alphaX = alphaY = roughness^2
PT_HDI float ggx_lambda(Ggx g, V3 w){
    const float c2 = w.z * w.z;
    if(c2 == 0.0f)return 0.0f;
    const float s2 = fmaxf(0.0f, 1.0f - c2);
    const float s = sqrtf(s2);
    const float cp = s == 0.0f ? 1.0f : clampf(w.x / s, -1.0f, 1.0f);
    const float sp = s == 0.0f ? 0.0f : clampf(w.y / s, -1.0f, 1.0f);
    const float a2 = (cp * g.a) * (cp * g.a) + (sp * g.a) * (sp * g.a);
    return (sqrtf(1.0f + a2 * s2 / c2) - 1.0f) / 2.0f;
}
PT_HDI float ggx_D(Ggx g, V3 wh){
    const float c2 = wh.z * wh.z;
    if(c2 == 0.0f)return 0.0f;
    const float c4 = c2 * c2;
    const float s2 = fmaxf(0.0f, 1.0f - c2);
    const float s = sqrtf(s2);
    const float cp = s == 0.0f ? 1.0f : clampf(wh.x / s, -1.0f, 1.0f);
    const float sp = s == 0.0f ? 0.0f : clampf(wh.y / s, -1.0f, 1.0f);
    const float e = s2 / c2 * ((cp / g.a) * (cp / g.a) + (sp / g.a) * (sp / g.a));
    const float denom = kPi * g.a * g.a * c4 * (1.0f + e) * (1.0f + e);
    if(denom <= 0.0f)return INFINITY;
    return 1.0f / denom;
}
PT_HDI float ggx_G1(Ggx g, V3 w){ return 1.0f / (1.0f + ggx_lambda(g, w)); }
PT_HDI float ggx_G(Ggx g, V3 wo, V3 wi){ return 1.0f / (1.0f + ggx_lambda(g, wo) + ggx_lambda(g, wi)); }
PT_HDI float ggx_pdf(Ggx g, V3 wo, V3 wh){ return ggx_D(g, wh) * ggx_G1(g, wo) * fabsf(dot(wo, wh) / wo.z); }
// sampleGGXVNDF (Heitz 2018), Material.hpp:114-139
PT_HDI V3 ggx_vndf(V3 Ve, float a, float U1, float U2){
    const V3 Vh = normalize(v3(a * Ve.x, a * Ve.y, Ve.z));
    const float lensq = Vh.x * Vh.x + Vh.y * Vh.y;
    const V3 T1 = lensq > 0.0f ? v3(-Vh.y, Vh.x, 0.0f) * (1.0f / sqrtf(lensq)) : v3(1, 0, 0);
    const V3 T2 = cross(Vh, T1);
    const float r = sqrtf(U1);
    const float phi = 2.0f * kPi * U2;
    const float t1 = r * cosf(phi);
    float t2 = r * sinf(phi);
    const float s = 0.5f * (1.0f + Vh.z);
    t2 = (1.0f - s) * sqrtf(1.0f - t1 * t1) + s * t2;
    const V3 Nh = t1 * T1 + t2 * T2 + sqrtf(fmaxf(0.0f, 1.0f - t1 * t1 - t2 * t2)) * Vh;
    return normalize(v3(a * Nh.x, a * Nh.y, fmaxf(0.0f, Nh.z)));
}
PT_HDI V3 ggx_sample_wh(Ggx g, V3 wo, float u1, float u2){
    const bool flip = wo.z < 0.0f;
    V3 wh = ggx_vndf(flip ? -wo : wo, g.a, u1, u2);
    if(flip)wh = -wh;
    return wh;
}

// FresnelDielectric / FresnelSchlick (Material.hpp:11-32)
PT_HDI float fresnel_dielectric(float ci, float eta){
    ci = clampf(ci, -1.0f, 1.0f);
    if(ci < 0.0f){ eta = 1.0f / eta; ci = -ci; }
    const float s2i = 1.0f - ci * ci;
    const float s2t = s2i / (eta * eta);
    if(s2t >= 1.0f)return 1.0f;
    const float ct = sqrtf(1.0f - s2t);
    const float rpl = (eta * ci - ct) / (eta * ci + ct);
    const float rpp = (ci - eta * ct) / (ci + eta * ct);
    return (rpl * rpl + rpp * rpp) / 2.0f;
}
PT_HDI V3 fresnel_schlick(float c, V3 F0){ return F0 + (v3(1.0f) - F0) * powf(1.0f - c, 5.0f); }

} // namespace ptb
