// shade_mat.cuh — Material::{scatter, PDF, calc_attenuation} for the four material
// classes (Material.hpp:200-673), dispatched on DevMaterial::kind.
#pragma once
#include "shade.cuh"

namespace ptb {

enum : uint32_t { BX_TRANSMISSIVE = 1u, BX_SPECULAR = 2u };   // BxDFFlags, Material.hpp:35-39

struct Scatter {
    V3 f; float pdf; uint32_t flags;
    V3 o, d;            // scattered ray
};

// The material of one hit with its textures already evaluated at the hit's uv.  scatter / PDF /
// calc_attenuation of the reference each look the same textures up again (Material.hpp:241-262,
// 272-278, 309-323 ...); the values are identical, so they are fetched once per hit here.
struct MatCtx {
    int kind;
    float ri;
    V3 color;           // albedo texture (white for a ThinDielectric without one); conductor: F0 tint
    float rough;        // Diffuse: max(texture.g, 1e-4), default 1 (Material.hpp:204,272-275); Dielectric: texture.g, default 0
    float metallic;     // Diffuse: texture.b, default 0 (Material.hpp:276-278)
};

PT_HDN void mat_fetch(const ShadeScene& S, const SurfInt& si, MatCtx& m){
    const DevMaterial& M = S.materials[si.mat];
    m.kind = M.kind; m.ri = M.ri; m.rough = 0.0f; m.metallic = 0.0f;
    if(M.kind == PT_MAT_CONDUCTOR){ m.color = ld3(M.albedo); return; }
    m.color = M.tex >= 0 ? tex_eval(S, M.tex, si.u, si.v, si.ns) : v3(1.0f);
    if(M.kind == PT_MAT_THIN_DIELECTRIC)return;
    // roughness in .g, metallic in .b (Material.hpp:272-278): one lookup when both name the same texture (the usual packing), the value is the same
    const V3 orm = M.rough >= 0 ? tex_eval(S, M.rough, si.u, si.v, si.ns) : v3(0.0f);
    const float g = M.rough >= 0 ? orm.y : (M.kind == PT_MAT_DIFFUSE ? 1.0f : 0.0f);
    if(M.kind == PT_MAT_DIFFUSE){
        m.rough = fmaxf(g, 0.0001f);
        m.metallic = M.metal >= 0 ? (M.metal == M.rough ? orm.z : tex_eval(S, M.metal, si.u, si.v, si.ns).z) : 0.0f;
    } else m.rough = g;
}

PT_HDI float diff_prob(float r){ return r >= 0.7f ? 1.0f : 0.5f; }

// shared tail of MicrofacetDiffuse::scatter / calc_attenuation (Material.hpp:241-262, 309-323)
PT_HDI bool diff_f(const MatCtx& M, Ggx g, V3 wo, V3 wi, V3 wh, V3& f){
    const V3 color = M.color;
    const float metallic = M.metallic;
    const V3 F0 = v3(0.04f) * (1.0f - metallic) + color * metallic;       // glm::mix
    const V3 F = fresnel_schlick(dot(wi, wh), F0);
    const V3 num = (ggx_D(g, wh) * ggx_G(g, wo, wi)) * F;
    const float den = fabsf(4.0f * wo.z * wi.z);
    if(den == 0.0f)return false;
    const V3 spec = num / den;
    const V3 kD = (v3(1.0f) - F) * (1.0f - metallic);
    f = kD * color * kInvPi + spec;
    return true;
}

__host__ __device__ inline bool diffuse_scatter(const MatCtx& M, const SurfInt& si, V3 rd, float u, float ux, float uy, Scatter& out){
    const float rough = M.rough;
    const Onb tbn = onb_from_normal(dot(rd, si.ns) > 0.0f ? -si.ns : si.ns);
    const Ggx g = { rough * rough };
    const float prob = diff_prob(rough);
    const V3 wo = to_local(tbn, -rd);
    V3 wi, wh;
    if(u >= prob){
        wh = ggx_sample_wh(g, wo, ux, uy);
        wi = reflect3(-wo, wh);
    } else{
        const float z = sqrtf(1.0f - uy);
        const float phi = 2.0f * kPi * ux;
        const float s2 = sqrtf(uy);
        wi = v3(cosf(phi) * s2, sinf(phi) * s2, z);
        wh = normalize(wo + wi);
    }
    if(wi.z <= 0.0f)return false;
    const float dpdf = prob * wi.z * kInvPi;
    const float spdf = (1.0f - prob) * ggx_pdf(g, wo, wh) / (4.0f * fabsf(dot(wo, wh)));
    out.pdf = dpdf + spdf;
    if(!diff_f(M, g, wo, wi, wh, out.f))return false;
    out.o = si.p; out.d = to_world(tbn, wi); out.flags = 0;
    return true;
}

__host__ __device__ inline float diffuse_pdf(const MatCtx& M, const SurfInt& si, V3 in, V3 outd){
    const float rough = M.rough;
    const Ggx g = { rough * rough };
    const Onb tbn = onb_from_normal(dot(in, si.ns) > 0.0f ? -si.ns : si.ns);
    const V3 wo = to_local(tbn, -in);
    const V3 wh = to_local(tbn, normalize(outd - in));
    const float prob = diff_prob(rough);
    const float d = prob * fabsf(dot(si.ns, outd)) * kInvPi;
    const float sp = ggx_pdf(g, wo, wh) / (4.0f * fabsf(dot(wo, wh)));   // no (1 - prob): Material.hpp:293
    return d + sp;
}

__host__ __device__ inline V3 diffuse_eval(const MatCtx& M, const SurfInt& si, V3 in, V3 outd){
    const Onb tbn = onb_from_normal(dot(in, si.ns) > 0.0f ? -si.ns : si.ns);
    const V3 wo = to_local(tbn, -in), wi = to_local(tbn, outd);
    const V3 wh = normalize(wo + wi);
    const float rough = M.rough;
    const Ggx g = { rough * rough };
    V3 f;
    if(!diff_f(M, g, wo, wi, wh, f))return v3(0.0f);
    return f;
}

// ---- MicrofacetDielectric (Material.hpp:363-598) -------------------------------------------
__host__ __device__ inline bool dielectric_scatter(const MatCtx& M, const SurfInt& si, V3 ro, V3 rd, float u, float ux, float uy, Scatter& out){
    const float rough = M.rough;
    const Ggx g = { rough * rough };
    const Onb tbn = onb_from_frame(si.ns, si.tangent);
    const V3 wo = to_local(tbn, -rd);
    const float eta = dot(-rd, si.ns) > 0.0f ? 1.0f / M.ri : M.ri;
    const V3 Ng = dot(rd, si.n) > 0.0f ? -si.n : si.n;
    const V3 at = ro + si.t * rd;
    const V3 color = M.color;
    if(M.ri == 1.0f || g.a < 1e-6f){
        const V3 N = dot(rd, si.ns) > 0.0f ? -si.ns : si.ns;
        const float R = fresnel_dielectric(wo.z, M.ri), T = 1.0f - R;
        if(u < R / (R + T)){
            out.d = to_world(tbn, v3(-wo.x, -wo.y, wo.z));
            out.o = at + kEps * Ng;
            out.f = color * R / fabsf(dot(si.ns, out.d));
            out.pdf = R / (R + T);
        } else{
            out.d = refract3(rd, N, eta);
            if(is_zero(out.d))return false;
            out.o = at - kEps * Ng;
            out.f = color * T / fabsf(dot(si.ns, out.d));
            out.pdf = T / (R + T);
        }
        out.flags = BX_TRANSMISSIVE | BX_SPECULAR;
        return true;
    }
    const V3 wh = ggx_sample_wh(g, wo, ux, uy);
    const float R = fresnel_dielectric(dot(wo, wh), 1.0f / eta), T = 1.0f - R;
    out.flags = BX_TRANSMISSIVE | (rough < 0.001f ? BX_SPECULAR : 0u);
    if(u < R / (R + T)){
        const V3 wi = reflect3(-wo, wh);
        if(wo.z * wi.z < 0.0f)return false;
        out.o = at + kEps * Ng; out.d = to_world(tbn, wi);
        out.pdf = ggx_pdf(g, wo, wh) / (4.0f * fabsf(dot(wo, wh))) * R / (R + T);
        out.f = color * (ggx_D(g, wh) * ggx_G(g, wo, wi) * R / fabsf(4.0f * wi.z * wo.z));
        return true;
    }
    const V3 wi = refract3(-wo, wh, eta);
    if(wo.z * wi.z > 0.0f || wi.z == 0.0f)return false;
    out.o = at - kEps * Ng; out.d = to_world(tbn, wi);
    const float s = dot(wi, wh) + dot(wo, wh) * eta;
    const float denom = s * s;
    const float dwh = fabsf(dot(wi, wh)) / denom;
    out.pdf = ggx_pdf(g, wo, wh) * dwh * T / (R + T);
    const float ft = T * ggx_D(g, wh) * ggx_G(g, wo, wi) * fabsf(dot(wi, wh) * dot(wo, wh) / (denom * wi.z * wo.z));
    out.f = color * ft;
    return true;
}

// common front of MicrofacetDielectric::PDF / calc_attenuation (Material.hpp:497-523, 540-566)
__host__ __device__ inline bool diel_half(const MatCtx& M, V3 wo, V3 wi, V3& wh, bool& reflect, float& etap){
    const float co = wo.z, ci = wi.z;
    reflect = ci * co > 0.0f;
    etap = 1.0f;
    if(!reflect)etap = co > 0.0f ? M.ri : 1.0f / M.ri;
    wh = wi * etap + wo;
    if(dot(wh, wh) == 0.0f)return false;
    wh = normalize(wh);
    if(wh.z < 0.0f)wh = -wh;
    if(dot(wh, wi) * ci <= 0.0f || dot(wh, wo) * co <= 0.0f)return false;
    return true;
}

__host__ __device__ inline float dielectric_pdf(const MatCtx& M, const SurfInt& si, V3 in, V3 outd){
    const float rough = M.rough;
    const Ggx g = { rough * rough };
    if(M.ri == 1.0f || g.a < 1e-6f)return 0.0f;
    const Onb tbn = onb_from_frame(si.ns, si.tangent);
    const V3 wo = to_local(tbn, -in), wi = to_local(tbn, outd);
    V3 wh; bool reflect; float etap;
    if(!diel_half(M, wo, wi, wh, reflect, etap))return 0.0f;
    const float R = fresnel_dielectric(dot(wo, wh), M.ri), T = 1.0f - R;
    const float pdf = ggx_pdf(g, wo, wh);
    if(reflect)return pdf / (4.0f * fabsf(dot(wo, wh))) * R / (R + T);
    const float s = dot(wi, wh) + dot(wo, wh) / etap;
    return pdf * (fabsf(dot(wi, wh)) / (s * s)) * T / (R + T);
}

__host__ __device__ inline V3 dielectric_eval(const MatCtx& M, const SurfInt& si, V3 in, V3 outd){
    const float rough = M.rough;
    const Ggx g = { rough * rough };
    if(M.ri == 1.0f || g.a < 1e-6f)return v3(0.0f);
    const Onb tbn = onb_from_frame(si.ns, si.tangent);
    const V3 wo = to_local(tbn, -in), wi = to_local(tbn, outd);
    V3 wh; bool reflect; float etap;
    if(!diel_half(M, wo, wi, wh, reflect, etap))return v3(0.0f);
    const float F = fresnel_dielectric(dot(wo, wh), M.ri);
    const V3 color = M.color;
    if(reflect)return color * (ggx_D(g, wh) * ggx_G(g, wo, wi) * F / fabsf(4.0f * wi.z * wo.z));
    const float s = dot(wi, wh) + dot(wo, wh) / etap;
    const float denom = s * s * wi.z * wo.z;
    const float ft = ggx_D(g, wh) * (1.0f - F) * ggx_G(g, wo, wi) * fabsf(dot(wi, wh) * dot(wo, wh) / denom);
    return color * ft;
}

// ---- ThinDielectric (Material.hpp:601-657), SpecularConductor (:660-673) -------------------
__host__ __device__ inline bool thin_scatter(const MatCtx& M, const SurfInt& si, V3 ro, V3 rd, float u, Scatter& out){
    const Onb tbn = onb_from_frame(si.ns, si.tangent);
    const V3 wo = to_local(tbn, -rd);
    const V3 Ng = dot(rd, si.n) > 0.0f ? -si.n : si.n;
    float R = fresnel_dielectric(wo.z, M.ri), T = 1.0f - R;
    if(R < 1.0f){ R += T * T * R / (1.0f - R * R); T = 1.0f - R; }
    const V3 at = ro + si.t * rd;
    if(u < R / (R + T)){
        out.d = to_world(tbn, v3(-wo.x, -wo.y, wo.z));
        out.o = at + kEps * Ng;
        out.f = v3(R / fabsf(dot(si.ns, out.d)));
        out.pdf = R / (R + T);
    } else{
        out.d = rd;
        out.o = at - kEps * Ng;
        out.f = v3(T / fabsf(dot(si.ns, out.d)));
        out.pdf = T / (R + T);
    }
    out.f *= M.color;
    out.flags = BX_TRANSMISSIVE | BX_SPECULAR;
    return true;
}

__host__ __device__ inline bool conductor_scatter(const MatCtx& M, const SurfInt& si, V3 rd, Scatter& out){
    out.o = si.p; out.d = reflect3(rd, si.ns);
    const float d = dot(out.d, si.ns);
    if(d <= 0.0f)return false;
    out.f = fresnel_schlick(dot(si.ns, -rd), M.color) / d;
    out.pdf = 1.0f; out.flags = BX_SPECULAR;
    return true;
}

// ---- dispatch (the reference's virtual calls) ------------------------------------------------
PT_HDN bool mat_scatter(const MatCtx& M, const SurfInt& si, V3 ro, V3 rd, float u, float ux, float uy, Scatter& out){
    switch(M.kind){
    case PT_MAT_DIFFUSE: return diffuse_scatter(M, si, rd, u, ux, uy, out);
    case PT_MAT_DIELECTRIC: return dielectric_scatter(M, si, ro, rd, u, ux, uy, out);
    case PT_MAT_THIN_DIELECTRIC: return thin_scatter(M, si, ro, rd, u, out);
    default: return conductor_scatter(M, si, rd, out);
    }
}
PT_HDN float mat_pdf(const MatCtx& M, const SurfInt& si, V3 in, V3 outd){
    if(M.kind == PT_MAT_DIFFUSE)return diffuse_pdf(M, si, in, outd);
    if(M.kind == PT_MAT_DIELECTRIC)return dielectric_pdf(M, si, in, outd);
    return 0.0f;                                                  // Material.hpp:157-159, 651-653
}
PT_HDN V3 mat_eval(const MatCtx& M, const SurfInt& si, V3 in, V3 outd){
    if(M.kind == PT_MAT_DIFFUSE)return diffuse_eval(M, si, in, outd);
    if(M.kind == PT_MAT_DIELECTRIC)return dielectric_eval(M, si, in, outd);
    if(M.kind == PT_MAT_THIN_DIELECTRIC)return v3(0.0f);
    return v3(1.0f);                                              // Material base class, Material.hpp:153-155
}

} // namespace ptb
