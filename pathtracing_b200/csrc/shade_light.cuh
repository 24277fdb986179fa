// shade_light.cuh — LightSampler::{Sample,PMF} (LightSampler.cpp), Light::{sample,
// PDF,L} for every light class (Light.cpp) incl. TransformedLight, and the
// homogeneous medium + Henyey-Greenstein phase function (Medium.hpp,
// PhaseFunction.cpp).
#pragma once
#include "shade.cuh"

namespace ptb {

// UniformLightSampler::Sample (LightSampler.cpp:7-11) / PowerLightSampler::Sample (:34-46)
__host__ __device__ inline int pick_light(const ShadeScene& S, float u){
    const int n = S.n_sampler_lights;
    if(n == 0)return -1;
    if(S.sampler_kind == 0){
        int i = (int)(u * n);
        if(i > n - 1)i = n - 1;
        return S.sampler_lights[i];
    }
    // first i with cdf[i] >= u * totalPower, else the last light; the cdf is the reference's running float sum
    const float power = u * S.total_power;
    int lo = 0, hi = n;
    while(lo < hi){
        const int mid = (lo + hi) >> 1;
        if(S.sampler_cdf[mid] >= power)hi = mid; else lo = mid + 1;
    }
    if(lo >= n)lo = n - 1;
    return S.sampler_lights[lo];
}

// LightSample (Light.hpp:6-14) with what SampleLd needs from it
struct LightSmp {
    V3 L;          // radiance carried by the sample (delta lights only)
    V3 p, n;       // sampled point / normal; n == 0 marks a direction sample (isDeltaInteraction)
    V3 dir;
    float u, v;    // texcoords of the sampled point
};

__host__ __device__ inline void tri_vertices(const ShadeScene& S, const DevLight& L, V3& v0, V3& v1, V3& v2, uint32_t& i0, uint32_t& i1, uint32_t& i2){
    const DevMesh& m = S.meshes[L.mesh];
    i0 = m.vtx_base + S.idx[m.idx_base + 3 * L.tri]; i1 = m.vtx_base + S.idx[m.idx_base + 3 * L.tri + 1]; i2 = m.vtx_base + S.idx[m.idx_base + 3 * L.tri + 2];
    v0 = ld3(S.pos + 3 * (size_t)i0); v1 = ld3(S.pos + 3 * (size_t)i1); v2 = ld3(S.pos + 3 * (size_t)i2);
}

// TransformedLight::sample (Light.cpp:306-312) of the light on instance I, then AnimatedLight::sample (:345-348)
PT_HDI void light_to_world(const DevInstShade& I, float time, LightSmp& o){
    if(I.has_xf){
        o.p = xf_point(I.fwd, o.p);
        o.n = m3_mul(I.nrm, o.n);
    }
    if(I.animated)o.p = o.p + anim_offset(I, time);
}

// Light::sample(uv, time).  `rng` feeds TextureInfiniteLight::sample's extra random_float (Light.cpp:120).
PT_HDN void light_sample(const ShadeScene& S, const DevLight& L, float ux, float uy, float time, Pcg& rng, LightSmp& o){
    o.L = v3(0.0f); o.p = v3(0.0f); o.n = v3(0.0f); o.dir = v3(0.0f); o.u = 0.0f; o.v = 0.0f;
    switch(L.kind){
    case DL_UNIFORM_INF: o.dir = sphere_dir(ux, uy); o.L = ld3(L.color); break;                  // Light.cpp:35-42
    case DL_GRADIENT_INF: o.dir = sphere_dir(ux, uy); o.L = inf_Le(S, L, o.dir); break;          // Light.cpp:61-68
    case DL_TEXTURE_INF: {                                                                       // Light.cpp:118-144
        const float weight = rnd_float(rng) * L.total_weight;
        const int n = S.env_w * S.env_h;
        int lo = 0, hi = n;                      // std::upper_bound: first cdf > weight
        while(lo < hi){ const int mid = (lo + hi) >> 1; if(S.env_cdf[mid] > weight)hi = mid; else lo = mid + 1; }
        const int cx = lo % S.env_h, cy = lo / S.env_h;     // the reference decodes both with ySamples
        o.dir = sphere_dir((cx + ux) / (float)S.env_w, (cy + uy) / (float)S.env_h);
        break;
    }
    case DL_DISTANT: o.dir = normalize(ld3(L.v) + sphere_dir(ux, uy) * 0.02f); o.L = ld3(L.color); o.u = ux; o.v = uy; break;   // Light.cpp:208-215
    case DL_POINT: o.p = ld3(L.v); o.n = v3(1.0f); o.L = ld3(L.color); o.u = ux; o.v = uy; break;                             // Light.cpp:235-237
    case DL_SPHERE: {                                                                            // SphereShape::Sample, Shape.cpp:73-81
        const float* g = S.inst[L.inst].shp;
        const float z = 1.0f - 2.0f * ux;
        const float r = sqrtf(1.0f - z * z);
        const float phi = 2.0f * kPi * uy;
        const V3 c = ld3(g);
        o.p = c + g[3] * v3(r * cosf(phi), r * sinf(phi), z);
        o.n = normalize(o.p - c);
        sphere_uv(o.p, o.u, o.v);                      // GetSphereUV of the point, not of the normal
        light_to_world(S.inst[L.inst], time, o);
        break;
    }
    case DL_QUAD: {                                                                              // QuadShape::Sample, Shape.hpp:139-141: uv stays (0,0)
        const float* g = S.inst[L.inst].shp;
        o.p = ld3(g) + ux * ld3(g + 3) + uy * ld3(g + 6);
        o.n = ld3(g + 9);
        light_to_world(S.inst[L.inst], time, o);
        break;
    }
    default: {
        // AreaLight::sample -> TriangleShape::Sample (Shape.cpp:277-297): no fold of u+v > 1
        V3 v0, v1, v2; uint32_t i0, i1, i2;
        tri_vertices(S, L, v0, v1, v2, i0, i1, i2);
        const float w = 1.0f - ux - uy;
        V3 n = normalize(cross(v1 - v0, v2 - v0));
        if(n.x != n.x)n = v3(0.0f);
        o.p = ux * v1 + uy * v2 + w * v0; o.n = n;
        o.u = ux * S.uv[2 * (size_t)i1] + uy * S.uv[2 * (size_t)i2] + w * S.uv[2 * (size_t)i0];
        o.v = ux * S.uv[2 * (size_t)i1 + 1] + uy * S.uv[2 * (size_t)i2 + 1] + w * S.uv[2 * (size_t)i0 + 1];
        light_to_world(S.inst[L.inst], time, o);
        break;
    }
    }
}

PT_HDI bool light_is_delta(const DevLight& L){ return L.kind == DL_DISTANT || L.kind == DL_POINT; }

// TriangleShape::PDF(interaction, ray) (Shape.cpp:308-315)
PT_HDI float tri_pdf(V3 p, V3 n, V3 ro, V3 rd, float area){
    const V3 to = p - ro;
    const float d2 = dot(to, to);
    const float c = fabsf(dot(-rd, n));
    if(area == 0.0f || c == 0.0f || n.x != n.x)return 0.0f;
    return d2 / (c * area);
}

// AreaLight::PDF(interaction, ray) (Light.cpp:267-272) for the light on a hit or sampled point, world space:
// TriangleShape::PDF (Shape.cpp:308-315), SphereShape::PDF (Shape.cpp:64-71), QuadShape::PDF (Shape.hpp:151-158)
PT_HDI float area_pdf(const DevLight& L, V3 p, V3 n, V3 ro, V3 rd){
    if(L.kind == DL_AREA)return tri_pdf(p, n, ro, rd, L.area);
    if(L.one_sided && !(dot(-rd, n) > 0.0f))return 0.0f;
    const V3 to = p - ro;
    const float d2 = dot(to, to);
    const float c = fabsf(dot(-rd, n));
    if(L.kind == DL_SPHERE ? (L.area * c == 0.0f) : (L.area == 0.0f))return 0.0f;
    return d2 / (c * L.area);
}

// AreaLight::L (Light.cpp:257-260)
PT_HDI V3 area_L(const ShadeScene& S, const DevLight& L, V3 n, V3 ns, float u, float v, V3 rd){
    if(L.one_sided && dot(rd, n) > 0.0f)return v3(0.0f);
    return tex_eval(S, L.tex, u, v, ns);
}

// Light::PDF(interaction, ray) of the light object the sampler holds: AnimatedLight::PDF (Light.cpp:353-356) ->
// TransformedLight::PDF (:318-325) -> AreaLight::PDF (:267-272), each wrapper taking p, n and the ray into its
// child's space
PT_HDN float light_pdf_sampled(const ShadeScene& S, const DevLight& L, V3 p, V3 n, V3 ro, V3 rd, float time){
    switch(L.kind){
    case DL_UNIFORM_INF: case DL_GRADIENT_INF: return 1.0f / (4.0f * kPi);
    case DL_TEXTURE_INF: {                                                                       // Light.cpp:146-150
        const float cell = 4.0f * kPi / (float)(S.env_w * S.env_h);
        return (lum(inf_Le(S, L, rd)) / L.total_weight) * (1.0f / cell);
    }
    case DL_DISTANT: case DL_POINT: return 0.0f;
    default: {
        const DevInstShade& I = S.inst[L.inst];
        if(I.animated){
            const V3 off = anim_offset(I, time);
            p = p - off; n = normalize(n); ro = ro - off; rd = normalize(rd);
        }
        if(I.has_xf){
            p = xf_point(I.inv, p); n = normalize(xf_vec(I.inv, n));
            ro = xf_point(I.inv, ro); rd = normalize(xf_vec(I.inv, rd));
        }
        return L.kind == DL_AREA ? tri_pdf(p, n, ro, rd, L.area) : area_pdf(L, p, n, ro, rd);
    }
    }
}

// Light::L(lightSample.interaction, shadow_ray) as SampleLd calls it on the sampled light
__host__ __device__ inline V3 light_L_sampled(const ShadeScene& S, const DevLight& L, const LightSmp& ls, V3 rd){
    if(L.kind <= DL_TEXTURE_INF)return inf_Le(S, L, rd);
    if(L.kind != DL_AREA && L.kind < DL_SPHERE)return v3(0.0f);
    const DevInstShade& I = S.inst[L.inst];
    // TransformedLight::L / AnimatedLight::L build a fresh SurfaceInteraction (Light.cpp:300-305, 341-344): its
    // uv is (0,0) and its normal the wrapper's normal matrix applied to the (already world-space) sample normal
    const bool wrapped = I.has_xf || I.animated;
    const V3 n = I.has_xf ? m3_mul(I.nrm, ls.n) : ls.n;
    if(L.kind >= DL_SPHERE)return area_L(S, L, n, v3(0.0f), wrapped ? 0.0f : ls.u, wrapped ? 0.0f : ls.v, rd);
    return tex_eval(S, L.tex, wrapped ? 0.0f : ls.u, wrapped ? 0.0f : ls.v, v3(0.0f));
}

// ---- media (Medium.hpp:14-61, PhaseFunction.cpp:4-26) ---------------------------------------
PT_HDI V3 medium_Tr(const DevMedium& M, float t){
    const float tt = fminf(t, 3.402823466e+38f);
    return v3(expf(-M.sigma_t[0] * tt), expf(-M.sigma_t[1] * tt), expf(-M.sigma_t[2] * tt));
}
// HomogeneusMedium::Sample: returns the throughput factor; sampled = a medium event before t, at `dist`
PT_HDI V3 medium_sample(const DevMedium& M, float t, Pcg& rng, bool& sampled, float& dist){
    int ch = (int)(3.0f * rnd_float(rng));
    if(ch > 2)ch = 2;
    dist = fminf((float)(-log(1.0 - (double)rnd_float(rng))) / M.sigma_t[ch], t);
    sampled = dist < t;
    const V3 tr = medium_Tr(M, dist);
    const V3 st = ld3(M.sigma_t);
    const V3 dens = sampled ? st * tr : tr;
    const float pdf = (dens.x + dens.y + dens.z) / 3.0f;
    return sampled ? (tr * ld3(M.sigma_s) / pdf) : (tr / pdf);
}
PT_HDI float phase_hg(float c, float g){
    const float denom = 1.0f + g * g + 2.0f * g * c;
    return 0.25f * kInvPi * (1.0f - g * g) / (denom * sqrtf(denom));
}
PT_HDI V3 hg_sample(float g, V3 in, float u0, float u1){
    float c;
    if(fabsf(g) < 1e-3f)c = 1.0f - 2.0f * u0;
    else{
        const float sqr = (1.0f - g * g) / (1.0f - g + 2.0f * g * u0);
        c = (1.0f + g * g - sqr * sqr) / (2.0f * g);
    }
    const float s = sqrtf(fmaxf(0.0f, 1.0f - c * c));
    const float phi = 2.0f * kPi * u1;
    return normalize(to_world(onb_from_normal(in), v3(s * cosf(phi), s * sinf(phi), c)));
}

} // namespace ptb
