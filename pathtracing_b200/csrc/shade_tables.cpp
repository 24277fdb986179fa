// shade_tables.cpp — builds the shading tables on the host: vertex data, per-triangle mesh/emission
// lookups, instance transforms, textures, materials, media, and the light list with the reference's
// pre-processing (AreaLight::PreProcess Light.cpp:277-287, InfiniteLight power Light.cpp:47-108,
// 166-200, PowerLightSampler::PreProcess LightSampler.cpp:52-64).
#include <cfloat>
#include <cmath>
#include <cstring>
#include <cstdlib>
#include <cstdio>
#include <chrono>
#include <thread>

#include "shade_tables.hpp"
#include "shade_light.cuh"

namespace ptb {

namespace {

float power_func(int k, float r){
    switch(k){
    case 1: return 4.0f * r;
    case 2: return 4.0f * std::sqrt(r);
    default: return std::sqrt(r);
    }
}

// glm::determinant(mat4), column-major
float mat4_det(const float m[16]){
#define M(c, r) m[(c) * 4 + (r)]
    const float s0 = M(2, 2) * M(3, 3) - M(3, 2) * M(2, 3);
    const float s1 = M(2, 1) * M(3, 3) - M(3, 1) * M(2, 3);
    const float s2 = M(2, 1) * M(3, 2) - M(3, 1) * M(2, 2);
    const float s3 = M(2, 0) * M(3, 3) - M(3, 0) * M(2, 3);
    const float s4 = M(2, 0) * M(3, 2) - M(3, 0) * M(2, 2);
    const float s5 = M(2, 0) * M(3, 1) - M(3, 0) * M(2, 1);
    const float c0 = +(M(1, 1) * s0 - M(1, 2) * s1 + M(1, 3) * s2);
    const float c1 = -(M(1, 0) * s0 - M(1, 2) * s3 + M(1, 3) * s4);
    const float c2 = +(M(1, 0) * s1 - M(1, 1) * s3 + M(1, 3) * s5);
    const float c3 = -(M(1, 0) * s2 - M(1, 1) * s4 + M(1, 2) * s5);
    return M(0, 0) * c0 + M(0, 1) * c1 + M(0, 2) * c2 + M(0, 3) * c3;
#undef M
}

} // namespace

static void build_geometry(pt_scene* s, ShadeTables& T){
    const bool dev = T.geometry_on_device;
    if(!dev){   // one allocation per table (a 10 M-triangle scene spent most of its 0.9 s here growing the vectors)
        size_t nv = 0, ni = 0, nt = 0;
        for(const HostMesh& m : s->meshes){ nv += m.pos.size() / 3; ni += m.idx.size(); }
        for(const HostBlas& hb : s->blas)nt += hb.n_tris;
        T.pos.reserve(3 * nv); T.nrm.reserve(3 * nv); T.uv.reserve(2 * nv); T.tan.reserve(3 * nv); T.idx.reserve(ni);
        T.stri.reserve(nt); T.tri_emis.reserve(nt);
    }
    for(const HostMesh& m : s->meshes){
        DevMesh d {};
        d.vtx_base = (uint32_t)T.n_vertices;
        d.idx_base = (uint32_t)T.n_indices;
        d.prim_base = 0;
        d.has_tangents = m.tan.empty() ? 0u : 1u;
        d.material = m.d.material; d.emissive_tex = m.d.emissive_tex; d.medium = m.d.medium;
        T.n_vertices += m.pos.size() / 3; T.n_indices += m.idx.size();
        if(!dev){
            T.pos.insert(T.pos.end(), m.pos.begin(), m.pos.end());
            T.nrm.insert(T.nrm.end(), m.nrm.begin(), m.nrm.end());
            T.uv.insert(T.uv.end(), m.uv.begin(), m.uv.end());
            if(m.tan.empty())T.tan.insert(T.tan.end(), m.pos.size(), 0.0f);
            else T.tan.insert(T.tan.end(), m.tan.begin(), m.tan.end());
            T.idx.insert(T.idx.end(), m.idx.begin(), m.idx.end());
        }
        T.meshes.push_back(d);
    }
    // per-BLAS triangle tables in original (BLAS-wide) order.  A mesh used by two BLASes keeps
    // the prim_base of the last one; shading resolves BLAS-wide ids through the ShadeTri table, which needs no prim_base.
    for(const HostBlas& hb : s->blas){
        DevBlasInfo bi {};
        bi.tri_base = (uint32_t)T.n_blas_tris;
        bi.n_tris = hb.n_tris;
        int32_t emis = 0;
        for(size_t k = 0; k < hb.meshes.size(); k++){
            const int mi = hb.meshes[k];
            const HostMesh& m = s->meshes[mi];
            T.meshes[mi].prim_base = hb.tri_mesh_base[k];
            T.tri_runs.push_back({ mi, (uint32_t)T.n_blas_tris, m.d.n_triangles, m.d.emissive_tex >= 0 ? emis : -1 });
            T.n_blas_tris += m.d.n_triangles;
            if(dev){ if(m.d.emissive_tex >= 0)emis += (int32_t)m.d.n_triangles; continue; }
            {
                const uint32_t vb = T.meshes[mi].vtx_base;
                const size_t at = T.stri.size();
                T.stri.resize(at + m.d.n_triangles);
                for(uint32_t t = 0; t < m.d.n_triangles; t++)T.stri[at + t] = { vb + m.idx[3 * (size_t)t], vb + m.idx[3 * (size_t)t + 1], vb + m.idx[3 * (size_t)t + 2], (uint32_t)mi };
            }
            if(m.d.emissive_tex >= 0)for(uint32_t t = 0; t < m.d.n_triangles; t++)T.tri_emis.push_back(emis++);
            else T.tri_emis.insert(T.tri_emis.end(), m.d.n_triangles, -1);
        }
        T.blas.push_back(bi);
    }
    for(const HostInstance& hi : s->instances){
        DevInstShade d {};
        d.shape = -1; d.material = -1; d.medium = -1; d.light_base = -1;
        d.animated = hi.animated ? 1u : 0u;
        for(int k = 0; k < 3; k++)d.adir[k] = hi.adir[k];
        d.atb[0] = hi.atb[0]; d.atb[1] = hi.atb[1];
        float inv[16], nm[9];
        mat4_inverse(hi.xf, inv);
        mat3_normal_matrix(hi.xf, nm);
        for(int c = 0; c < 4; c++)for(int r = 0; r < 3; r++){ d.fwd[c * 3 + r] = hi.xf[c * 4 + r]; d.inv[c * 3 + r] = inv[c * 4 + r]; }
        for(int k = 0; k < 9; k++)d.nrm[k] = nm[k];
        d.det = mat4_det(hi.xf);
        d.has_xf = hi.has_xf ? 1u : 0u;
        if(hi.blas < 0){
            const pt_shape_desc& sd = hi.shape;
            d.shape = sd.kind; d.material = sd.material; d.medium = sd.medium;
            for(int k = 0; k < 3; k++)d.shp[k] = sd.p[k];
            if(sd.kind == PT_SHAPE_SPHERE)d.shp[3] = sd.radius;
            else{
                const V3 u = ld3(sd.u), v = ld3(sd.v);
                const V3 n = normalize(cross(u, v));                      // QuadShape ctor, Shape.hpp:121-122
                for(int k = 0; k < 3; k++){ d.shp[3 + k] = sd.u[k]; d.shp[6 + k] = sd.v[k]; }
                d.shp[9] = n.x; d.shp[10] = n.y; d.shp[11] = n.z;
            }
        } else{
            d.blas = (uint32_t)hi.blas;
            d.tri_base = T.blas[hi.blas].tri_base;
        }
        T.inst.push_back(d);
    }
}

static void build_appearance(pt_scene* s, ShadeTables& T){
    for(const HostTexture& ht : s->textures){
        DevTexture d {};
        d.kind = ht.d.kind;
        for(int k = 0; k < 3; k++){ d.color[k] = ht.d.color[k]; d.scale[k] = ht.d.color_scale[k]; }
        d.w = ht.d.width; d.h = ht.d.height; d.ch = ht.d.channels;
        d.tex_a = ht.d.tex_a; d.tex_b = ht.d.tex_b;
        d.inv_scale[0] = 1.0f / ht.d.uv_scale[0]; d.inv_scale[1] = 1.0f / ht.d.uv_scale[1];
        while(T.texels.size() % 16)T.texels.push_back(0);
        d.texels = T.texels.size();
        T.texels.insert(T.texels.end(), ht.pixels.begin(), ht.pixels.end());
        T.textures.push_back(d);
    }
    for(const pt_material_desc& md : s->materials){
        DevMaterial d {};
        d.kind = md.kind; d.tex = md.tex; d.norm = md.norm; d.rough = md.rough; d.metal = md.metal; d.alpha = md.alpha;
        d.ri = md.ri;
        for(int k = 0; k < 3; k++)d.albedo[k] = md.albedo[k];
        d.alpha_mode = md.alpha_mode; d.alpha_cutoff = md.alpha_cutoff;
        T.materials.push_back(d);
    }
    for(const pt_medium_desc& md : s->media){
        DevMedium d {};
        for(int k = 0; k < 3; k++){
            d.sigma_a[k] = md.density * md.sigma_a[k];
            d.sigma_s[k] = md.density * md.sigma_s[k];
            d.sigma_t[k] = md.density * (md.sigma_a[k] + md.sigma_s[k]);   // Medium.hpp:18-20
            d.Le[k] = md.Le[k] * md.le_density;
        }
        d.g = std::fmin(std::fmax(md.g, -0.99f), 0.99f);                   // PhaseFunction.hpp:19
        T.media.push_back(d);
    }
}

// ---- lights -------------------------------------------------------------------------------

// AreaLight::PreProcess (Light.cpp:277-287): 2 * Area * luminance(mean emissive over 16x16
// strata of TriangleShape::Sample, which does not fold u+v>1, Shape.cpp:277-297).  The strata are
// sampled at their centres here (the reference jitters them with an unseeded RNG).
static float area_light_power(const ShadeTables& T, const DevMesh& m, const HostMesh& hm, uint32_t tri, float& area_out){
    // vertex data straight from the mesh (mesh-local indices): the staged copies may not exist (ShadeTables::geometry_on_device)
    const uint32_t* ix = &hm.idx[3 * (size_t)tri];
    const V3 p0 = ld3(&hm.pos[3 * (size_t)ix[0]]), p1 = ld3(&hm.pos[3 * (size_t)ix[1]]), p2 = ld3(&hm.pos[3 * (size_t)ix[2]]);
    const float* t0 = &hm.uv[2 * (size_t)ix[0]]; const float* t1 = &hm.uv[2 * (size_t)ix[1]]; const float* t2 = &hm.uv[2 * (size_t)ix[2]];
    const float area = length(cross(p0 - p2, p1 - p2)) * 0.5f;           // TriangleShape::Area, Shape.cpp:298-301
    area_out = area;
    double acc[3] = { 0, 0, 0 };
    for(int j = 0; j < 16; j++)for(int i = 0; i < 16; i++){
        const float ux = (i + 0.5f) / 16.0f, uy = (j + 0.5f) / 16.0f, w = 1.0f - ux - uy;
        const float u = ux * t1[0] + uy * t2[0] + w * t0[0], v = ux * t1[1] + uy * t2[1] + w * t0[1];
        const V3 e = tex_eval(T.view, m.emissive_tex, u, v, v3(0.0f));
        acc[0] += e.x; acc[1] += e.y; acc[2] += e.z;
    }
    const double l = 0.2126 * (acc[0] / 256.0) + 0.7152 * (acc[1] / 256.0) + 0.0722 * (acc[2] / 256.0);
    return (float)(2.0 * area * l);
}

// AreaLight::PreProcess for a sphere / quad light: (oneSided ? 1 : 2) * Area * luminance(mean emissive)
static float shape_light_power(const ShadeTables& T, const DevLight& L, float& area_out){
    const float* g = T.inst[L.inst].shp;
    const float area = L.kind == DL_SPHERE ? 4.0f * kPi * g[3] * g[3]                     // Shape.cpp:58-60
                                           : length(cross(ld3(g + 3), ld3(g + 6)));       // Shape.hpp:143-145
    area_out = area;
    double acc[3] = { 0, 0, 0 };
    Pcg unused {};
    for(int j = 0; j < 16; j++)for(int i = 0; i < 16; i++){
        LightSmp ls;
        light_sample(T.view, L, (i + 0.5f) / 16.0f, (j + 0.5f) / 16.0f, 0.0f, unused, ls);   // texcoords are the inner shape's
        const V3 e = tex_eval(T.view, L.tex, ls.u, ls.v, v3(0.0f));
        acc[0] += e.x; acc[1] += e.y; acc[2] += e.z;
    }
    const double l = 0.2126 * (acc[0] / 256.0) + 0.7152 * (acc[1] / 256.0) + 0.0722 * (acc[2] / 256.0);
    return (float)((L.one_sided ? 1.0 : 2.0) * area * l);
}

static void build_env_table(const ShadeTables& T, DevLight& L, std::vector<float>& cdf, int xs, int ys){
    // TextureInfiniteLight::PreProcess (Light.cpp:166-200), including its cell decode
    // (x = k % ySamples, y = k / ySamples for both axes).  2x2 strata centres per cell
    // instead of 64 jittered samples.
    const int n = xs * ys;
    std::vector<float> w((size_t)n);
    unsigned nt = std::thread::hardware_concurrency(); if(nt < 1)nt = 1;
    std::vector<std::thread> pool;
    for(unsigned t = 0; t < nt; t++)pool.emplace_back([&, t]{
        for(int k = (int)t; k < n; k += (int)nt){
            const int y = k / ys, x = k % ys;
            double acc = 0;
            for(int sy = 0; sy < 2; sy++)for(int sx = 0; sx < 2; sx++){
                const float u = (x + (sx + 0.5f) * 0.5f) / (float)xs, v = (y + (sy + 0.5f) * 0.5f) / (float)ys;
                acc += lum(inf_Le(T.view, L, sphere_dir(u, v)));
            }
            w[k] = (float)(acc / 4.0);
        }
    });
    for(auto& th : pool)th.join();
    cdf.resize((size_t)n);
    float run = 0.0f;                      // std::partial_sum over vector<float>
    for(int k = 0; k < n; k++){ run += w[k]; cdf[k] = run; }
    L.total_weight = run;
}

static int build_lights(pt_scene* s, ShadeTables& T, const std::vector<uint32_t>& inst_order){
    const float* lo = s->bounds.lo; const float* hi = s->bounds.hi;
    const V3 c = (v3(hi[0], hi[1], hi[2]) + v3(lo[0], lo[1], lo[2])) * 0.5f;
    const float radius = length(v3(hi[0], hi[1], hi[2]) - c);    // InfiniteLight::PreProcess, Light.cpp:9-12
    std::vector<float> power, power_inner;

    // area lights: Scene::GetLights walks the TLAS leaves (BVH.hpp:69-81); every emissive triangle of
    // every instance is one light (Model::BuildBlas, Model.hpp:47-59; TransformedPrimitive::GetLights)
    for(uint32_t id : inst_order){
        const HostInstance& hinst = s->instances[id];
        if(hinst.blas < 0){
            // GeometricPrimitive::GetLights (Primitive.cpp:28-30): its AreaLight, if any
            if(hinst.shape.emissive_tex < 0)continue;
            DevLight L {};
            L.kind = hinst.shape.kind == PT_SHAPE_SPHERE ? DL_SPHERE : DL_QUAD;
            L.tex = hinst.shape.emissive_tex; L.inst = id; L.valid = 1u; L.one_sided = hinst.shape.one_sided ? 1u : 0u;
            T.view.inst = T.inst.data();
            const float p = shape_light_power(T, L, L.area);
            T.inst[id].light_base = (int32_t)T.lights.size();
            T.lights.push_back(L);
            // TransformedLight::Power = inner * det(transform) (Light.cpp:334-336); AnimatedLight::Power = inner (:362-364)
            power.push_back(hinst.has_xf ? p * T.inst[id].det : p); power_inner.push_back(p);
            continue;
        }
        const HostBlas& hb = s->blas[hinst.blas];
        bool any = false;
        for(int mi : hb.meshes)if(s->meshes[mi].d.emissive_tex >= 0)any = true;
        if(!any)continue;
        T.inst[id].light_base = (int32_t)T.lights.size();
        for(int mi : hb.meshes){
            const DevMesh& m = T.meshes[mi];
            if(m.emissive_tex < 0)continue;
            for(uint32_t t = 0; t < s->meshes[mi].d.n_triangles; t++){
                DevLight L {};
                L.kind = DL_AREA; L.tex = m.emissive_tex; L.inst = id; L.mesh = (uint32_t)mi; L.tri = t;
                const float p = area_light_power(T, m, s->meshes[mi], t, L.area);
                L.valid = p > FLT_EPSILON ? 1u : 0u;              // Model.hpp:53-56
                T.lights.push_back(L);
                power_inner.push_back(p);
                power.push_back(hinst.has_xf ? p * T.inst[id].det : p);   // TransformedLight::Power
            }
        }
    }
    // user lights
    for(const pt_light_desc& d : s->lights){
        DevLight L {};
        L.kind = d.kind; L.valid = 1; L.tex = d.tex; L.scale = d.scale;
        for(int k = 0; k < 3; k++){ L.color[k] = d.color[k]; L.color_b[k] = d.color_b[k]; L.v[k] = d.dir_or_pos[k]; }
        float p = 0;
        const float pf = power_func(d.power_func, radius);
        switch(d.kind){
        case PT_LIGHT_UNIFORM_INFINITE: case PT_LIGHT_DISTANT: case PT_LIGHT_POINT:
            p = (d.color[0] + d.color[1] + d.color[2]) * pf; break;
        case PT_LIGHT_GRADIENT_INFINITE: {
            // FunctionInfiniteLight::PreProcess (Light.cpp:78-108): mean luminance over the sphere
            double acc = 0;
            for(int y = 0; y < 100; y++)for(int x = 0; x < 100; x++)
                acc += lum(inf_Le(T.view, L, sphere_dir((x + 0.5f) / 100.0f, (y + 0.5f) / 100.0f)));
            p = (float)(acc / 10000.0) * pf;
            break;
        }
        case PT_LIGHT_TEXTURE_INFINITE: {
            if(!T.env_cdf.empty()){ set_error("only one TextureInfiniteLight per scene is supported"); return PT_ERR_INVALID; }
            build_env_table(T, L, T.env_cdf, 1920, 1080);
            p = (float)((double)L.total_weight / (1920.0 * 1080.0)) * pf;
            break;
        }
        default: break;
        }
        if(d.kind <= PT_LIGHT_TEXTURE_INFINITE)T.inf_lights.push_back((int32_t)T.lights.size());
        T.lights.push_back(L);
        power.push_back(p); power_inner.push_back(p);
    }
    // LightSampler::PreProcess: cull Power < 0.01, accumulate totalPower in float (LightSampler.cpp:16-26,52-64)
    float total = 0.0f;
    for(size_t i = 0; i < T.lights.size(); i++){
        if(!T.lights[i].valid || power[i] < 0.01f)continue;
        T.sampler_lights.push_back((int32_t)i);
        total += power[i];
    }
    float run = 0.0f;
    for(int32_t i : T.sampler_lights){ run += power[i]; T.sampler_cdf.push_back(run); }
    const size_t ns = T.sampler_lights.size();
    for(size_t i = 0; i < T.lights.size(); i++){
        DevLight& L = T.lights[i];
        if(s->light_sampler == 0){      // UniformLightSampler::PMF
            L.pmf_sample = L.pmf_hit = ns ? 1.0f / (float)ns : 0.0f;
        } else{                         // PowerLightSampler::PMF: light->Power() / totalPower (1 if totalPower == 0)
            L.pmf_sample = total == 0.0f ? 1.0f : power[i] / total;
            L.pmf_hit = total == 0.0f ? 1.0f : power_inner[i] / total;
        }
    }
    T.view.total_power = total;
    return PT_OK;
}

int build_shade_tables_host(pt_scene* s, const std::vector<uint32_t>& inst_order, ShadeTables& T){
    static const bool prof = getenv("PTB_PROFILE") != nullptr;
    const auto t0 = std::chrono::steady_clock::now();
    build_geometry(s, T);
    const auto t1 = std::chrono::steady_clock::now();
    build_appearance(s, T);
    const auto t2 = std::chrono::steady_clock::now();
    ShadeScene& V = T.view;
    V.meshes = T.meshes.data(); V.blas = T.blas.data(); V.inst = T.inst.data();
    V.pos = T.pos.data(); V.nrm = T.nrm.data(); V.uv = T.uv.data(); V.tan = T.tan.data(); V.idx = T.idx.data();
    V.stri = T.stri.data(); V.tri_emis = T.tri_emis.data();
    V.textures = T.textures.data(); V.texels = T.texels.data(); V.materials = T.materials.data(); V.media = T.media.data();
    V.env_w = 1920; V.env_h = 1080;
    int rc = build_lights(s, T, inst_order);
    if(rc != PT_OK)return rc;
    if(prof){
        auto ms = [](auto a, auto b){ return std::chrono::duration<double, std::milli>(b - a).count(); };
        fprintf(stderr, "[ptb] shade tables (host): geometry %.1f ms, appearance %.1f ms, lights %.1f ms\n", ms(t0, t1), ms(t1, t2), ms(t2, std::chrono::steady_clock::now()));
    }
    V.inst = T.inst.data();          // light_base was filled in
    V.lights = T.lights.data(); V.sampler_lights = T.sampler_lights.data(); V.sampler_cdf = T.sampler_cdf.data();
    V.inf_lights = T.inf_lights.data(); V.env_cdf = T.env_cdf.data();
    V.n_lights = (int32_t)T.lights.size(); V.n_sampler_lights = (int32_t)T.sampler_lights.size(); V.n_inf_lights = (int32_t)T.inf_lights.size();
    V.sampler_kind = s->light_sampler;
    V.scene_medium = s->scene_medium;
    return PT_OK;
}

} // namespace ptb
