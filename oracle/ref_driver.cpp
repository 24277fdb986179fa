// oracle/ref_driver.cpp — TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// A C-ABI harness around the UNMODIFIED reference sources of
// marko176/PathTracing, compiled in place from /root/reference by
// oracle/Makefile into oracle/_ref/libptref*.so.  It lets the tests and
// bench.py's cpu_baseline / --impl reference leg drive the reference's own
// CPU implementation of the hot path (BLAS/TLAS Intersect, IntersectPred,
// {Simple,,Vol}PathIntegrator::Li) with the same POD scene description the
// product's C-ABI (include/pt_b200.h) takes.
//
// No reference source is copied: everything below only *calls* reference
// classes.  The render harness mirrors TileIntegrator::Render
// (Integrators.cpp:23-129) with the adaptive loop disabled (exactly spp
// samples per pixel), because the stock Render() keeps no sample counter and
// its Film pixels are private.
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <unistd.h>

#include "Scene.hpp"
#include "Mesh.hpp"
#include "Sampler.hpp"      // brings Util.hpp (luminance) in before Film.hpp needs it
#include "Integrators.hpp"
#include "Camera.hpp"
#include "LightSampler.hpp"
#include "Film.hpp"

#include "../include/pt_b200.h"

namespace {

struct RefMesh {
    pt_mesh_desc d;
    std::vector<float> pos, nrm, uv, tan;
    std::vector<uint32_t> idx;
};

struct RefBlasDesc { std::vector<int> meshes; };
struct RefInstDesc {   // blas < 0: an analytic shape
    int blas; bool has_xf; glm::mat4 xf; pt_shape_desc shape;
    bool animated = false; glm::vec3 adir { 0 }; glm::vec2 atb { 0, 1 };   // AnimatedPrimitive around it
};

struct RefScene {
    std::vector<std::shared_ptr<Texture>> textures;
    std::vector<std::shared_ptr<Material>> materials;
    std::vector<std::shared_ptr<Medium>> media;
    std::vector<RefMesh> meshDescs;
    std::vector<RefBlasDesc> blasDescs;
    std::vector<RefInstDesc> instDescs;
    std::vector<pt_light_desc> lightDescs;
    int sceneMedium = -1;
    int tmpCounter = 0;

    // built at commit
    std::vector<std::shared_ptr<Mesh>> meshes;           // keep alive (meshList registry)
    std::vector<std::shared_ptr<Primitive>> blas;        // per blas desc
    std::vector<uint32_t> blasTris;                      // triangles per blas
    std::vector<std::shared_ptr<Primitive>> topLevel;    // per instance
    std::shared_ptr<Scene> scene;
    std::shared_ptr<LightSampler> lightSampler;
    std::vector<std::shared_ptr<Light>> allLights;       // what was handed to lightSampler->Add, in order
    int width = 0;
    bool idMode = false;
    bool hasShapes = false;
};

std::function<float(float)> powerFunc(int k){
    switch(k){
    case 1: return [](float r) -> float{ return 4 * r; };
    case 2: return [](float r) -> float{ return 4 * std::sqrt(r); };
    default: return [](float r) -> float{ return std::sqrt(r); };
    }
}

std::shared_ptr<Texture> texOrNull(RefScene* s, int id){
    return (id >= 0 && id < (int)s->textures.size()) ? s->textures[id] : nullptr;
}

template <typename BLAS>
std::shared_ptr<Primitive> buildBlas(RefScene* s, const RefBlasDesc& bd, uint32_t& triCount){
    // Model::BuildBlas recipe, Model.hpp:41-60
    std::vector<GeometricPrimitive> prims;
    uint32_t base = 0;
    for(int mi : bd.meshes){
        const RefMesh& md = s->meshDescs[mi];
        uint32_t n = md.d.n_triangles;
        std::vector<uint32_t> indices;
        std::vector<glm::vec3> vertices, normals, tangents;
        std::vector<glm::vec2> texCoords;
        if(s->idMode){
            // private vertices per triangle; texcoords encode the BLAS-wide
            // triangle id so that SurfaceInteraction::uv returns it
            indices.resize(3 * (size_t)n);
            vertices.resize(3 * (size_t)n);
            normals.resize(3 * (size_t)n);
            texCoords.resize(3 * (size_t)n);
            for(uint32_t t = 0;t < n;t++){
                uint32_t id = base + t;
                for(int k = 0;k < 3;k++){
                    uint32_t vi = md.idx[3 * (size_t)t + k];
                    size_t o = 3 * (size_t)t + k;
                    indices[o] = (uint32_t)o;
                    vertices[o] = glm::vec3(md.pos[3 * vi], md.pos[3 * vi + 1], md.pos[3 * vi + 2]);
                    normals[o] = glm::vec3(md.nrm[3 * vi], md.nrm[3 * vi + 1], md.nrm[3 * vi + 2]);
                    texCoords[o] = glm::vec2(float(id & 4095u), float(id >> 12));
                }
            }
        } else{
            indices = md.idx;
            size_t nv = md.d.n_vertices;
            vertices.resize(nv); normals.resize(nv); texCoords.resize(nv);
            for(size_t v = 0;v < nv;v++){
                vertices[v] = glm::vec3(md.pos[3 * v], md.pos[3 * v + 1], md.pos[3 * v + 2]);
                normals[v] = glm::vec3(md.nrm[3 * v], md.nrm[3 * v + 1], md.nrm[3 * v + 2]);
                texCoords[v] = glm::vec2(md.uv[2 * v], md.uv[2 * v + 1]);
            }
            if(!md.tan.empty()){
                tangents.resize(nv);
                for(size_t v = 0;v < nv;v++)tangents[v] = glm::vec3(md.tan[3 * v], md.tan[3 * v + 1], md.tan[3 * v + 2]);
            }
        }
        std::shared_ptr<Material> mat = md.d.material >= 0 ? s->materials[md.d.material] : nullptr;
        std::shared_ptr<Texture> emissive = texOrNull(s, md.d.emissive_tex);
        std::shared_ptr<Medium> medium = md.d.medium >= 0 ? s->media[md.d.medium] : nullptr;
        auto m = std::make_shared<Mesh>(indices, vertices, tangents, normals, texCoords, mat, emissive, medium);
        s->meshes.push_back(m);
        for(uint32_t j = 0;j < n;j++){
            std::shared_ptr<Shape> shape = std::shared_ptr<Shape>(m->GetControlPtr(), m->GetShape(j));
            std::shared_ptr<AreaLight> area = emissive != nullptr ? std::make_shared<AreaLight>(shape, emissive) : nullptr;
            if(area){
                area->PreProcess({});
                if(area->Power() <= std::numeric_limits<float>::epsilon())area = nullptr;
            }
            prims.emplace_back(shape, mat, area, medium);
        }
        base += n;
    }
    triCount = base;
    return std::make_shared<BLAS>(prims);
}

} // namespace

extern "C" {

PT_API void* ref_scene_create(){ return new RefScene(); }

PT_API void ref_scene_destroy(void* p){ delete static_cast<RefScene*>(p); }

PT_API int ref_hardware_threads(){ return (int)std::thread::hardware_concurrency(); }

// Image textures can only be built from a file in the reference
// (Texture.hpp:137,169), so the pixels are written to a temporary PNG/HDR with
// the reference's own stb_image_write and loaded back through ImageTexture.
PT_API int ref_scene_add_texture(void* p, const pt_texture_desc* d){
    RefScene* s = static_cast<RefScene*>(p);
    glm::vec3 scale(d->color_scale[0], d->color_scale[1], d->color_scale[2]);
    std::shared_ptr<Texture> t;
    switch(d->kind){
    case PT_TEX_SOLID:
        t = std::make_shared<SolidColor>(glm::vec3(d->color[0], d->color[1], d->color[2]), scale);
        break;
    case PT_TEX_IMAGE_U8: {
        char name[256];
        snprintf(name, sizeof name, "/tmp/ptref_%d_%p_%d.png", (int)getpid(), p, s->tmpCounter++);
        if(!stbi_write_png(name, d->width, d->height, d->channels, d->pixels, d->width * d->channels))return PT_ERR_INVALID;
        t = std::make_shared<ImageTexture>(std::string(name), false, scale);
        remove(name);
        break;
    }
    case PT_TEX_IMAGE_F32: {
        char name[256];
        snprintf(name, sizeof name, "/tmp/ptref_%d_%p_%d.hdr", (int)getpid(), p, s->tmpCounter++);
        if(!stbi_write_hdr(name, d->width, d->height, d->channels, (const float*)d->pixels))return PT_ERR_INVALID;
        t = std::make_shared<FloatImageTexture>(std::string(name), scale);
        remove(name);
        break;
    }
    case PT_TEX_CHECKER:
        t = std::make_shared<CheckerTexture>(s->textures.at(d->tex_a), s->textures.at(d->tex_b), glm::vec2(d->uv_scale[0], d->uv_scale[1]), scale);
        break;
    // UVTexture / NormalTexture only have the implicit default constructor (Texture.hpp:215-232): colorScale is 1
    case PT_TEX_UV: if(scale != glm::vec3(1))return PT_ERR_INVALID; t = std::make_shared<UVTexture>(); break;
    case PT_TEX_NORMAL: if(scale != glm::vec3(1))return PT_ERR_INVALID; t = std::make_shared<NormalTexture>(); break;
    default: return PT_ERR_INVALID;
    }
    s->textures.push_back(t);
    return (int)s->textures.size() - 1;
}

// RGBE round trip of an .hdr file as the reference would see it
// (stbi_write_hdr -> stbi_loadf), so that the GPU can be handed the very same
// decoded texels.
PT_API int ref_hdr_roundtrip(const float* in, int w, int h, int ch, float* out){
    char name[256];
    snprintf(name, sizeof name, "/tmp/ptref_rt_%d_%p.hdr", (int)getpid(), (const void*)in);
    if(!stbi_write_hdr(name, w, h, ch, in))return PT_ERR_INVALID;
    int ww, hh, cc;
    float* data = stbi_loadf(name, &ww, &hh, &cc, 0);
    remove(name);
    if(!data || ww != w || hh != h || cc != ch){ if(data)stbi_image_free(data); return PT_ERR_INVALID; }
    memcpy(out, data, sizeof(float) * (size_t)w * h * ch);
    stbi_image_free(data);
    return PT_OK;
}

PT_API int ref_scene_add_material(void* p, const pt_material_desc* d){
    RefScene* s = static_cast<RefScene*>(p);
    std::shared_ptr<Material> m;
    AlphaTester tester(d->alpha_mode == PT_ALPHA_OPAQUE ? AlphaMode::Opaque : d->alpha_mode == PT_ALPHA_MASK ? AlphaMode::Mask : AlphaMode::Blend, d->alpha_cutoff);
    switch(d->kind){
    case PT_MAT_DIFFUSE: {
        auto rough = d->rough >= 0 ? s->textures.at(d->rough) : std::make_shared<SolidColor>(glm::vec3(1));
        auto metal = d->metal >= 0 ? s->textures.at(d->metal) : std::make_shared<SolidColor>(glm::vec3(0));
        auto md = std::make_shared<MicrofacetDiffuse>(s->textures.at(d->tex), texOrNull(s, d->norm), rough, metal, texOrNull(s, d->alpha));
        // default-constructed tester is Blend (Material.hpp:196); only call
        // setAlphaTester when the caller asks for something else, as user code would
        if(d->alpha_mode != PT_ALPHA_BLEND)md->setAlphaTester(tester);
        m = md;
        break;
    }
    case PT_MAT_DIELECTRIC: {
        auto rough = d->rough >= 0 ? s->textures.at(d->rough) : std::make_shared<SolidColor>(glm::vec3(0.0));
        auto md = std::make_shared<MicrofacetDielectric>(d->ri, s->textures.at(d->tex), texOrNull(s, d->norm), rough, texOrNull(s, d->alpha));
        if(d->alpha_mode != PT_ALPHA_BLEND)md->setAlphaTester(tester);
        m = md;
        break;
    }
    case PT_MAT_THIN_DIELECTRIC:
        m = std::make_shared<ThinDielectric>(d->ri, texOrNull(s, d->tex));
        break;
    case PT_MAT_CONDUCTOR:
        m = std::make_shared<SpecularConductor>(glm::vec3(d->albedo[0], d->albedo[1], d->albedo[2]));
        break;
    default: return PT_ERR_INVALID;
    }
    s->materials.push_back(m);
    return (int)s->materials.size() - 1;
}

PT_API int ref_scene_add_medium(void* p, const pt_medium_desc* d){
    RefScene* s = static_cast<RefScene*>(p);
    s->media.push_back(std::make_shared<HomogeneusMedium>(
        glm::vec3(d->sigma_a[0], d->sigma_a[1], d->sigma_a[2]), glm::vec3(d->sigma_s[0], d->sigma_s[1], d->sigma_s[2]),
        std::make_shared<HenyeyGreenstein>(d->g), d->density, glm::vec3(d->Le[0], d->Le[1], d->Le[2]), d->le_density));
    return (int)s->media.size() - 1;
}

PT_API int ref_scene_add_mesh(void* p, const pt_mesh_desc* d){
    RefScene* s = static_cast<RefScene*>(p);
    RefMesh m;
    m.d = *d;
    m.pos.assign(d->positions, d->positions + 3 * (size_t)d->n_vertices);
    m.nrm.assign(d->normals, d->normals + 3 * (size_t)d->n_vertices);
    m.uv.assign(d->texcoords, d->texcoords + 2 * (size_t)d->n_vertices);
    if(d->tangents)m.tan.assign(d->tangents, d->tangents + 3 * (size_t)d->n_vertices);
    m.idx.assign(d->indices, d->indices + 3 * (size_t)d->n_triangles);
    s->meshDescs.push_back(std::move(m));
    return (int)s->meshDescs.size() - 1;
}

PT_API int ref_scene_add_blas(void* p, const int32_t* mesh_ids, int n){
    RefScene* s = static_cast<RefScene*>(p);
    RefBlasDesc b;
    b.meshes.assign(mesh_ids, mesh_ids + n);
    s->blasDescs.push_back(b);
    return (int)s->blasDescs.size() - 1;
}

PT_API int ref_scene_add_instance(void* p, int blas, const float* xf){
    RefScene* s = static_cast<RefScene*>(p);
    RefInstDesc i;
    i.blas = blas;
    i.has_xf = xf != nullptr;
    if(xf)memcpy(&i.xf[0][0], xf, 16 * sizeof(float));
    s->instDescs.push_back(i);
    return (int)s->instDescs.size() - 1;
}

// GeometricPrimitive(shape, material, areaLight, medium) handed to Scene::Add (examples/example_1.cpp:37-66)
PT_API int ref_scene_add_shape(void* p, const pt_shape_desc* d, const float* xf){
    RefScene* s = static_cast<RefScene*>(p);
    RefInstDesc i;
    i.blas = -1; i.has_xf = xf != nullptr; i.shape = *d;
    if(xf)memcpy(&i.xf[0][0], xf, 16 * sizeof(float));
    s->instDescs.push_back(i);
    s->hasShapes = true;
    return (int)s->instDescs.size() - 1;
}

// AnimatedPrimitive(primitive, direction, timeBounds) around top-level primitive `id` (main.cpp:189)
PT_API int ref_scene_animate(void* p, int id, const float* dir, const float* tb){
    RefScene* s = static_cast<RefScene*>(p);
    if(id < 0 || id >= (int)s->instDescs.size())return PT_ERR_INVALID;
    RefInstDesc& i = s->instDescs[id];
    i.animated = true; i.adir = glm::vec3(dir[0], dir[1], dir[2]); i.atb = glm::vec2(tb[0], tb[1]);
    return PT_OK;
}

PT_API int ref_scene_add_light(void* p, const pt_light_desc* d){
    RefScene* s = static_cast<RefScene*>(p);
    s->lightDescs.push_back(*d);
    return (int)s->lightDescs.size() - 1;
}

PT_API int ref_scene_set_medium(void* p, int id){
    static_cast<RefScene*>(p)->sceneMedium = id;
    return PT_OK;
}

} // extern "C"

namespace {

thread_local uint64_t tlClosest = 0, tlAny = 0;

// Optional ray counter placed at the scene root: forwards to
// the real TLAS, counts calls in thread-local counters.
class CountingPrimitive : public Primitive{
public:
    CountingPrimitive(const std::shared_ptr<Primitive>& inner) : inner(inner){}
    AABB BoundingBox() const final{ return inner->BoundingBox(); }
    bool IntersectPred(const Ray& ray, float max) const final{ ++tlAny; return inner->IntersectPred(ray, max); }
    bool Intersect(const Ray& ray, SurfaceInteraction& interaction, float max) const final{ ++tlClosest; return inner->Intersect(ray, interaction, max); }
    std::vector<std::shared_ptr<Light>> GetLights() const final{ return inner->GetLights(); }
private:
    std::shared_ptr<Primitive> inner;
};

} // namespace

extern "C" {

// id_mode: 1 = ray-cast parity mode (texcoords replaced by triangle ids);
// count_rays: 1 = wrap the TLAS in a CountingPrimitive.
PT_API int ref_scene_commit(void* p, int bvh_width, int light_sampler, int id_mode, int count_rays){
    RefScene* s = static_cast<RefScene*>(p);
    if(bvh_width != 2 && bvh_width != 4 && bvh_width != 8)return PT_ERR_INVALID;
    s->width = bvh_width;
    s->idMode = id_mode != 0;
    std::streambuf* old = std::cout.rdbuf(nullptr); // the BVH ctors print timings
    s->blasTris.resize(s->blasDescs.size());
    for(size_t b = 0;b < s->blasDescs.size();b++){
        s->blas.push_back(bvh_width == 2 ? buildBlas<BLAS2>(s, s->blasDescs[b], s->blasTris[b])
            : bvh_width == 4 ? buildBlas<BLAS4>(s, s->blasDescs[b], s->blasTris[b])
            : buildBlas<BLAS8>(s, s->blasDescs[b], s->blasTris[b]));
    }
    std::shared_ptr<Medium> sceneMed = s->sceneMedium >= 0 ? s->media[s->sceneMedium] : nullptr;
    s->scene = std::make_shared<Scene>(sceneMed);
    for(const RefInstDesc& i : s->instDescs){
        if(i.blas < 0){
            const pt_shape_desc& d = i.shape;
            std::shared_ptr<Shape> shape;
            if(d.kind == PT_SHAPE_SPHERE)shape = std::make_shared<SphereShape>(glm::vec3(d.p[0], d.p[1], d.p[2]), d.radius);
            else shape = std::make_shared<QuadShape>(glm::vec3(d.p[0], d.p[1], d.p[2]), glm::vec3(d.u[0], d.u[1], d.u[2]), glm::vec3(d.v[0], d.v[1], d.v[2]));
            std::shared_ptr<AreaLight> area = d.emissive_tex >= 0 ? std::make_shared<AreaLight>(shape, s->textures.at(d.emissive_tex), d.one_sided != 0) : nullptr;
            std::shared_ptr<Primitive> gp = std::make_shared<GeometricPrimitive>(shape, d.material >= 0 ? s->materials.at(d.material) : nullptr, area,
                                                                                 d.medium >= 0 ? s->media.at(d.medium) : nullptr);
            if(i.has_xf)gp = std::make_shared<TransformedPrimitive>(gp, i.xf);
            if(i.animated)gp = std::make_shared<AnimatedPrimitive>(gp, i.adir, i.atb);
            s->topLevel.push_back(gp);
            continue;
        }
        std::shared_ptr<Primitive> prim = s->blas.at(i.blas);
        if(i.has_xf)prim = std::make_shared<TransformedPrimitive>(prim, i.xf);
        if(i.animated)prim = std::make_shared<AnimatedPrimitive>(prim, i.adir, i.atb);
        s->topLevel.push_back(prim);
    }
    if(count_rays){
        std::shared_ptr<Primitive> inner;
        if(bvh_width == 2)inner = std::make_shared<TLAS2>(s->topLevel);
        else if(bvh_width == 4)inner = std::make_shared<TLAS4>(s->topLevel);
        else inner = std::make_shared<TLAS8>(s->topLevel);
        s->scene->Add(std::make_shared<CountingPrimitive>(inner));
    } else{
        for(auto& prim : s->topLevel)s->scene->Add(prim);
    }
    if(bvh_width == 2)s->scene->BuildTlas<TLAS2>();
    else if(bvh_width == 4)s->scene->BuildTlas<TLAS4>();
    else s->scene->BuildTlas<TLAS8>();

    if(light_sampler == 0)s->lightSampler = std::make_shared<UniformLightSampler>();
    else s->lightSampler = std::make_shared<PowerLightSampler>();
    std::vector<std::shared_ptr<Light>> extra;
    for(const pt_light_desc& d : s->lightDescs){
        glm::vec3 c(d.color[0], d.color[1], d.color[2]);
        glm::vec3 cb(d.color_b[0], d.color_b[1], d.color_b[2]);
        glm::vec3 v(d.dir_or_pos[0], d.dir_or_pos[1], d.dir_or_pos[2]);
        float scale = d.scale;
        switch(d.kind){
        case PT_LIGHT_UNIFORM_INFINITE:
            s->scene->infiniteLights.push_back(std::make_shared<UniformInfiniteLight>(c, powerFunc(d.power_func)));
            break;
        case PT_LIGHT_GRADIENT_INFINITE:
            s->scene->infiniteLights.push_back(std::make_shared<FunctionInfiniteLight>([=](const Ray& ray){
                float a = 0.5f * (ray.dir.y + 1.0f);
                return scale * ((1.0f - a) * c + a * cb);
                }, powerFunc(d.power_func)));
            break;
        case PT_LIGHT_TEXTURE_INFINITE:
            s->scene->infiniteLights.push_back(std::make_shared<TextureInfiniteLight>(s->textures.at(d.tex), scale, powerFunc(d.power_func)));
            break;
        case PT_LIGHT_DISTANT: extra.push_back(std::make_shared<DistantLight>(v, c, powerFunc(d.power_func))); break;
        case PT_LIGHT_POINT: extra.push_back(std::make_shared<PointLight>(v, c, powerFunc(d.power_func))); break;
        default: std::cout.rdbuf(old); return PT_ERR_INVALID;
        }
    }
    // examples/example_1.cpp:99-101
    s->allLights = s->scene->GetLights();
    for(auto& l : extra)s->allLights.push_back(l);
    s->lightSampler->Add(s->allLights);
    s->lightSampler->PreProcess(s->scene->BoundingBox());
    std::cout.rdbuf(old);
    return PT_OK;
}

PT_API int ref_scene_bounds(void* p, float bmin[3], float bmax[3]){
    RefScene* s = static_cast<RefScene*>(p);
    if(!s->scene)return PT_ERR_STATE;
    AABB b = s->scene->BoundingBox();
    for(int i = 0;i < 3;i++){ bmin[i] = b.min[i]; bmax[i] = b.max[i]; }
    return PT_OK;
}

// Scene::Intersect over a ray batch on `threads` std::threads (strided slices,
// BASELINE.md §3).  In id mode hits[i].prim is the triangle id decoded from
// si.uv and u,v are NaN; otherwise prim is PT_MISS-1 and (u,v) = si.uv.
// inst (optional) is found by a linear scan over the top-level primitives.
PT_API int ref_trace_closest(void* p, const pt_ray* rays, size_t n, pt_hit* hits, uint32_t* inst, int threads, double* seconds){
    RefScene* s = static_cast<RefScene*>(p);
    if(!s->scene)return PT_ERR_STATE;
    if(threads < 1)threads = 1;
    auto work = [&](int tid){
        for(size_t i = tid;i < n;i += threads){
            const pt_ray& r = rays[i];
            Ray ray(glm::vec3(r.ox, r.oy, r.oz), glm::vec3(r.dx, r.dy, r.dz), r.time);
            SurfaceInteraction si;
            bool hit = s->scene->Intersect(ray, si, r.tmax);
            pt_hit h;
            if(!hit){
                h.t = std::numeric_limits<float>::infinity(); h.u = h.v = 0; h.prim = PT_MISS;
                if(inst)inst[i] = PT_MISS;
            } else{
                h.t = si.t;
                uint32_t best = PT_MISS;
                if(inst || s->hasShapes){
                    float max = r.tmax;
                    for(size_t k = 0;k < s->topLevel.size();k++){
                        SurfaceInteraction tmp;
                        if(s->topLevel[k]->Intersect(ray, tmp, max)){ max = tmp.t; best = (uint32_t)k; }
                    }
                    if(inst)inst[i] = best;
                }
                if(best != PT_MISS && s->instDescs[best].blas < 0){
                    // a sphere / quad: prim 0; u, v = the quad's (alpha, beta), 0 for a sphere (include/pt_b200.h)
                    h.prim = 0;
                    h.u = s->instDescs[best].shape.kind == PT_SHAPE_QUAD ? si.uv.x : 0.0f;
                    h.v = s->instDescs[best].shape.kind == PT_SHAPE_QUAD ? si.uv.y : 0.0f;
                } else if(s->idMode){
                    uint32_t lo = (uint32_t)std::lround(si.uv.x), hi = (uint32_t)std::lround(si.uv.y);
                    h.prim = lo | (hi << 12);
                    h.u = h.v = std::numeric_limits<float>::quiet_NaN();
                } else{
                    h.prim = PT_MISS - 1; h.u = si.uv.x; h.v = si.uv.y;
                }
            }
            hits[i] = h;
        }
        };
    auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> pool;
    for(int t = 0;t < threads;t++)pool.emplace_back(work, t);
    for(auto& t : pool)t.join();
    if(seconds)*seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return PT_OK;
}

PT_API int ref_trace_any(void* p, const pt_ray* rays, size_t n, uint8_t* occluded, int threads, double* seconds){
    RefScene* s = static_cast<RefScene*>(p);
    if(!s->scene)return PT_ERR_STATE;
    if(threads < 1)threads = 1;
    auto work = [&](int tid){
        for(size_t i = tid;i < n;i += threads){
            const pt_ray& r = rays[i];
            Ray ray(glm::vec3(r.ox, r.oy, r.oz), glm::vec3(r.dx, r.dy, r.dz), r.time);
            occluded[i] = s->scene->IntersectPred(ray, r.tmax) ? 1 : 0;
        }
        };
    auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> pool;
    for(int t = 0;t < threads;t++)pool.emplace_back(work, t);
    for(auto& t : pool)t.join();
    if(seconds)*seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return PT_OK;
}

typedef struct ref_render_out {
    uint64_t samples, closest_rays, shadow_rays;
    double seconds;
} ref_render_out;

// Fixed-spp mirror of TileIntegrator::Render (Integrators.cpp:23-129): same
// 32x32 tile queue, same per-sample calls (StartPixelSample, getPixel2D, get1D,
// GenerateRay, Li, FilmTile::Add), adaptive loop disabled.  rgbw: 4*W*H doubles
// (sum RGB*w, sum w) merged from the reference's own FilmTile objects; sumsq
// (optional): 3*W*H per-pixel sum of squared sample radiance (sample assigned
// to pixel floor(p)), for the 3-sigma test.
PT_API int ref_render(void* p, const pt_camera_desc* cd, const pt_render_params* rp, int threads, double* rgbw, double* sumsq, ref_render_out* out){
    RefScene* s = static_cast<RefScene*>(p);
    if(!s->scene)return PT_ERR_STATE;
    std::shared_ptr<Filter> filter;
    glm::vec2 rad(rp->filter_radius[0], rp->filter_radius[1]);
    switch(rp->filter){
    case PT_FILTER_BOX: filter = std::make_shared<BoxFilter>(rad); break;
    case PT_FILTER_GAUSSIAN: filter = std::make_shared<GaussianFilter>(rad, rp->filter_param[0]); break;
    case PT_FILTER_MITCHELL: filter = std::make_shared<MitchellFilter>(rad, rp->filter_param[0], rp->filter_param[1]); break;
    case PT_FILTER_LANCZOS: filter = std::make_shared<LanczosFilter>(rad, rp->filter_param[0]); break;
    default: return PT_ERR_INVALID;
    }
    auto film = std::make_shared<Film>(glm::ivec2 { rp->width, rp->height }, filter);
    glm::vec3 from(cd->look_from[0], cd->look_from[1], cd->look_from[2]), at(cd->look_at[0], cd->look_at[1], cd->look_at[2]);
    std::shared_ptr<Medium> camMed = cd->medium >= 0 ? s->media.at(cd->medium) : nullptr;
    std::shared_ptr<Camera> camera;
    if(cd->focus_angle == 0 || cd->focus_distance == 0){
        camera = std::make_shared<Camera>(from, at, cd->fov, film, glm::vec2(cd->shutter_start, cd->shutter_end));
        camera->SetMedium(camMed);
    } else{
        camera = std::make_shared<Camera>(from, at, cd->fov, film, cd->focus_angle, cd->focus_distance, camMed);
    }
    std::shared_ptr<Sampler> sampler;
    if(rp->sampler == 0)sampler = std::make_shared<UniformSampler>(rp->spp);
    else sampler = std::make_shared<StratifiedSampler>(rp->strata_x, rp->strata_y);
    std::shared_ptr<Integrator> integrator;
    switch(rp->integrator){
    case PT_INTEGRATOR_SIMPLE: integrator = std::make_shared<SimplePathIntegrator>(s->scene, camera, sampler, rp->max_depth); break;
    case PT_INTEGRATOR_PATH: integrator = std::make_shared<PathIntegrator>(s->scene, camera, sampler, s->lightSampler, rp->max_depth); break;
    case PT_INTEGRATOR_VOLPATH: integrator = std::make_shared<VolPathIntegrator>(s->scene, camera, sampler, s->lightSampler, rp->max_depth); break;
    default: return PT_ERR_INVALID;
    }
    const int W = rp->width, H = rp->height;
    memset(rgbw, 0, sizeof(double) * 4 * (size_t)W * H);
    if(sumsq)memset(sumsq, 0, sizeof(double) * 3 * (size_t)W * H);
    if(threads < 1)threads = 1;
    constexpr int tileSize = 32;
    const int tilesX = (W + tileSize - 1) / tileSize, tilesY = (H + tileSize - 1) / tileSize;
    const int tileCount = tilesX * tilesY;
    std::atomic<int> next { 0 };
    std::atomic<uint64_t> nSamples { 0 }, nClosest { 0 }, nAny { 0 };
    std::mutex mergeLock;
    const unsigned spp = sampler->SamplesPerPixel();
    auto work = [&](){
        std::shared_ptr<Sampler> cloned = sampler->Clone();
        tlClosest = tlAny = 0;
        uint64_t mySamples = 0;
        int tileNum;
        while((tileNum = next.fetch_add(1, std::memory_order_relaxed)) < tileCount){
            int tileX = tileNum % tilesX, tileY = tileNum / tilesX;
            int minX = tileX * tileSize, minY = tileY * tileSize;
            int maxX = std::min((tileX + 1) * tileSize, W), maxY = std::min((tileY + 1) * tileSize, H);
            FilmTile tile = film->GetFilmTile({ {minX,minY},{maxX,maxY} });
            for(int y = minY;y < maxY;y++){
                for(int x = minX;x < maxX;x++){
                    // rp->adaptive: the stopping rule of TileIntegrator::Render (Integrators.cpp:59-86) with the
                    // reference's own VarianceEstimator; otherwise exactly one round of spp samples
                    VarianceEstimator estimator[3];
                    while(estimator[0].Samples() < (rp->adaptive ? 128u : 1u) * spp){
                        for(uint32_t si = 0;si < spp;si++){
                            cloned->StartPixelSample({ x,y }, si);
                            glm::dvec2 pp = glm::dvec2 { x,y } + cloned->getPixel2D();
                            float time = cloned->get1D();
                            Ray ray = camera->GenerateRay(pp, time, cloned->get2D());
                            glm::dvec3 color = integrator->Li(ray);
                            tile.Add(pp, color);
                            if(sumsq){
                                // owner pixel is this thread's tile: no race
                                double* q = sumsq + 3 * ((size_t)y * W + x);
                                q[0] += color.x * color.x; q[1] += color.y * color.y; q[2] += color.z * color.z;
                            }
                            ++mySamples;
                            color *= glm::dvec3(0.2126f, 0.7152f, 0.0722f);
                            estimator[0].Add(color[0]); estimator[1].Add(color[1]); estimator[2].Add(color[2]);
                        }
                        const float minRelativeVariance = 1.5;
                        if(estimator[0].RelativeVariance() <= minRelativeVariance && estimator[1].RelativeVariance() <= minRelativeVariance &&
                           estimator[2].RelativeVariance() <= minRelativeVariance)break;
                    }
                }
            }
            Bounds2i tb = tile.TileBounds();
            std::lock_guard<std::mutex> g(mergeLock); // Film::Merge (Film.hpp:125-132) into a readable buffer
            for(int y = tb.min.y;y < tb.max.y;y++){
                for(int x = tb.min.x;x < tb.max.x;x++){
                    const FilmTilePixel& px = tile.At({ x,y });
                    double* q = rgbw + 4 * ((size_t)y * W + x);
                    q[0] += px.RGB.x; q[1] += px.RGB.y; q[2] += px.RGB.z; q[3] += px.weight;
                }
            }
        }
        nSamples += mySamples; nClosest += tlClosest; nAny += tlAny;
        };
    auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> pool;
    for(int t = 0;t < threads;t++)pool.emplace_back(work);
    for(auto& t : pool)t.join();
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if(out){ out->samples = nSamples; out->closest_rays = nClosest; out->shadow_rays = nAny; out->seconds = sec; }
    return PT_OK;
}

// The stock, unmodified Render() (adaptive sampling on), wall time only.
PT_API int ref_render_stock(void* p, const pt_camera_desc* cd, const pt_render_params* rp, int threads, double* seconds){
    RefScene* s = static_cast<RefScene*>(p);
    if(!s->scene)return PT_ERR_STATE;
    auto film = std::make_shared<Film>(glm::ivec2 { rp->width, rp->height }, std::make_shared<BoxFilter>());
    glm::vec3 from(cd->look_from[0], cd->look_from[1], cd->look_from[2]), at(cd->look_at[0], cd->look_at[1], cd->look_at[2]);
    auto camera = std::make_shared<Camera>(from, at, cd->fov, film, glm::vec2(cd->shutter_start, cd->shutter_end));
    std::shared_ptr<Sampler> sampler = std::make_shared<UniformSampler>(rp->spp);
    std::shared_ptr<Integrator> integrator = std::make_shared<PathIntegrator>(s->scene, camera, sampler, s->lightSampler, rp->max_depth);
    std::streambuf* old = std::cout.rdbuf(nullptr);
    auto t0 = std::chrono::steady_clock::now();
    integrator->Render(threads);
    if(seconds)*seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::cout.rdbuf(old);
    return PT_OK;
}

} // extern "C"


// ---- deterministic unit probes (tests/test_shading_units.py): the reference's own shading routines at
// caller-chosen inputs, next to tests/emul's host build of the product's shade*.cuh ------------------------
extern "C" {

// per ray, PT_UNIT_STRIDE floats (layout mirrored by tests/emul/emul.cpp emul_unit_shade):
//  0 hit | 1-3 p | 4-6 n | 7-9 ns | 10-12 tangent | 13-14 uv | 15 t | 16 material id | 17 has area light
//  18 scatter ok | 19-21 f | 22 pdf | 23 flags | 24-26 scattered origin | 27-29 scattered dir
//  30-32 calc_attenuation(sampled dir) | 33 PDF(sampled dir) | 34-36 calc_attenuation(w2) | 37 PDF(w2)
//  38-40 AreaLight::L(si, ray) | 41 AreaLight PMF | 42 AreaLight PDF(si, ray)
// in: per ray u, ux, uy, w2.xyz (6 floats)
PT_API int ref_unit_shade(void* p, const pt_ray* rays, size_t n, const float* in, float* out, int stride){
    RefScene* s = static_cast<RefScene*>(p);
    if(!s->scene || stride < 43)return PT_ERR_STATE;
    for(size_t i = 0;i < n;i++){
        float* o = out + i * (size_t)stride;
        for(int k = 0;k < stride;k++)o[k] = 0.0f;
        const pt_ray& r = rays[i];
        Ray ray(glm::vec3(r.ox, r.oy, r.oz), glm::vec3(r.dx, r.dy, r.dz), r.time);
        SurfaceInteraction si;
        if(!s->scene->Intersect(ray, si, r.tmax))continue;
        o[0] = 1.0f;
        for(int k = 0;k < 3;k++){ o[1 + k] = si.p[k]; o[4 + k] = si.n[k]; o[7 + k] = si.ns[k]; o[10 + k] = si.tangent[k]; }
        o[13] = si.uv.x; o[14] = si.uv.y; o[15] = si.t;
        o[16] = -1.0f;
        for(size_t m = 0;m < s->materials.size();m++)if(s->materials[m] == si.mat)o[16] = (float)m;
        o[17] = si.AreaLight ? 1.0f : 0.0f;
        const float* q = in + 6 * i;
        if(si.mat){
            Ray scattered;
            std::optional<BxDFSample> bs = si.mat->scatter(ray, si, scattered, q[0], glm::vec2(q[1], q[2]));
            if(bs){
                o[18] = 1.0f;
                for(int k = 0;k < 3;k++){ o[19 + k] = bs->f[k]; o[24 + k] = scattered.origin[k]; o[27 + k] = scattered.dir[k]; }
                o[22] = bs->pdf; o[23] = (float)bs->flags;
                glm::vec3 a = si.mat->calc_attenuation(ray, si, scattered);
                for(int k = 0;k < 3;k++)o[30 + k] = a[k];
                o[33] = si.mat->PDF(ray, si, scattered);
            }
            Ray other(si.p, glm::normalize(glm::vec3(q[3], q[4], q[5])), r.time);
            glm::vec3 a2 = si.mat->calc_attenuation(ray, si, other);
            for(int k = 0;k < 3;k++)o[34 + k] = a2[k];
            o[37] = si.mat->PDF(ray, si, other);
        }
        if(si.AreaLight){
            glm::vec3 L = si.AreaLight->L(si, ray);
            for(int k = 0;k < 3;k++)o[38 + k] = L[k];
            o[41] = s->lightSampler->PMF(si.AreaLight);
            o[42] = si.AreaLight->PDF(si, ray);
        }
    }
    return PT_OK;
}

PT_API int ref_unit_light_count(void* p){ return (int)static_cast<RefScene*>(p)->allLights.size(); }

// per light: 0 Power | 1 PMF | 2 isDelta
PT_API int ref_unit_lights(void* p, float* out){
    RefScene* s = static_cast<RefScene*>(p);
    for(size_t k = 0;k < s->allLights.size();k++){
        out[3 * k] = s->allLights[k]->Power(); out[3 * k + 1] = s->lightSampler->PMF(s->allLights[k]); out[3 * k + 2] = s->allLights[k]->isDelta() ? 1.0f : 0.0f;
    }
    return PT_OK;
}

// Light::sample(uv) of light k and what SampleLd evaluates next for a shading point ip (Integrators.cpp:260-294):
// in: ux, uy, ip.xyz | out: 0-2 sample.L | 3-5 p | 6-8 n | 9-10 uv | 11-13 dir | 14-16 shadow dir | 17 t | 18 PDF(sample, shadow ray) | 19-21 L(sample, shadow ray)
PT_API int ref_unit_light_sample(void* p, int k, size_t n, const float* in, float* out, int stride, float time){
    RefScene* s = static_cast<RefScene*>(p);
    if(k < 0 || k >= (int)s->allLights.size() || stride < 22)return PT_ERR_INVALID;
    const std::shared_ptr<Light>& light = s->allLights[k];
    for(size_t i = 0;i < n;i++){
        const float* q = in + 5 * i;
        float* o = out + i * (size_t)stride;
        LightSample ls = light->sample(glm::vec2(q[0], q[1]), time);
        glm::vec3 ip(q[2], q[3], q[4]), lightDir; float t;
        if(ls.isDeltaInteraction()){ lightDir = ls.dir; t = std::numeric_limits<float>::infinity(); }
        else{ lightDir = ls.interaction.p - ip; t = glm::length(lightDir) - shadowEpsilon; }
        Ray shadow(ip, glm::normalize(lightDir), time);
        for(int c = 0;c < 3;c++){ o[c] = ls.L[c]; o[3 + c] = ls.interaction.p[c]; o[6 + c] = ls.interaction.n[c]; o[11 + c] = ls.dir[c]; o[14 + c] = shadow.dir[c]; }
        o[9] = ls.interaction.uv.x; o[10] = ls.interaction.uv.y; o[17] = t;
        o[18] = light->PDF(ls.interaction, shadow);
        glm::vec3 L = light->L(ls.interaction, shadow);
        for(int c = 0;c < 3;c++)o[19 + c] = L[c];
    }
    return PT_OK;
}

// infinite lights seen by an escaping ray: sum over scene->infiniteLights of Le(ray), and each one's PDF * PMF
PT_API int ref_unit_escape(void* p, size_t n, const float* dirs, float* out){
    RefScene* s = static_cast<RefScene*>(p);
    for(size_t i = 0;i < n;i++){
        Ray ray(glm::vec3(0), glm::normalize(glm::vec3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2])), 0.0f);
        glm::vec3 le(0); float pp = 0;
        for(auto& l : s->scene->infiniteLights){ le += l->Le(ray); pp += s->lightSampler->PMF(l) * l->PDF(GeometricInteraction {}, ray); }
        out[4 * i] = le.x; out[4 * i + 1] = le.y; out[4 * i + 2] = le.z; out[4 * i + 3] = pp;
    }
    return PT_OK;
}

// Camera::GenerateRay(p, time, lens) for n samples: in px, py, time, lx, ly | out origin, dir, time
PT_API int ref_unit_camera(const pt_camera_desc* cd, int W, int H, size_t n, const double* in, float* out){
    auto film = std::make_shared<Film>(glm::ivec2 { W, H }, std::make_shared<BoxFilter>());
    glm::vec3 from(cd->look_from[0], cd->look_from[1], cd->look_from[2]), at(cd->look_at[0], cd->look_at[1], cd->look_at[2]);
    std::shared_ptr<Camera> camera;
    if(cd->focus_angle == 0 || cd->focus_distance == 0)camera = std::make_shared<Camera>(from, at, cd->fov, film, glm::vec2(cd->shutter_start, cd->shutter_end));
    else camera = std::make_shared<Camera>(from, at, cd->fov, film, cd->focus_angle, cd->focus_distance, nullptr);
    for(size_t i = 0;i < n;i++){
        const double* q = in + 5 * i;
        Ray r = camera->GenerateRay(glm::dvec2(q[0], q[1]), (float)q[2], glm::vec2((float)q[3], (float)q[4]));
        for(int c = 0;c < 3;c++){ out[7 * i + c] = r.origin[c]; out[7 * i + 3 + c] = r.dir[c]; }
        out[7 * i + 6] = r.time;
    }
    return PT_OK;
}

// StratifiedSampler's building blocks (Util.hpp:45-168) and the phase function (PhaseFunction.cpp:4-26)
PT_API uint32_t ref_unit_permutation_element(uint32_t i, uint32_t l, uint32_t p){ return (uint32_t)PermutationElement(i, l, p); }
PT_API uint64_t ref_unit_hash(uint32_t px, uint32_t py, uint64_t dim){ return Hash(px, py, dim); }
PT_API int ref_unit_hg(float g, size_t n, const float* in, float* out){
    HenyeyGreenstein hg(g);
    for(size_t i = 0;i < n;i++){
        const float* q = in + 5 * i;
        glm::vec3 wi = glm::normalize(glm::vec3(q[0], q[1], q[2])), wo;
        float pdf = hg.Sample(wi, wo, glm::vec2(q[3], q[4]));
        out[5 * i] = wo.x; out[5 * i + 1] = wo.y; out[5 * i + 2] = wo.z; out[5 * i + 3] = pdf; out[5 * i + 4] = hg.PDF(wi, wo);
    }
    return PT_OK;
}

} // extern "C"

// ---- the reference's own film writers (Film.hpp:154-217) on caller-chosen pixel values -------------------------------
// rgb: W*H*3 floats, row 0 = the film's row 0 (bottom row of the written image).  Every pixel receives ONE sample of that
// radiance through the reference's FilmTile::Add / Film::Merge with a box filter (weight 1), so the film holds exactly
// these values; then Film::WritePPM / WritePNG / WriteJPG with the reference's default tone mapper write
// <path_noext>.ppm / .png / .jpg.  fmt: 0 ppm, 1 png, 2 jpg.
extern "C" int ref_film_write(const float* rgb, int W, int H, const char* path_noext, int fmt, int quality){
    auto film = std::make_shared<Film>(glm::ivec2 { W, H }, std::make_shared<BoxFilter>());
    FilmTile tile = film->GetFilmTile({ {0,0},{W,H} });
    for(int y = 0;y < H;y++)for(int x = 0;x < W;x++){
        const float* c = rgb + 3 * ((size_t)y * W + x);
        tile.Add(glm::dvec2 { x + 0.5, y + 0.5 }, glm::dvec3 { c[0], c[1], c[2] });
    }
    film->Merge(tile);
    film->setOutputPath(std::filesystem::path(path_noext));
    if(fmt == 0)film->WritePPM("unused");
    else if(fmt == 1)film->WritePNG("unused");
    else film->WriteJPG("unused", quality);
    return 0;
}
