// PathTracing.hpp — the reference's C++20 object API (Scene / Mesh / Model / BLASx / TLASx / Camera /
// Film / Sampler / *Integrator, examples/example_1.cpp:17-122) as thin host objects over the C-ABI of
// include/pt_b200.h.  The objects hold plain descriptions; nothing is traced on the CPU.  The GPU scene
// is assembled and committed the first time an integrator renders (or a ray is cast), because the
// reference lets the light sampler be attached after Scene::BuildTlas (example_1.cpp:97-102).
//
// Same class names, constructor signatures and method names as the reference headers cited per class.
// What has no GPU counterpart yet is a compile-time absence (SphereShape / QuadShape, BVH8, file-loading
// texture ctors, per-triangle GeometricPrimitive lists), not a silent approximation.
#pragma once
#include <algorithm>
#include <cstdlib>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <functional>
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#include <glm/glm.hpp>

#include "pt_b200.h"
#include "ImageIO.hpp"

namespace ptb_host {

inline void check(int rc, const char* what){
    if(rc < 0)throw std::runtime_error(std::string(what) + ": " + pt_last_error());
}

// Collects ids while the object graph is lowered into one pt_scene.
struct Builder {
    pt_scene* scene = nullptr;
    std::unordered_map<const void*, int> ids;     // texture / material / medium / mesh / blas object -> id
    int lookup(const void* p) const{ auto it = ids.find(p); return it == ids.end() ? -1 : it->second; }
};

} // namespace ptb_host

// ---- Ray.hpp:14-41, AABB.hpp:9-70, Interaction.hpp:36-51 ------------------------------------------
class Medium;
struct Ray {
    Ray() = default;
    Ray(const glm::vec3& origin, const glm::vec3& direction, float time = 0.0f, const std::shared_ptr<Medium>& medium = nullptr)
        : origin(origin), dir(direction), time(time), medium(medium){}
    glm::vec3 at(float t) const{ return origin + dir * t; }
    glm::vec3 origin { 0, 0, 0 }, dir { 0, 0, 1 };
    float time = 0.0f;
    std::shared_ptr<Medium> medium;
};

struct AABB {
    glm::vec3 min { std::numeric_limits<float>::infinity() }, max { -std::numeric_limits<float>::infinity() };
    void Expand(const glm::vec3& p){
        for(int a = 0; a < 3; a++){ if(p[a] < min[a])min[a] = p[a]; if(p[a] > max[a])max[a] = p[a]; }
    }
};

// What one GPU hit record carries back; the shading frame itself is rebuilt on the device
// (csrc/shade.cuh build_interaction) and never crosses the boundary.
struct SurfaceInteraction {
    glm::vec3 p { 0, 0, 0 };
    float t = std::numeric_limits<float>::infinity();
    glm::vec2 barycentric { 0, 0 };
    uint32_t primitive = PT_MISS;    // triangle index inside its BLAS
    uint32_t instance = PT_MISS;     // index in Scene::Add order
};

// ---- Texture.hpp:108-232 --------------------------------------------------------------------------
class Texture {
public:
    Texture(const glm::vec3& colorScale = glm::vec3(1, 1, 1)) : colorScale(colorScale){}
    virtual ~Texture() = default;
    virtual int Channels() const{ return 3; }
    int Emit(ptb_host::Builder& b) const{
        int id = b.lookup(this);
        if(id >= 0)return id;
        pt_texture_desc d;
        std::memset(&d, 0, sizeof d);
        d.color_scale[0] = colorScale.x; d.color_scale[1] = colorScale.y; d.color_scale[2] = colorScale.z;
        d.tex_a = d.tex_b = -1;
        Describe(b, d);
        id = pt_scene_add_texture(b.scene, &d);
        ptb_host::check(id, "pt_scene_add_texture");
        b.ids[this] = id;
        return id;
    }
protected:
    virtual void Describe(ptb_host::Builder& b, pt_texture_desc& d) const = 0;
    glm::vec3 colorScale;
};

class SolidColor : public Texture {
public:
    SolidColor(const glm::vec3& color, const glm::vec3& colorScale = glm::vec3(1)) : Texture(colorScale), color(color){}
    SolidColor(float r, float g, float b) : SolidColor(glm::vec3(r, g, b)){}
protected:
    void Describe(ptb_host::Builder&, pt_texture_desc& d) const override{
        d.kind = PT_TEX_SOLID; d.color[0] = color.x; d.color[1] = color.y; d.color[2] = color.z;
    }
private:
    glm::vec3 color;
};

// The reference builds image textures from a file name (Texture.hpp:139, Image::Image Texture.cpp:4-19: stbi_load, then the
// sRGB -> linear LUT of Texture.hpp:26-34 on the colour channels when gammaCorrection is set).  Same constructor here, decoding
// through ImageIO.hpp (PNG / PPM / BMP / TGA); a second constructor takes texels that are already decoded (row 0 first, 1-4
// channels).  `invertTex` is accepted and, as in the reference (Texture.hpp:110-119: stored, never read by Evaluate), has no effect.
class ImageTexture : public Texture {
public:
    ImageTexture(const std::string& filename, bool gammaCorrection = false, const glm::vec3& colorScale = glm::vec3(1), bool /*invertTex*/ = false)
        : Texture(colorScale){
        ptb_io::ImageU8 img = ptb_io::load_u8(filename);        // throws where the reference aborts (Texture.cpp:6-9)
        width = img.width; height = img.height; channels = img.channels; texels = std::move(img.data);
        if(gammaCorrection){
            uint8_t lut[256];
            for(int i = 0; i < 256; i++){
                const double srgb = i / 255.0, lin = srgb <= 0.04045 ? srgb / 12.92 : std::pow((srgb + 0.055) / 1.055, 2.4);
                lut[i] = (uint8_t)std::lround(std::min(1.0, std::max(0.0, lin)) * 255.0);
            }
            for(size_t i = 0; i < texels.size(); i++)if((int)(i % (size_t)channels) < std::min(channels, 3))texels[i] = lut[texels[i]];
        }
    }
    ImageTexture(std::vector<uint8_t> texels, int width, int height, int channels, const glm::vec3& colorScale = glm::vec3(1))
        : Texture(colorScale), texels(std::move(texels)), width(width), height(height), channels(channels){
        if((size_t)width * height * channels != this->texels.size())throw std::invalid_argument("ImageTexture: texel count");
    }
    int Channels() const override{ return channels; }
protected:
    void Describe(ptb_host::Builder&, pt_texture_desc& d) const override{
        d.kind = PT_TEX_IMAGE_U8; d.width = width; d.height = height; d.channels = channels; d.pixels = texels.data();
    }
private:
    std::vector<uint8_t> texels;
    int width, height, channels;
};

class FloatImageTexture : public Texture {
public:
    // Texture.hpp:170 / FloatImage::FloatImage (Texture.cpp:26-32: stbi_loadf): Radiance .hdr and PFM natively, LDR files as (v / 255)^2.2
    FloatImageTexture(const std::string& filename, const glm::vec3& colorScale = glm::vec3(1), bool /*invertTex*/ = false) : Texture(colorScale){
        ptb_io::ImageF32 img = ptb_io::load_f32(filename);
        width = img.width; height = img.height; channels = img.channels; texels = std::move(img.data);
    }
    FloatImageTexture(std::vector<float> texels, int width, int height, int channels, const glm::vec3& colorScale = glm::vec3(1))
        : Texture(colorScale), texels(std::move(texels)), width(width), height(height), channels(channels){
        if((size_t)width * height * channels != this->texels.size())throw std::invalid_argument("FloatImageTexture: texel count");
    }
    int Channels() const override{ return channels; }
protected:
    void Describe(ptb_host::Builder&, pt_texture_desc& d) const override{
        d.kind = PT_TEX_IMAGE_F32; d.width = width; d.height = height; d.channels = channels; d.pixels = texels.data();
    }
private:
    std::vector<float> texels;
    int width, height, channels;
};

class CheckerTexture : public Texture {
public:
    CheckerTexture(const std::shared_ptr<Texture>& textureA, const std::shared_ptr<Texture>& textureB, const glm::vec2& uvscale,
                   const glm::vec3& colorScale = glm::vec3(1)) : Texture(colorScale), tex1(textureA), tex2(textureB), uvscale(uvscale){}
    int Channels() const override{ return tex1->Channels(); }
protected:
    void Describe(ptb_host::Builder& b, pt_texture_desc& d) const override{
        d.kind = PT_TEX_CHECKER; d.tex_a = tex1->Emit(b); d.tex_b = tex2->Emit(b); d.uv_scale[0] = uvscale.x; d.uv_scale[1] = uvscale.y;
    }
private:
    std::shared_ptr<Texture> tex1, tex2;
    glm::vec2 uvscale;
};

class UVTexture : public Texture {
protected:
    void Describe(ptb_host::Builder&, pt_texture_desc& d) const override{ d.kind = PT_TEX_UV; }
};
class NormalTexture : public Texture {
protected:
    void Describe(ptb_host::Builder&, pt_texture_desc& d) const override{ d.kind = PT_TEX_NORMAL; }
};

// ---- Material.hpp:147-673 -------------------------------------------------------------------------
enum class AlphaMode { Opaque, Blend, Mask };
struct AlphaTester {
    AlphaTester(AlphaMode mode, float cutoff = 0.5f) : mode(mode), cutoff(cutoff){}
    AlphaTester() = default;
    AlphaMode mode = AlphaMode::Blend;      // the reference's default (Material.hpp:196)
    float cutoff = 0.5f;
};

class Material {
public:
    virtual ~Material() = default;
    int Emit(ptb_host::Builder& b) const{
        int id = b.lookup(this);
        if(id >= 0)return id;
        pt_material_desc d;
        std::memset(&d, 0, sizeof d);
        d.tex = d.norm = d.rough = d.metal = d.alpha = -1;
        d.ri = 1.0f; d.alpha_mode = PT_ALPHA_BLEND; d.alpha_cutoff = 0.5f;
        Describe(b, d);
        id = pt_scene_add_material(b.scene, &d);
        ptb_host::check(id, "pt_scene_add_material");
        b.ids[this] = id;
        return id;
    }
protected:
    virtual void Describe(ptb_host::Builder& b, pt_material_desc& d) const = 0;
    static int opt(ptb_host::Builder& b, const std::shared_ptr<Texture>& t){ return t ? t->Emit(b) : -1; }
    static int mode(const AlphaTester& t){ return t.mode == AlphaMode::Opaque ? PT_ALPHA_OPAQUE : t.mode == AlphaMode::Mask ? PT_ALPHA_MASK : PT_ALPHA_BLEND; }
};

class MicrofacetDiffuse : public Material {
public:
    MicrofacetDiffuse(const glm::vec3& albedo) : MicrofacetDiffuse(std::make_shared<SolidColor>(albedo)){}
    MicrofacetDiffuse(const std::shared_ptr<Texture>& tex, const std::shared_ptr<Texture>& norm = nullptr,
                      const std::shared_ptr<Texture>& roughnessTexture = std::make_shared<SolidColor>(glm::vec3(1)),
                      const std::shared_ptr<Texture>& metallicTexture = std::make_shared<SolidColor>(glm::vec3(0)),
                      const std::shared_ptr<Texture>& alpha = nullptr)
        : tex(tex), norm(norm), roughnessTexture(roughnessTexture), metallicTexture(metallicTexture), alpha(alpha){}
    // Material.hpp:350-353: without an alpha texture or a 4-channel albedo the tester is demoted to Opaque
    void setAlphaTester(AlphaTester tester){
        alphaTester = tester;
        if(alpha == nullptr && tex->Channels() != 4)alphaTester.mode = AlphaMode::Opaque;
    }
protected:
    void Describe(ptb_host::Builder& b, pt_material_desc& d) const override{
        d.kind = PT_MAT_DIFFUSE; d.tex = tex->Emit(b); d.norm = opt(b, norm); d.rough = opt(b, roughnessTexture);
        d.metal = opt(b, metallicTexture); d.alpha = opt(b, alpha); d.alpha_mode = mode(alphaTester); d.alpha_cutoff = alphaTester.cutoff;
    }
private:
    std::shared_ptr<Texture> tex, norm, roughnessTexture, metallicTexture, alpha;
    AlphaTester alphaTester;
};

class MicrofacetDielectric : public Material {
public:
    MicrofacetDielectric(float refIndex, const glm::vec3& albedo) : MicrofacetDielectric(refIndex, std::make_shared<SolidColor>(albedo)){}
    MicrofacetDielectric(float refIndex, float roughness, const glm::vec3& albedo)
        : MicrofacetDielectric(refIndex, std::make_shared<SolidColor>(albedo), nullptr, std::make_shared<SolidColor>(glm::vec3(roughness))){}
    MicrofacetDielectric(float refIndex, const std::shared_ptr<Texture>& tex, const std::shared_ptr<Texture>& norm = nullptr,
                         const std::shared_ptr<Texture>& roughnessTexture = std::make_shared<SolidColor>(glm::vec3(0.0f)),
                         const std::shared_ptr<Texture>& alpha = nullptr)
        : ri(refIndex), tex(tex), norm(norm), roughnessTexture(roughnessTexture), alpha(alpha){}
    void setAlphaTester(AlphaTester tester){
        alphaTester = tester;
        if(alpha == nullptr && tex->Channels() != 4)alphaTester.mode = AlphaMode::Opaque;
    }
protected:
    void Describe(ptb_host::Builder& b, pt_material_desc& d) const override{
        d.kind = PT_MAT_DIELECTRIC; d.ri = ri; d.tex = tex->Emit(b); d.norm = opt(b, norm); d.rough = opt(b, roughnessTexture);
        d.alpha = opt(b, alpha); d.alpha_mode = mode(alphaTester); d.alpha_cutoff = alphaTester.cutoff;
    }
private:
    float ri;
    std::shared_ptr<Texture> tex, norm, roughnessTexture, alpha;
    AlphaTester alphaTester;
};

class ThinDielectric : public Material {
public:
    ThinDielectric(float eta, const std::shared_ptr<Texture>& tex) : ri(eta), albedo(tex != nullptr ? tex : std::make_shared<SolidColor>(glm::vec3(1))){}
protected:
    void Describe(ptb_host::Builder& b, pt_material_desc& d) const override{ d.kind = PT_MAT_THIN_DIELECTRIC; d.ri = ri; d.tex = albedo->Emit(b); }
private:
    float ri;
    std::shared_ptr<Texture> albedo;
};

class SpecularConductor : public Material {
public:
    SpecularConductor(const glm::vec3& albedo) : albedo(albedo){}
protected:
    void Describe(ptb_host::Builder&, pt_material_desc& d) const override{
        d.kind = PT_MAT_CONDUCTOR; d.albedo[0] = albedo.x; d.albedo[1] = albedo.y; d.albedo[2] = albedo.z;
    }
private:
    glm::vec3 albedo;
};

// ---- PhaseFunction.hpp:10-30, Medium.hpp:5-61 ------------------------------------------------------
class PhaseFunction {
public:
    virtual ~PhaseFunction() = default;
    virtual float G() const = 0;
};
class HenyeyGreenstein : public PhaseFunction {
public:
    HenyeyGreenstein(float G) : g(glm::clamp(G, -0.99f, 0.99f)){}
    float G() const override{ return g; }
private:
    float g;
};

class Medium {
public:
    virtual ~Medium() = default;
    virtual int Emit(ptb_host::Builder& b) const = 0;
};
class HomogeneusMedium : public Medium {
public:
    HomogeneusMedium(const glm::vec3& sigma_a, const glm::vec3& sigma_s, const std::shared_ptr<PhaseFunction> phaseFunction, float density = 1.0f,
                     const glm::vec3& Le = glm::vec3 { 0, 0, 0 }, float LeDensity = 1.0f)
        : sigma_a(sigma_a), sigma_s(sigma_s), phase(phaseFunction), density(density), Le(Le), LeDensity(LeDensity){}
    int Emit(ptb_host::Builder& b) const override{
        int id = b.lookup(this);
        if(id >= 0)return id;
        pt_medium_desc d;
        for(int a = 0; a < 3; a++){ d.sigma_a[a] = sigma_a[a]; d.sigma_s[a] = sigma_s[a]; d.Le[a] = Le[a]; }
        d.g = phase ? phase->G() : 0.0f; d.density = density; d.le_density = LeDensity;
        id = pt_scene_add_medium(b.scene, &d);
        ptb_host::check(id, "pt_scene_add_medium");
        b.ids[this] = id;
        return id;
    }
private:
    glm::vec3 sigma_a, sigma_s;
    std::shared_ptr<PhaseFunction> phase;
    float density;
    glm::vec3 Le;
    float LeDensity;
};

// ---- Mesh.hpp:14-95 ---------------------------------------------------------------------------------
// Shape.hpp:60-116: one triangle of a Mesh.  The reference's Mesh owns a vector of them and hands them out through
// GetShape(j) + GetControlPtr() (Mesh.hpp:64-69) so that user code can build GeometricPrimitive lists by hand (the recipe of
// Model::BuildBlas, Model.hpp:47-59).  Defined after Shape below; here only what Mesh needs.
class Mesh;
class TriangleShape;

class Mesh {
public:
    Mesh(const std::vector<uint32_t>& indices, const std::vector<glm::vec3>& vertices, const std::vector<glm::vec3>& tangents,
         const std::vector<glm::vec3>& normals, const std::vector<glm::vec2>& texCoords, const std::shared_ptr<Material>& mat,
         const std::shared_ptr<Texture>& emissiveTex, const std::shared_ptr<Medium>& meshMedium)
        : indices(indices), vertices(vertices), tangents(tangents), normals(normals), texCoords(texCoords), material(mat),
          emissiveTexture(emissiveTex), medium(meshMedium){
        if(normals.size() != vertices.size() || texCoords.size() != vertices.size())throw std::invalid_argument("Mesh: normals and texCoords are per vertex");
        if(!tangents.empty() && tangents.size() != vertices.size())throw std::invalid_argument("Mesh: tangents are per vertex or empty");
    }
    uint32_t GetTriangleCount() const{ return (uint32_t)(indices.size() / 3); }
    uint32_t GetVertexCount() const{ return (uint32_t)vertices.size(); }
    const std::vector<uint32_t>& GetIndices() const{ return indices; }
    const std::vector<glm::vec3>& GetVertices() const{ return vertices; }
    const std::vector<glm::vec3>& GetTangents() const{ return tangents; }
    const std::vector<glm::vec3>& GetNormals() const{ return normals; }
    const std::vector<glm::vec2>& GetTexCoords() const{ return texCoords; }
    std::shared_ptr<Material> GetMaterial() const{ return material; }
    std::shared_ptr<Texture> GetEmissiveTexture() const{ return emissiveTexture; }
    std::shared_ptr<Medium> GetMedium() const{ return medium; }
    // Mesh.hpp:64-69: the triangle shapes of this mesh and the control block that keeps them alive (aliasing shared_ptr recipe)
    inline const std::shared_ptr<std::vector<TriangleShape>>& GetControlPtr() const;
    inline TriangleShape* GetShape(std::size_t index) const;
    // the triangles [first, first + count) as a mesh of their own with this mesh's vertex arrays (BLAS lists, below)
    std::shared_ptr<Mesh> Slice(const std::vector<uint32_t>& tris, const std::shared_ptr<Material>& mat, const std::shared_ptr<Texture>& emissive,
                                const std::shared_ptr<Medium>& med) const{
        std::vector<uint32_t> idx;
        idx.reserve(tris.size() * 3);
        for(uint32_t t : tris){ idx.push_back(indices[3 * t]); idx.push_back(indices[3 * t + 1]); idx.push_back(indices[3 * t + 2]); }
        return std::make_shared<Mesh>(idx, vertices, tangents, normals, texCoords, mat, emissive, med);
    }

    // material / medium overrides as Model::BuildBlas(material, medium) applies them (Model.hpp:62-80)
    int Emit(ptb_host::Builder& b, const std::shared_ptr<Material>& matOverride, const std::shared_ptr<Medium>& medOverride, bool overridden) const{
        static_assert(sizeof(glm::vec3) == 12 && sizeof(glm::vec2) == 8, "glm vectors must be tightly packed");
        pt_mesh_desc d;
        d.positions = &vertices[0].x; d.normals = &normals[0].x; d.texcoords = &texCoords[0].x;
        d.tangents = tangents.empty() ? nullptr : &tangents[0].x;
        d.indices = indices.data();
        d.n_vertices = GetVertexCount(); d.n_triangles = GetTriangleCount();
        const std::shared_ptr<Material>& m = overridden ? matOverride : material;
        const std::shared_ptr<Medium>& md = overridden ? medOverride : medium;
        d.material = m ? m->Emit(b) : -1;
        d.emissive_tex = emissiveTexture ? emissiveTexture->Emit(b) : -1;
        d.medium = md ? md->Emit(b) : -1;
        const int id = pt_scene_add_mesh(b.scene, &d);
        ptb_host::check(id, "pt_scene_add_mesh");
        return id;
    }
private:
    mutable std::shared_ptr<std::vector<TriangleShape>> shapes;       // built on first use
    std::vector<uint32_t> indices;
    std::vector<glm::vec3> vertices, tangents, normals;
    std::vector<glm::vec2> texCoords;
    std::shared_ptr<Material> material;
    std::shared_ptr<Texture> emissiveTexture;
    std::shared_ptr<Medium> medium;
};

// ---- Light.hpp:16-251 -------------------------------------------------------------------------------
// Power functions are std::function<float(float)> in the reference; the three every reference scene
// uses are selectable by value here (pt_light_desc::power_func).
enum class PowerFunction { SqrtR = 0, FourR = 1, FourSqrtR = 2 };

class Light {
public:
    virtual ~Light() = default;
    virtual bool IsSceneLights() const{ return false; }
    virtual void Describe(ptb_host::Builder&, pt_light_desc&) const{}
    int Emit(ptb_host::Builder& b) const{
        pt_light_desc d;
        std::memset(&d, 0, sizeof d);
        d.tex = -1; d.scale = 1.0f;
        Describe(b, d);
        const int id = pt_scene_add_light(b.scene, &d);
        ptb_host::check(id, "pt_scene_add_light");
        return id;
    }
protected:
    static void set3(float* o, const glm::vec3& v){ o[0] = v.x; o[1] = v.y; o[2] = v.z; }
};
class InfiniteLight : public Light {};

class UniformInfiniteLight : public InfiniteLight {
public:
    UniformInfiniteLight(const glm::vec3& light_color, PowerFunction powerFunc = PowerFunction::SqrtR) : color(light_color), pf(powerFunc){}
    void Describe(ptb_host::Builder&, pt_light_desc& d) const override{ d.kind = PT_LIGHT_UNIFORM_INFINITE; set3(d.color, color); d.power_func = (int)pf; }
private:
    glm::vec3 color; PowerFunction pf;
};

// The sky every reference scene passes to FunctionInfiniteLight (main.cpp:108-111):
// scale * ((1-a) * horizon + a * zenith), a = (dir.y + 1) / 2.  Callable, so the same object also
// satisfies the reference's std::function<glm::vec3(const Ray&)> parameter.
struct GradientSky {
    glm::vec3 horizon { 1, 1, 1 }, zenith { 0.5f, 0.7f, 1.0f };
    float scale = 1.0f;
    glm::vec3 operator()(const Ray& ray) const{
        const float a = 0.5f * (ray.dir.y + 1.0f);
        return scale * ((1.0f - a) * horizon + a * zenith);
    }
};
// FunctionInfiniteLight (Light.hpp:70-93, Light.cpp:53-108) holds a host closure.  The GradientSky functor is evaluated on
// the device in closed form.  Any other std::function<glm::vec3(const Ray&)> cannot run on the GPU: it is BAKED, once, into a
// lat-long float table (kBakeW x kBakeH texels at the centres of GetSphereUV cells) that the device reads bilinearly; sampling
// stays the reference's (uniform on the sphere, pdf 1 / 4 pi), Power() is the reference's mean luminance over the sphere.
// BakeError() reports what the table costs: the largest and the mean absolute difference between the function and the
// table's bilinear value at 20 000 directions, relative to the function's mean luminance.
class FunctionInfiniteLight : public InfiniteLight {
public:
    static constexpr int kBakeW = 1024, kBakeH = 512;
    struct Error { double max_rel = 0, mean_rel = 0; };
    FunctionInfiniteLight(const GradientSky& lightFunc, PowerFunction powerFunc = PowerFunction::SqrtR) : sky(lightFunc), pf(powerFunc){}
    FunctionInfiniteLight(const std::function<glm::vec3(const Ray&)>& lightFunc, PowerFunction powerFunc = PowerFunction::SqrtR) : pf(powerFunc){ Bake(lightFunc); }
    void Describe(ptb_host::Builder& b, pt_light_desc& d) const override{
        d.kind = PT_LIGHT_GRADIENT_INFINITE; d.power_func = (int)pf;
        if(table){ d.tex = table->Emit(b); d.scale = 1.0f; }
        else{ set3(d.color, sky.horizon); set3(d.color_b, sky.zenith); d.scale = sky.scale; }
    }
    bool Baked() const{ return table != nullptr; }
    Error BakeError() const{ return err; }
private:
    static glm::vec3 Dir(double u, double v){          // inverse of SphereShape::GetSphereUV (Shape.hpp): u = phi / 2 pi, v = theta / pi, y up
        const double phi = 2.0 * 3.14159265358979323846 * u, theta = 3.14159265358979323846 * v;
        return glm::vec3((float)(std::sin(theta) * std::cos(phi)), (float)std::cos(theta), (float)(std::sin(theta) * std::sin(phi)));
    }
    void Bake(const std::function<glm::vec3(const Ray&)>& f){
        std::vector<float> tex((size_t)kBakeW * kBakeH * 3);
        for(int y = 0; y < kBakeH; y++)for(int x = 0; x < kBakeW; x++){
            const glm::vec3 c = f(Ray(glm::vec3(0, 0, 0), Dir((x + 0.5) / kBakeW, (y + 0.5) / kBakeH)));
            float* t = &tex[((size_t)y * kBakeW + x) * 3];
            t[0] = c.x; t[1] = c.y; t[2] = c.z;
        }
        // error of the bilinear read-back (repeat wrap in u, clamp in v as the device texture fetch does near the poles)
        auto fetch = [&](double u, double v, int ch){
            const double fx = u * kBakeW - 0.5, fy = v * kBakeH - 0.5;
            const int x0 = (int)std::floor(fx), y0 = (int)std::floor(fy);
            const double ax = fx - x0, ay = fy - y0;
            auto at = [&](int x, int y){ x = ((x % kBakeW) + kBakeW) % kBakeW; y = ((y % kBakeH) + kBakeH) % kBakeH; return (double)tex[((size_t)y * kBakeW + x) * 3 + ch]; };
            return (1 - ay) * ((1 - ax) * at(x0, y0) + ax * at(x0 + 1, y0)) + ay * ((1 - ax) * at(x0, y0 + 1) + ax * at(x0 + 1, y0 + 1));
        };
        double lum = 0;
        for(size_t i = 0; i < tex.size(); i += 3)lum += 0.2126 * tex[i] + 0.7152 * tex[i + 1] + 0.0722 * tex[i + 2];
        lum = std::max(lum / ((double)kBakeW * kBakeH), 1e-30);
        uint64_t st = 0x9E3779B97F4A7C15ull;
        auto rnd = [&](){ st ^= st << 13; st ^= st >> 7; st ^= st << 17; return (double)(st >> 11) / 9007199254740992.0; };
        double worst = 0, sum = 0; const int n = 20000;
        for(int k = 0; k < n; k++){
            const double u = rnd(), v = std::acos(1.0 - 2.0 * rnd()) / 3.14159265358979323846;
            const glm::vec3 c = f(Ray(glm::vec3(0, 0, 0), Dir(u, v)));
            double e = 0;
            for(int ch = 0; ch < 3; ch++)e = std::max(e, std::fabs(fetch(u, v, ch) - (double)c[ch]));
            worst = std::max(worst, e / lum); sum += e / lum;
        }
        err.max_rel = worst; err.mean_rel = sum / n;
        table = std::make_shared<FloatImageTexture>(std::move(tex), kBakeW, kBakeH, 3);
    }
    GradientSky sky; PowerFunction pf;
    std::shared_ptr<FloatImageTexture> table;
    Error err;
};

class TextureInfiniteLight : public InfiniteLight {
public:
    TextureInfiniteLight(const std::shared_ptr<Texture>& tex, float LeScale = 1.0f, PowerFunction powerFunc = PowerFunction::SqrtR) : tex(tex), LeScale(LeScale), pf(powerFunc){}
    void Describe(ptb_host::Builder& b, pt_light_desc& d) const override{ d.kind = PT_LIGHT_TEXTURE_INFINITE; d.tex = tex->Emit(b); d.scale = LeScale; d.power_func = (int)pf; }
private:
    std::shared_ptr<Texture> tex; float LeScale; PowerFunction pf;
};

class DistantLight : public Light {
public:
    DistantLight(const glm::vec3& light_dir, const glm::vec3& light_color, PowerFunction powerFunc = PowerFunction::SqrtR) : dir(light_dir), color(light_color), pf(powerFunc){}
    void Describe(ptb_host::Builder&, pt_light_desc& d) const override{ d.kind = PT_LIGHT_DISTANT; set3(d.dir_or_pos, dir); set3(d.color, color); d.power_func = (int)pf; }
private:
    glm::vec3 dir, color; PowerFunction pf;
};

class PointLight : public Light {
public:
    PointLight(const glm::vec3& p, const glm::vec3& light_color, PowerFunction powerFunc = PowerFunction::FourR) : p(p), color(light_color), pf(powerFunc){}
    void Describe(ptb_host::Builder&, pt_light_desc& d) const override{ d.kind = PT_LIGHT_POINT; set3(d.dir_or_pos, p); set3(d.color, color); d.power_func = (int)pf; }
private:
    glm::vec3 p, color; PowerFunction pf;
};

// ---- Shape.hpp:8-175: the analytic shapes a GeometricPrimitive can carry into the TLAS ---------------------
class Shape {
public:
    virtual ~Shape() = default;
    virtual void Describe(pt_shape_desc& d) const = 0;
    virtual AABB BoundingBox() const = 0;
    virtual float Area() const = 0;
};

class SphereShape : public Shape {
public:
    SphereShape(const glm::vec3& sphereCenter, float sphereRadius) : center(sphereCenter), radius(sphereRadius){}
    void Describe(pt_shape_desc& d) const override{
        d.kind = PT_SHAPE_SPHERE; d.radius = radius;
        for(int a = 0; a < 3; a++)d.p[a] = center[a];
    }
    AABB BoundingBox() const override{ AABB b; b.Expand(center - glm::vec3(radius)); b.Expand(center + glm::vec3(radius)); return b; }
    float Area() const override{ return 4.0f * 3.14159265358979323846f * radius * radius; }
private:
    glm::vec3 center; float radius;
};

class QuadShape : public Shape {
public:
    QuadShape(const glm::vec3& origin, const glm::vec3& u, const glm::vec3& v) : Q(origin), u(u), v(v){}
    void Describe(pt_shape_desc& d) const override{
        d.kind = PT_SHAPE_QUAD;
        for(int a = 0; a < 3; a++){ d.p[a] = Q[a]; d.u[a] = u[a]; d.v[a] = v[a]; }
    }
    AABB BoundingBox() const override{ AABB b; b.Expand(Q); b.Expand(Q + u + v); b.Expand(Q + u); b.Expand(Q + v); return b; }
    float Area() const override{ return glm::length(glm::cross(u, v)); }
private:
    glm::vec3 Q, u, v;
};

// Light.hpp:179-206.  An AreaLight reaches the light sampler through Scene::GetLights() (the
// GeometricPrimitive that carries it), exactly as in the reference; it cannot be added on its own.
class TriangleShape : public Shape {
public:
    TriangleShape(const Mesh* mesh, uint32_t triangleIndex) : mesh(mesh), tri(triangleIndex){}
    void Describe(pt_shape_desc&) const override{ throw std::logic_error("TriangleShape: triangles reach the GPU through a BLAS (BLAS2/4/8 over a GeometricPrimitive list, or Model::BuildBlas)"); }
    AABB BoundingBox() const override{
        AABB b;
        for(int k = 0; k < 3; k++)b.Expand(mesh->GetVertices()[mesh->GetIndices()[3 * tri + k]]);
        return b;
    }
    float Area() const override{
        const glm::vec3 a = mesh->GetVertices()[mesh->GetIndices()[3 * tri]], e1 = mesh->GetVertices()[mesh->GetIndices()[3 * tri + 1]] - a, e2 = mesh->GetVertices()[mesh->GetIndices()[3 * tri + 2]] - a;
        return 0.5f * glm::length(glm::cross(e1, e2));
    }
    const Mesh* GetMesh() const{ return mesh; }
    uint32_t Index() const{ return tri; }
private:
    const Mesh* mesh;
    uint32_t tri;
};
inline const std::shared_ptr<std::vector<TriangleShape>>& Mesh::GetControlPtr() const{
    if(!shapes){
        shapes = std::make_shared<std::vector<TriangleShape>>();
        shapes->reserve(GetTriangleCount());
        for(uint32_t j = 0; j < GetTriangleCount(); j++)shapes->emplace_back(this, j);
    }
    return shapes;
}
inline TriangleShape* Mesh::GetShape(std::size_t index) const{ return &(*GetControlPtr())[index]; }

class AreaLight : public Light {
public:
    AreaLight(const std::shared_ptr<Shape>& light_shape, const glm::vec3& light_color, bool oneSided = false)
        : shape(light_shape), emissiveTexture(std::make_shared<SolidColor>(light_color)), oneSided(oneSided){}
    AreaLight(const std::shared_ptr<Shape>& light_shape, const std::shared_ptr<Texture>& emissiveTex, bool oneSided = false)
        : shape(light_shape), emissiveTexture(emissiveTex), oneSided(oneSided){}
    void Describe(ptb_host::Builder&, pt_light_desc&) const override{
        throw std::logic_error("AreaLight: hand it to a GeometricPrimitive and add scene->GetLights() to the light sampler");
    }
    std::shared_ptr<Shape> getShape() const{ return shape; }
    // the recipe of Model::BuildBlas calls these per triangle (Model.hpp:52-55); powers are computed at commit on this path
    // (csrc/shade_tables.cpp, incl. the Power <= epsilon drop), so here they only keep the recipe compiling
    void PreProcess(const AABB&){}
    float Power() const{ return 1.0f; }
    const std::shared_ptr<Texture>& EmissiveTexture() const{ return emissiveTexture; }
    bool OneSided() const{ return oneSided; }
private:
    std::shared_ptr<Shape> shape;
    std::shared_ptr<Texture> emissiveTexture;
    bool oneSided;
};

// What Scene::GetLights() hands to LightSampler::Add: "every area light of every emissive mesh (through
// its instance transform) plus the scene's infinite lights".  The GPU scene derives that list itself at
// commit (csrc/shade_tables.cpp build_lights), in the reference's order.
class SceneLights : public Light {
public:
    bool IsSceneLights() const override{ return true; }
};

// ---- Primitive.hpp:7-50, Model.hpp:14-80, BVH.hpp:554-555,1219-1220,1550-1551 -----------------------------------
struct TLAS2 { static constexpr int width = 2; };
struct TLAS4 { static constexpr int width = 4; };
struct TLAS8 { static constexpr int width = 8; };

class Primitive {
public:
    virtual ~Primitive() = default;
    // adds this primitive to the scene being assembled and returns its top-level id; transform = column-major mat4 or null
    virtual int Lower(ptb_host::Builder& b, const glm::mat4* transform) const = 0;
    virtual int RequiredWidth() const{ return 0; }     // BLAS width this primitive was built with (0: any)
};

// Primitive.hpp:17-33: shape + material + area light + medium, added to the Scene directly (examples/example_1.cpp:52-66).
// material == nullptr: a medium boundary.  The area light must be built over the same shape.
class GeometricPrimitive : public Primitive {
public:
    const std::shared_ptr<Shape>& GetShape() const{ return shape; }
    const std::shared_ptr<Material>& GetMaterial() const{ return material; }
    const std::shared_ptr<AreaLight>& GetAreaLight() const{ return areaLight; }
    const std::shared_ptr<Medium>& GetMedium() const{ return medium; }
    GeometricPrimitive(const std::shared_ptr<Shape>& shape, const std::shared_ptr<Material>& material, const std::shared_ptr<AreaLight>& areaLight = nullptr,
                       const std::shared_ptr<Medium>& medium = nullptr)
        : shape(shape), material(material), areaLight(areaLight), medium(medium){
        if(!shape)throw std::invalid_argument("GeometricPrimitive: null shape");
        if(areaLight && areaLight->getShape() != shape)throw std::invalid_argument("GeometricPrimitive: the AreaLight must be built over the primitive's own shape");
    }
    int Lower(ptb_host::Builder& b, const glm::mat4* transform) const override{
        pt_shape_desc d;
        std::memset(&d, 0, sizeof d);
        shape->Describe(d);
        d.material = material ? material->Emit(b) : -1;
        d.medium = medium ? medium->Emit(b) : -1;
        d.emissive_tex = areaLight ? areaLight->EmissiveTexture()->Emit(b) : -1;
        d.one_sided = areaLight && areaLight->OneSided() ? 1 : 0;
        // inside a TransformedPrimitive (Primitive.cpp:32-73) the shape keeps its own space and the wrapper's matrix travels with it
        const int id = transform ? pt_scene_add_shape_xf(b.scene, &d, &(*transform)[0][0]) : pt_scene_add_shape(b.scene, &d);
        ptb_host::check(id, "pt_scene_add_shape");
        return id;
    }
private:
    std::shared_ptr<Shape> shape;
    std::shared_ptr<Material> material;
    std::shared_ptr<AreaLight> areaLight;
    std::shared_ptr<Medium> medium;
};

// A bottom-level structure over meshes.  The reference's Model imports a file through assimp
// (Model.cpp:22-357, host I/O outside the hot path); here the meshes are handed over directly — the
// constructor the reference keeps commented out at Model.hpp:18.
// BVH.hpp:554-555,743,1219-1220,1550-1551: BLASx = BVHx<GeometricPrimitive>.  As a tag (BLAS4::width) it selects the node width in
// Model::BuildBlas<BLAS4>() / Scene::BuildTlas<TLAS4>(); constructed from a hand-built primitive list -- the public constructor
// BVH4(const std::vector<GeometricPrimitive>&), used with the recipe of Model.hpp:47-59 -- it is a Primitive the Scene can hold.
// The list must consist of TriangleShape primitives (what Mesh::GetShape hands out); runs of consecutive triangles that share
// mesh, material, emitter and medium become one sub-mesh each, in list order, so primitive ids are list positions as in the
// reference.  Analytic shapes go into the Scene directly (GeometricPrimitive), as in examples/example_1.cpp.
template <int W>
class BLASx : public Primitive {
public:
    static constexpr int width = W;
    BLASx() = default;
    BLASx(const std::vector<GeometricPrimitive>& prims){
        size_t i = 0;
        while(i < prims.size()){
            const TriangleShape* t0 = dynamic_cast<const TriangleShape*>(prims[i].GetShape().get());
            if(!t0)throw std::invalid_argument("BLAS: only TriangleShape primitives can be put into a BLAS on this path (spheres / quads go into the Scene directly)");
            const std::shared_ptr<Texture> em = prims[i].GetAreaLight() ? prims[i].GetAreaLight()->EmissiveTexture() : nullptr;
            std::vector<uint32_t> tris;
            size_t j = i;
            for(; j < prims.size(); j++){
                const TriangleShape* t = dynamic_cast<const TriangleShape*>(prims[j].GetShape().get());
                const std::shared_ptr<Texture> ej = prims[j].GetAreaLight() ? prims[j].GetAreaLight()->EmissiveTexture() : nullptr;
                if(!t || t->GetMesh() != t0->GetMesh() || prims[j].GetMaterial() != prims[i].GetMaterial() || prims[j].GetMedium() != prims[i].GetMedium() || ej != em)break;
                tris.push_back(t->Index());
            }
            meshes.push_back(t0->GetMesh()->Slice(tris, prims[i].GetMaterial(), em, prims[i].GetMedium()));
            i = j;
        }
        if(meshes.empty())throw std::invalid_argument("BLAS: empty primitive list");
    }
    int RequiredWidth() const override{ return W; }
    int Lower(ptb_host::Builder& b, const glm::mat4* transform) const override{
        if(meshes.empty())throw std::logic_error("BLAS: default-constructed (a width tag), nothing to add to a Scene");
        int blas = b.lookup(this);
        if(blas < 0){
            std::vector<int32_t> ids;
            for(const std::shared_ptr<Mesh>& m : meshes)ids.push_back(m->Emit(b, nullptr, nullptr, false));
            blas = pt_scene_add_blas(b.scene, ids.data(), (int)ids.size());
            ptb_host::check(blas, "pt_scene_add_blas");
            b.ids[this] = blas;
        }
        const int id = pt_scene_add_instance(b.scene, blas, transform ? &(*transform)[0][0] : nullptr);
        ptb_host::check(id, "pt_scene_add_instance");
        return id;
    }
    const std::vector<std::shared_ptr<Mesh>>& GetMeshes() const{ return meshes; }
private:
    std::vector<std::shared_ptr<Mesh>> meshes;
};
using BLAS2 = BLASx<2>;
using BLAS4 = BLASx<4>;
using BLAS8 = BLASx<8>;

class Model : public Primitive {
public:
    Model(std::vector<std::shared_ptr<Mesh>> meshes) : meshes(std::move(meshes)){}
    // Model(path) (Model.cpp:14-18): the reference imports through assimp; here Wavefront OBJ / MTL (ObjImport.hpp) and glTF 2.0 (GltfImport.hpp).
    // Throws where the reference prints "Failed to load model" and carries on with an empty model.
    explicit Model(const std::string& path);
    std::vector<std::shared_ptr<Mesh>> GetMeshes() const{ return meshes; }
    template <typename T> void BuildBlas(){ width = T::width; overridden = false; }
    template <typename T> void BuildBlas(const std::shared_ptr<Material>& material, const std::shared_ptr<Medium>& medium){
        width = T::width; overridden = true; matOverride = material; medOverride = medium;
    }
    int RequiredWidth() const override{ return width; }
    int Lower(ptb_host::Builder& b, const glm::mat4* transform) const override{
        if(width == 0)throw std::logic_error("Model::BuildBlas<BLAS2|BLAS4|BLAS8>() was not called");
        int blas = b.lookup(this);
        if(blas < 0){
            std::vector<int32_t> ids;
            for(const std::shared_ptr<Mesh>& m : meshes)ids.push_back(m->Emit(b, matOverride, medOverride, overridden));
            blas = pt_scene_add_blas(b.scene, ids.data(), (int)ids.size());
            ptb_host::check(blas, "pt_scene_add_blas");
            b.ids[this] = blas;
        }
        const int id = pt_scene_add_instance(b.scene, blas, transform ? &(*transform)[0][0] : nullptr);
        ptb_host::check(id, "pt_scene_add_instance");
        return id;
    }
private:
    std::vector<std::shared_ptr<Mesh>> meshes;
    int width = 0;
    bool overridden = false;
    std::shared_ptr<Material> matOverride;
    std::shared_ptr<Medium> medOverride;
};

class TransformedPrimitive : public Primitive {
public:
    TransformedPrimitive(const std::shared_ptr<Primitive>& primitive, const glm::mat4& transform) : primitive(primitive), transform(transform){}
    int RequiredWidth() const override{ return primitive->RequiredWidth(); }
    int Lower(ptb_host::Builder& b, const glm::mat4* outer) const override{
        // Nested wrappers (Primitive.hpp:37 takes any Primitive): the reference applies the levels one after the other, the GPU TLAS
        // holds one transform per instance, so the matrices are multiplied here.  Same geometry; the per-level renormalisation of the
        // ray direction is not associative in floating point, so hit distances of a NESTED wrapper agree with the reference to
        // rounding, not bit for bit (a single level stays bit-exact).
        if(outer){ const glm::mat4 m = *outer * transform; return primitive->Lower(b, &m); }
        return primitive->Lower(b, &transform);
    }
private:
    std::shared_ptr<Primitive> primitive;
    glm::mat4 transform;
};

// Primitive.hpp:52-66, Primitive.cpp:77-96 (main.cpp:187-190): motion blur.  Per ray the child is seen under
// translate(direction * t), t = clamp(ray.time - timeBounds.x, timeBounds.x, timeBounds.y) / (timeBounds.y -
// timeBounds.x); its lights become AnimatedLights.  The child may be a Model, a GeometricPrimitive or a
// TransformedPrimitive over either.
class AnimatedPrimitive : public Primitive {
public:
    AnimatedPrimitive(const std::shared_ptr<Primitive>& primitive, const glm::vec3& direction, const glm::vec2& timeBounds)
        : primitive(primitive), dir(direction), timeBounds(timeBounds){}
    int RequiredWidth() const override{ return primitive->RequiredWidth(); }
    int Lower(ptb_host::Builder& b, const glm::mat4* outer) const override{
        // TransformedPrimitive(AnimatedPrimitive(x, dir), M) sees x under M then translate(dir t): the same rays as
        // AnimatedPrimitive(TransformedPrimitive(x, M), M_linear dir), the chain the device understands
        const int id = primitive->Lower(b, outer);
        glm::vec3 dw = dir;
        if(outer){ const glm::vec4 q = *outer * glm::vec4(dir.x, dir.y, dir.z, 0.0f); dw = glm::vec3(q.x, q.y, q.z); }
        const float d[3] = { dw.x, dw.y, dw.z }, tb[2] = { timeBounds.x, timeBounds.y };
        ptb_host::check(pt_scene_animate(b.scene, id, d, tb), "pt_scene_animate");
        return id;
    }
private:
    std::shared_ptr<Primitive> primitive;
    glm::vec3 dir;
    glm::vec2 timeBounds;
};

// ---- LightSampler.hpp:5-60 ----------------------------------------------------------------------------
class LightSampler {
public:
    virtual ~LightSampler() = default;
    virtual int Kind() const = 0;                      // pt_scene_commit's light_sampler argument
    void Add(const std::shared_ptr<Light>& light){
        if(light->IsSceneLights())sceneLights = true;
        else lights.push_back(light);
    }
    void Add(const std::vector<std::shared_ptr<Light>>& ls){ for(const std::shared_ptr<Light>& l : ls)Add(l); }
    // powers, culling (< 0.01) and the cdf are computed at commit (csrc/shade_tables.cpp)
    void PreProcess(const AABB&){}
    bool HasSceneLights() const{ return sceneLights; }
    const std::vector<std::shared_ptr<Light>>& ExtraLights() const{ return lights; }
private:
    std::vector<std::shared_ptr<Light>> lights;       // lights the user added on top of Scene::GetLights()
    bool sceneLights = false;
};
class UniformLightSampler : public LightSampler { public: int Kind() const override{ return 0; } };
class PowerLightSampler : public LightSampler { public: int Kind() const override{ return 1; } };

// ---- Scene.hpp:5-40 -------------------------------------------------------------------------------------
class Scene {
public:
    Scene() = default;
    Scene(const std::shared_ptr<Medium>& medium) : sceneMedium(medium){}
    Scene(const Scene&) = delete;
    Scene& operator=(const Scene&) = delete;
    ~Scene(){ DropGpu(); }

    void Add(const std::shared_ptr<Primitive>& ptr){ primitives.push_back(ptr); Invalidate(); }
    std::vector<std::shared_ptr<Light>> GetLights() const{ return { std::make_shared<SceneLights>() }; }
    template <typename T> void BuildTlas(){ tlasWidth = T::width; Invalidate(); }
    std::shared_ptr<Medium> GetMedium() const{ return sceneMedium; }
    void SetMedium(const std::shared_ptr<Medium>& medium){ sceneMedium = medium; Invalidate(); }
    void SetDevice(int ordinal){ device = ordinal; Invalidate(); }

    AABB BoundingBox(){
        AABB b;
        ptb_host::check(pt_scene_bounds(Commit(nullptr), &b.min.x, &b.max.x), "pt_scene_bounds");
        return b;
    }

    // one ray through the GPU (a kernel launch per call: use the batch forms for throughput)
    bool Intersect(const Ray& ray, SurfaceInteraction& interaction, float max = std::numeric_limits<float>::infinity()){
        const pt_ray r = { ray.origin.x, ray.origin.y, ray.origin.z, max, ray.dir.x, ray.dir.y, ray.dir.z, ray.time };
        pt_hit h; uint32_t inst = PT_MISS;
        IntersectBatch(&r, 1, &h, &inst);
        if(h.prim == PT_MISS)return false;
        interaction.t = h.t; interaction.p = ray.at(h.t); interaction.barycentric = { h.u, h.v };
        interaction.primitive = h.prim; interaction.instance = inst;
        return true;
    }
    bool IntersectPred(const Ray& ray, float max = std::numeric_limits<float>::infinity()){
        const pt_ray r = { ray.origin.x, ray.origin.y, ray.origin.z, max, ray.dir.x, ray.dir.y, ray.dir.z, ray.time };
        uint8_t occ = 0;
        IntersectPredBatch(&r, 1, &occ);
        return occ != 0;
    }
    void IntersectBatch(const pt_ray* rays, size_t n, pt_hit* hits, uint32_t* instances = nullptr){
        ptb_host::check(pt_trace_closest(Commit(nullptr), rays, n, hits, instances), "pt_trace_closest");
    }
    void IntersectPredBatch(const pt_ray* rays, size_t n, uint8_t* occluded){
        ptb_host::check(pt_trace_any(Commit(nullptr), rays, n, occluded), "pt_trace_any");
    }

    std::vector<std::shared_ptr<InfiniteLight>> infiniteLights;

    // Lowers the object graph into a pt_scene and commits it; cached until the graph or the light
    // sampler changes.  sampler == nullptr: ray casting only (PowerLightSampler tables are built anyway).
    pt_scene* Commit(const LightSampler* sampler, const std::shared_ptr<Medium>& cameraMedium = nullptr){
        if(tlasWidth == 0)throw std::logic_error("Scene::BuildTlas<TLAS2|TLAS4|TLAS8>() was not called");
        const int kind = sampler ? sampler->Kind() : 1;
        std::vector<const void*> key;
        for(const auto& l : infiniteLights)key.push_back(l.get());
        if(sampler)for(const auto& l : sampler->ExtraLights())key.push_back(l.get());
        if(sampler && !sampler->HasSceneLights())throw std::logic_error("LightSampler::Add(scene->GetLights()) was not called");
        const bool newMedium = cameraMedium && MediumId(cameraMedium.get()) < 0;
        if(gpu && !newMedium && (sampler == nullptr || (kind == committedKind && key == committedKey)))return gpu;
        DropGpu();
        ptb_host::Builder b;
        b.scene = pt_scene_create(device);
        if(!b.scene)throw std::runtime_error(std::string("pt_scene_create: ") + pt_last_error());
        try{
            for(const std::shared_ptr<Primitive>& p : primitives){
                const int w = p->RequiredWidth();
                if(w != 0 && w != tlasWidth)throw std::logic_error("BLAS and TLAS widths must match (BLAS2+TLAS2, BLAS4+TLAS4 or BLAS8+TLAS8)");
                p->Lower(b, nullptr);
            }
            for(const auto& l : infiniteLights)l->Emit(b);
            if(sampler)for(const auto& l : sampler->ExtraLights())l->Emit(b);
            if(cameraMedium)cameraMedium->Emit(b);
            if(sceneMedium)ptb_host::check(pt_scene_set_medium(b.scene, sceneMedium->Emit(b)), "pt_scene_set_medium");
            // PTB_LAYOUT=fast: BVH4 scenes on the quantised 64 B clusters (same hits up to exact ties; 8-12 % faster frames on large scenes,
            // DESIGN.md 4.1).  The default stays the reference-exact layout.
            if(const char* e = std::getenv("PTB_LAYOUT"))
                if(tlasWidth == 4 && std::string(e) == "fast")ptb_host::check(pt_scene_set_layout(b.scene, PT_LAYOUT_FAST), "pt_scene_set_layout");
            ptb_host::check(pt_scene_commit(b.scene, tlasWidth, kind), "pt_scene_commit");
        } catch(...){
            pt_scene_destroy(b.scene);
            throw;
        }
        gpu = b.scene; committedKind = kind; committedKey = key; media = b.ids;
        return gpu;
    }
    // id of a medium object in the committed scene (camera medium), -1 if it is not part of it
    int MediumId(const Medium* m) const{ auto it = media.find(m); return it == media.end() ? -1 : it->second; }
    // The committed scene on `gpus` GPUs (the primary's device first, then the other visible ones): replicas, an NCCL
    // communicator and one host thread per GPU inside libptb200.so (pt_multi_create), cached like the scene itself.
    // gpus <= 1 or a single visible GPU: nullptr (Render then calls pt_render on the primary).
    pt_multi* Multi(pt_scene* committed, int gpus){
        const int visible = pt_device_count();
        if(gpus > visible)gpus = visible;
        if(gpus <= 1)return nullptr;
        if(multi && multiGpus == gpus)return multi;
        if(multi){ pt_multi_destroy(multi); multi = nullptr; }
        std::vector<int> devs { device };
        for(int d = 0; d < visible && (int)devs.size() < gpus; d++)if(d != device)devs.push_back(d);
        multi = pt_multi_create(committed, devs.data(), (int)devs.size());
        if(!multi)throw std::runtime_error(std::string("pt_multi_create: ") + pt_last_error());
        multiGpus = gpus;
        return multi;
    }
private:
    void DropGpu(){
        if(multi){ pt_multi_destroy(multi); multi = nullptr; }       // before the primary it refers to
        if(gpu){ pt_scene_destroy(gpu); gpu = nullptr; }
    }
    void Invalidate(){ DropGpu(); }
    pt_multi* multi = nullptr;
    int multiGpus = 0;
    std::vector<std::shared_ptr<Primitive>> primitives;
    std::shared_ptr<Medium> sceneMedium;
    int tlasWidth = 0, device = 0;
    pt_scene* gpu = nullptr;
    int committedKind = -1;
    std::vector<const void*> committedKey;
    std::unordered_map<const void*, int> media;
};

// ---- Filter.hpp:29-144 ------------------------------------------------------------------------------------
class Filter {
public:
    virtual ~Filter() = default;
    virtual glm::vec2 Radius() const = 0;
    virtual void Describe(pt_render_params& rp) const = 0;
protected:
    static void radius(pt_render_params& rp, const glm::vec2& r){ rp.filter_radius[0] = r.x; rp.filter_radius[1] = r.y; }
};
class BoxFilter : public Filter {
public:
    BoxFilter(const glm::vec2& filterRadius = glm::vec2 { 0.5f }) : r(filterRadius){}
    glm::vec2 Radius() const override{ return r; }
    void Describe(pt_render_params& rp) const override{ rp.filter = PT_FILTER_BOX; radius(rp, r); }
private:
    glm::vec2 r;
};
class GaussianFilter : public Filter {
public:
    GaussianFilter(const glm::vec2& filterRadius = glm::vec2 { 1.5f }, double sigma = 0.5) : r(filterRadius), sigma(sigma){}
    glm::vec2 Radius() const override{ return r; }
    void Describe(pt_render_params& rp) const override{ rp.filter = PT_FILTER_GAUSSIAN; radius(rp, r); rp.filter_param[0] = (float)sigma; }
private:
    glm::vec2 r; double sigma;
};
class MitchellFilter : public Filter {
public:
    MitchellFilter(const glm::vec2& radius = glm::vec2 { 1.5f }, double b = 1.0 / 3.0, double c = 1.0 / 3.0) : r(radius), b(b), c(c){}
    glm::vec2 Radius() const override{ return r; }
    void Describe(pt_render_params& rp) const override{ rp.filter = PT_FILTER_MITCHELL; radius(rp, r); rp.filter_param[0] = (float)b; rp.filter_param[1] = (float)c; }
private:
    glm::vec2 r; double b, c;
};
class LanczosFilter : public Filter {
public:
    LanczosFilter(const glm::vec2& filterRadius = glm::vec2 { 1.5f }, double tau = 3) : r(filterRadius), tau(tau){}
    glm::vec2 Radius() const override{ return r; }
    void Describe(pt_render_params& rp) const override{ rp.filter = PT_FILTER_LANCZOS; radius(rp, r); rp.filter_param[0] = (float)tau; }
private:
    glm::vec2 r; double tau;
};

// ---- Film.hpp:22-47,112-225 ---------------------------------------------------------------------------------
inline double luminance(const glm::dvec3& v){ return 0.2126 * v.x + 0.7152 * v.y + 0.0722 * v.z; }
// Reinhard-Jodie: blend of luminance-based and per-channel Reinhard, weighted by the per-channel result
inline glm::vec3 reinhard_jodie(glm::vec3 c){
    const glm::dvec3 color(c);
    const double l = luminance(color);
    glm::dvec3 out;
    for(int a = 0; a < 3; a++){
        const double t = color[a] / (1.0 + color[a]);
        out[a] = (color[a] / (1.0 + l)) * (1.0 - t) + t * t;
    }
    return glm::vec3(out);
}
inline glm::vec3 ACESFilm(glm::vec3 c){
    glm::vec3 out;
    for(int a = 0; a < 3; a++){
        const double x = c[a];
        out[a] = (float)glm::clamp((x * (2.51 * x + 0.03)) / (x * (2.43 * x + 0.59) + 0.14), 0.0, 1.0);
    }
    return out;
}
inline double linear_to_sRGB(double linear){
    linear = glm::clamp(linear, 0.0, 1.0);
    return linear < 0.0031308 ? 12.92 * linear : 1.055 * std::pow(linear, 1.0 / 2.4) - 0.055;
}

class Film {
public:
    Film(const glm::ivec2& resolution, const std::shared_ptr<Filter>& filter, double maxComponentValue = std::numeric_limits<double>::infinity())
        : xResolution(resolution.x), yResolution(resolution.y), screen(4 * (size_t)resolution.x * resolution.y, 0.0), filter(filter), maxComponent(maxComponentValue){}
    // stored and never read, exactly as in the reference (Film.hpp:114,270: no sample is ever clamped there)
    double MaxComponentValue() const{ return maxComponent; }
    void Clear(){ screen.assign(screen.size(), 0.0); }
    glm::ivec2 Resolution() const{ return { xResolution, yResolution }; }
    const std::shared_ptr<Filter>& GetFilter() const{ return filter; }

    // Film::Merge (Film.hpp:125-132) for a whole-frame accumulator downloaded from the device
    void Merge(const std::vector<double>& rgbw){
        if(rgbw.size() != screen.size())throw std::invalid_argument("Film::Merge: size");
        for(size_t i = 0; i < screen.size(); i++)screen[i] += rgbw[i];
    }
    // sum(RGB * w) / sum(w) (Film.hpp:161), row 0 = bottom row as the reference stores it
    glm::dvec3 Pixel(int x, int y) const{
        const double* q = &screen[4 * ((size_t)y * xResolution + x)];
        return glm::dvec3(q[0], q[1], q[2]) * (1.0 / q[3]);
    }
    const std::vector<double>& Accumulators() const{ return screen; }

    void setOutputPath(const std::string& path){ outputPath = path; }
    // binary PPM, top row first, tone-mapped then sRGB-encoded (Film.hpp:154-170)
    void WritePPM(const std::string& filename, const std::function<glm::vec3(glm::vec3)>& toneMapper = reinhard_jodie) const{
        std::ofstream out((outputPath.empty() ? filename : outputPath) + ".ppm", std::ios::binary);
        if(!out)throw std::runtime_error("Film::WritePPM: cannot open output file");
        out << "P6\n" << xResolution << " " << yResolution << "\n255\n";
        for(int i = yResolution - 1; i >= 0; i--)for(int j = 0; j < xResolution; j++){
            const glm::vec3 c = toneMapper(glm::vec3(Pixel(j, i)));
            for(int a = 0; a < 3; a++)out << (char)(255.999 * std::max(0.0, std::min(1.0, linear_to_sRGB((double)c[a]))));
        }
    }
    // PNG, 8-bit RGB, top row first (Film.hpp:172-194).  The image data goes into stored (uncompressed)
    // deflate blocks: a valid PNG without a compressor.
    void WritePNG(const std::string& filename, const std::function<glm::vec3(glm::vec3)>& toneMapper = reinhard_jodie) const{
        std::vector<uint8_t> raw;
        raw.reserve((size_t)yResolution * (3 * (size_t)xResolution + 1));
        for(int i = yResolution - 1; i >= 0; i--){
            raw.push_back(0);                                       // filter type: none
            for(int j = 0; j < xResolution; j++){
                const glm::vec3 c = toneMapper(glm::vec3(Pixel(j, i)));
                for(int a = 0; a < 3; a++)raw.push_back((uint8_t)(255.999 * std::max(0.0, std::min(1.0, linear_to_sRGB((double)c[a])))));
            }
        }
        std::vector<uint8_t> z = { 0x78, 0x01 };                    // zlib header, then stored blocks of <= 65535 bytes
        uint32_t s1 = 1, s2 = 0;
        for(size_t off = 0; off < raw.size() || off == 0; ){
            const size_t n = std::min<size_t>(65535, raw.size() - off);
            z.push_back(off + n >= raw.size() ? 1 : 0);
            z.push_back((uint8_t)(n & 255)); z.push_back((uint8_t)(n >> 8)); z.push_back((uint8_t)(~n & 255)); z.push_back((uint8_t)((~n >> 8) & 255));
            for(size_t k = 0; k < n; k++){ const uint8_t v = raw[off + k]; z.push_back(v); s1 = (s1 + v) % 65521u; s2 = (s2 + s1) % 65521u; }
            off += n;
            if(n == 0)break;
        }
        const uint32_t adler = (s2 << 16) | s1;
        for(int k = 3; k >= 0; k--)z.push_back((uint8_t)(adler >> (8 * k)));
        std::ofstream out((outputPath.empty() ? filename : outputPath) + ".png", std::ios::binary);
        if(!out)throw std::runtime_error("Film::WritePNG: cannot open output file");
        auto chunk = [&](const char* type, const std::vector<uint8_t>& data){
            std::vector<uint8_t> buf(type, type + 4);
            buf.insert(buf.end(), data.begin(), data.end());
            uint32_t crc = 0xFFFFFFFFu;
            for(uint8_t v : buf){ crc ^= v; for(int k = 0; k < 8; k++)crc = (crc >> 1) ^ (0xEDB88320u & (0u - (crc & 1u))); }
            crc = ~crc;
            const uint32_t len = (uint32_t)data.size();
            for(int k = 3; k >= 0; k--)out.put((char)(len >> (8 * k)));
            out.write((const char*)buf.data(), (std::streamsize)buf.size());
            for(int k = 3; k >= 0; k--)out.put((char)(crc >> (8 * k)));
        };
        out.write("\x89PNG\r\n\x1a\n", 8);
        std::vector<uint8_t> hdr;
        for(uint32_t v : { (uint32_t)xResolution, (uint32_t)yResolution })for(int k = 3; k >= 0; k--)hdr.push_back((uint8_t)(v >> (8 * k)));
        for(uint8_t v : { 8, 2, 0, 0, 0 })hdr.push_back(v);        // 8 bits, RGB, deflate, adaptive filtering, no interlace
        chunk("IHDR", hdr);
        chunk("IDAT", z);
        chunk("IEND", {});
    }
    // JPEG, quality 1..100 (Film.hpp:196-217): the same tone-mapped, sRGB-encoded 8-bit image as WritePNG, through the baseline
    // encoder of ImageIO.hpp (the reference calls stb_image_write; the bytes differ, the decoded image agrees to JPEG accuracy)
    void WriteJPG(const std::string& filename, int quality, const std::function<glm::vec3(glm::vec3)>& toneMapper = reinhard_jodie) const{
        const std::vector<uint8_t> rgb = ToneMapped8(toneMapper);
        ptb_io::write_jpg((outputPath.empty() ? filename : outputPath) + ".jpg", xResolution, yResolution, rgb.data(), quality);
    }
    // the 8-bit image every writer encodes: top row first, tone-mapped, sRGB-encoded, 255.999 * clamp (Film.hpp:160-168)
    std::vector<uint8_t> ToneMapped8(const std::function<glm::vec3(glm::vec3)>& toneMapper = reinhard_jodie) const{
        std::vector<uint8_t> rgb((size_t)xResolution * yResolution * 3);
        size_t k = 0;
        for(int i = yResolution - 1; i >= 0; i--)for(int j = 0; j < xResolution; j++){
            const glm::vec3 c = toneMapper(glm::vec3(Pixel(j, i)));
            for(int a = 0; a < 3; a++)rgb[k++] = (uint8_t)(255.999 * std::max(0.0, std::min(1.0, linear_to_sRGB((double)c[a]))));
        }
        return rgb;
    }
private:
    int xResolution, yResolution;
    std::vector<double> screen;             // RGB sums * w and weight sums, 4 doubles per pixel
    std::shared_ptr<Filter> filter;
    double maxComponent;
    std::string outputPath;
};

// ---- Camera.hpp:4-60 --------------------------------------------------------------------------------------------
class Camera {
public:
    Camera(const glm::vec3& lookFrom, const glm::vec3& lookAt, float fov, const std::shared_ptr<Film>& film) : Camera(lookFrom, lookAt, fov, film, 0, 0){}
    Camera(const glm::vec3& lookFrom, const glm::vec3& lookAt, float fov, const std::shared_ptr<Film>& film, float FocusAngle, float FocusDistance,
           const std::shared_ptr<Medium>& medium = nullptr)
        : lookFrom(lookFrom), lookAt(lookAt), Fov(fov), film(film), focusAngle(FocusAngle), focusDistance(FocusDistance), medium(medium){}
    Camera(const glm::vec3& lookFrom, const glm::vec3& lookAt, float fov, const std::shared_ptr<Film>& film, const glm::vec2& shutterBounds)
        : Camera(lookFrom, lookAt, fov, film, 0, 0){ shutter = shutterBounds; }
    std::shared_ptr<Film> GetFilm() const{ return film; }
    std::shared_ptr<Medium> GetMedium() const{ return medium; }
    void SetMedium(const std::shared_ptr<Medium>& m){ medium = m; }
    pt_camera_desc Describe(int mediumId) const{
        pt_camera_desc c;
        for(int a = 0; a < 3; a++){ c.look_from[a] = lookFrom[a]; c.look_at[a] = lookAt[a]; }
        c.fov = Fov; c.focus_angle = focusAngle; c.focus_distance = focusDistance; c.medium = mediumId;
        c.shutter_start = shutter.x; c.shutter_end = shutter.y;
        return c;
    }
private:
    glm::vec3 lookFrom, lookAt;
    float Fov;                               // radians (Camera.hpp:12)
    std::shared_ptr<Film> film;
    float focusAngle, focusDistance;
    std::shared_ptr<Medium> medium;
    glm::vec2 shutter { 0.0f, 1.0f };
};

// ---- Sampler.hpp:9-151 --------------------------------------------------------------------------------------------
class Sampler {
public:
    virtual ~Sampler() = default;
    virtual unsigned int SamplesPerPixel() const = 0;
    virtual void Describe(pt_render_params& rp) const = 0;
};
class UniformSampler : public Sampler {
public:
    UniformSampler(unsigned int samples) : samples(samples){}
    unsigned int SamplesPerPixel() const override{ return samples; }
    void Describe(pt_render_params& rp) const override{ rp.sampler = 0; rp.spp = (int32_t)samples; rp.strata_x = rp.strata_y = 1; }
private:
    unsigned int samples;
};
class StratifiedSampler : public Sampler {
public:
    StratifiedSampler(unsigned int xSamples, unsigned int ySamples) : xSamples(xSamples), ySamples(ySamples){}
    unsigned int SamplesPerPixel() const override{ return xSamples * ySamples; }
    void Describe(pt_render_params& rp) const override{ rp.sampler = 1; rp.spp = (int32_t)(xSamples * ySamples); rp.strata_x = (int32_t)xSamples; rp.strata_y = (int32_t)ySamples; }
private:
    unsigned int xSamples, ySamples;
};

// ---- Integrators.hpp:10-70 ---------------------------------------------------------------------------------------------
class Integrator {
public:
    Integrator(const std::shared_ptr<Scene>& scene, const std::shared_ptr<Camera>& camera, const std::shared_ptr<Sampler>& sampler)
        : scene(scene), camera(camera), sampler(sampler){}
    virtual ~Integrator() = default;
    // Runs the whole frame on the GPUs and merges it into the camera's Film.  The reference's workers are host threads
    // (Integrators.cpp:23-41); here a worker is a GPU: threadCount = 0 (the default) uses every visible GPU, a smaller
    // number that many (PTB_GPUS in the environment overrides both); the wavefront kernels size themselves to each device.
    virtual void Render(unsigned int threadCount = 0) const = 0;
    bool Intersect(const Ray& ray, SurfaceInteraction& interaction, float max = std::numeric_limits<float>::infinity()) const{ return scene->Intersect(ray, interaction, max); }
    bool Unoccluded(const Ray& ray, float t) const{ return !scene->IntersectPred(ray, t); }
protected:
    std::shared_ptr<Scene> scene;
    std::shared_ptr<Camera> camera;
    std::shared_ptr<Sampler> sampler;
};

class TileIntegrator : public Integrator {
public:
    TileIntegrator(const std::shared_ptr<Scene>& scene, const std::shared_ptr<Camera>& camera, const std::shared_ptr<Sampler>& sampler)
        : Integrator(scene, camera, sampler){}
    void Render(unsigned int threadCount = 0) const override{
        const std::shared_ptr<Film> film = camera->GetFilm();
        pt_scene* s = scene->Commit(Sampler_(), camera->GetMedium());
        int gpus = threadCount == 0 ? pt_device_count() : (int)std::min<unsigned int>(threadCount, (unsigned int)pt_device_count());
        if(const char* e = std::getenv("PTB_GPUS"))gpus = std::max(1, std::atoi(e));
        pt_multi* multi = scene->Multi(s, gpus);
        pt_render_params rp;
        std::memset(&rp, 0, sizeof rp);
        const glm::ivec2 res = film->Resolution();
        rp.width = res.x; rp.height = res.y;
        rp.max_depth = (int32_t)MaxDepth(); rp.integrator = Kind();
        sampler->Describe(rp);
        film->GetFilter()->Describe(rp);
        rp.seed = seed; rp.pass_begin = 0; rp.pass_stride = 1;
        rp.adaptive = adaptive ? 1 : 0;
        const pt_camera_desc cd = camera->Describe(camera->GetMedium() ? scene->MediumId(camera->GetMedium().get()) : -1);
        pt_render_stats st;
        if(multi)ptb_host::check(pt_multi_render(multi, &cd, &rp, 1, &st), "pt_multi_render");   // sample passes sharded over the GPUs, one NCCL reduce of the film
        else ptb_host::check(pt_render(s, &cd, &rp, 1, &st), "pt_render");
        std::vector<double> rgbw(4 * (size_t)res.x * res.y);
        ptb_host::check(pt_film_download(s, rgbw.data()), "pt_film_download");
        film->Merge(rgbw);
        stats = st;
        ++seed;     // a second Render() into the same Film adds independent samples, as the reference's unseeded generators do
    }
    // extensions: the reference's generators are seeded from std::random_device and its adaptive loop
    // (Integrators.cpp:59-86) is always on
    void SetSeed(uint64_t s){ seed = s; }
    void SetAdaptive(bool on){ adaptive = on; }
    const pt_render_stats& LastStats() const{ return stats; }
protected:
    virtual int Kind() const = 0;
    virtual uint32_t MaxDepth() const = 0;
    virtual const LightSampler* Sampler_() const{ return nullptr; }
private:
    mutable uint64_t seed = 1;
    bool adaptive = true;
    mutable pt_render_stats stats {};
};

class SimplePathIntegrator : public TileIntegrator {
public:
    SimplePathIntegrator(const std::shared_ptr<Scene>& scene, const std::shared_ptr<Camera>& camera, const std::shared_ptr<Sampler>& sampler, uint32_t maxDepth)
        : TileIntegrator(scene, camera, sampler), maxDepth(maxDepth){}
protected:
    int Kind() const override{ return PT_INTEGRATOR_SIMPLE; }
    uint32_t MaxDepth() const override{ return maxDepth; }
private:
    uint32_t maxDepth;
};

class PathIntegrator : public TileIntegrator {
public:
    PathIntegrator(const std::shared_ptr<Scene>& scene, const std::shared_ptr<Camera>& camera, const std::shared_ptr<Sampler>& sampler,
                   const std::shared_ptr<LightSampler>& lightSampler, uint32_t maxDepth)
        : TileIntegrator(scene, camera, sampler), lightSampler(lightSampler), maxDepth(maxDepth){}
protected:
    int Kind() const override{ return PT_INTEGRATOR_PATH; }
    uint32_t MaxDepth() const override{ return maxDepth; }
    const LightSampler* Sampler_() const override{ return lightSampler.get(); }
private:
    std::shared_ptr<LightSampler> lightSampler;
    uint32_t maxDepth;
};

class VolPathIntegrator : public TileIntegrator {
public:
    VolPathIntegrator(const std::shared_ptr<Scene>& scene, const std::shared_ptr<Camera>& camera, const std::shared_ptr<Sampler>& sampler,
                      const std::shared_ptr<LightSampler>& lightSampler, uint32_t maxDepth)
        : TileIntegrator(scene, camera, sampler), lightSampler(lightSampler), maxDepth(maxDepth){}
protected:
    int Kind() const override{ return PT_INTEGRATOR_VOLPATH; }
    uint32_t MaxDepth() const override{ return maxDepth; }
    const LightSampler* Sampler_() const override{ return lightSampler.get(); }
private:
    std::shared_ptr<LightSampler> lightSampler;
    uint32_t maxDepth;
};

// ---- ResourceManager.hpp:8-105: name-keyed caches of textures and models ----------------------------------------
// Textures and meshes are plain shared objects here (the scene is lowered to the device at commit), so the
// manager is just the caches; the file-importing entry points (GetImageTexture, CacheModel: stb / assimp)
// are host I/O outside this path.
class ResourceManager {
public:
    static ResourceManager& get_instance(){ static ResourceManager instance; return instance; }
    ResourceManager(const ResourceManager&) = delete;
    ResourceManager& operator=(const ResourceManager&) = delete;

    template <typename T, typename... Args>
    [[nodiscard]] std::shared_ptr<Texture> GetTexture(const std::string& name, Args&&... args){
        static_assert(std::is_base_of_v<Texture, T>, "GetTexture<T>: T must be a Texture");
        if(name.empty())return nullptr;
        auto it = textureCache.find(name);
        if(it != textureCache.end())return it->second;
        return textureCache.try_emplace(name, std::make_shared<T>(std::forward<Args>(args)...)).first->second;
    }
    template <typename... Args>
    [[nodiscard]] std::shared_ptr<Mesh> getMesh(Args&&... args){
        meshCache.push_back(std::make_shared<Mesh>(std::forward<Args>(args)...));
        return meshCache.back();
    }
    // ResourceManager::CacheModel<BLASx>(name, path[, material, medium]) (ResourceManager.hpp:45-55): import + BuildBlas
    template <typename T, typename... Args>
    std::shared_ptr<Model> CacheModel(const std::string& name, const std::string& path, Args&&... args){
        auto it = modelCache.find(name);
        if(it != modelCache.end())return it->second;
        std::shared_ptr<Model> ptr = std::make_shared<Model>(path);
        ptr->template BuildBlas<T>(std::forward<Args>(args)...);
        return modelCache.try_emplace(name, ptr).first->second;
    }
    template <typename T, typename... Args>
    std::shared_ptr<Model> CacheOverwriteModel(const std::string& name, const std::string& path, Args&&... args){
        modelCache.erase(name);
        return CacheModel<T>(name, path, std::forward<Args>(args)...);
    }
    // extension: a model that was assembled in memory
    std::shared_ptr<Model> CacheModel(const std::string& name, const std::shared_ptr<Model>& model){ return modelCache.try_emplace(name, model).first->second; }
    [[nodiscard]] std::shared_ptr<Model> GetModel(const std::string& name){
        auto it = modelCache.find(name);
        return it != modelCache.end() ? it->second : nullptr;
    }
    void releaseTextureCache(){ textureCache.clear(); }
    void releaseMeshCache(){ meshCache.clear(); }
    void releaseModelCache(){ modelCache.clear(); }
    void releaseAll(){ releaseTextureCache(); releaseMeshCache(); releaseModelCache(); }
private:
    ResourceManager() = default;
    std::unordered_map<std::string, std::shared_ptr<Texture>> textureCache;
    std::unordered_map<std::string, std::shared_ptr<Model>> modelCache;
    std::vector<std::shared_ptr<Mesh>> meshCache;
};

#include "ObjImport.hpp"
#include "GltfImport.hpp"
inline Model::Model(const std::string& path) : meshes(ptb_host::load_model(path)){}
