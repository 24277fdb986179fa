// shade_types.h — device tables read by the shade stage (materials, textures,
// lights, media, per-mesh vertex data).  Plain PODs, uploaded once at commit.
#pragma once
#include <stdint.h>

namespace ptb {

struct DevMesh {
    uint32_t vtx_base;       // first vertex in the global vertex arrays
    uint32_t idx_base;       // first index (uint32) in the global index array
    uint32_t prim_base;      // BLAS-wide id of this mesh's triangle 0
    uint32_t has_tangents;
    int32_t material;        // -1: medium boundary (no material)
    int32_t emissive_tex;    // -1: not emissive
    int32_t medium;          // -1: none
    uint32_t pad;
};

struct DevBlasInfo {
    uint32_t tri_base;       // offset of this BLAS in the per-triangle tables (original order)
    uint32_t n_tris;
};

struct DevInstShade {
    float fwd[12];           // rows 0..2 of transform, fwd[c*3+r] (c = 0..3, translation c = 3)
    float nrm[9];            // transpose(inverse(mat3(transform))), nrm[c*3+r]
    float inv[12];           // rows 0..2 of inverse transform
    float det;               // glm::determinant(transform) (TransformedLight::Power)
    uint32_t blas;
    uint32_t has_xf;
    int32_t light_base;      // first DevLight of this instance's emissive triangles (shape: its AreaLight), or -1
    // GeometricPrimitive over an analytic shape (shape >= 0): no BLAS; shp[] holds the shape constants in the
    // layout of the traversal record (scene_build.cpp pack_shape):
    //   sphere: shp[0..2] center, shp[3] radius        quad: shp[0..2] Q, [3..5] u, [6..8] v, [9..11] normal
    // has_xf then means TransformedPrimitive(GeometricPrimitive(shape), M) and fwd / nrm / inv / det are M's.
    int32_t shape;           // -1 | pt_shape_kind
    int32_t material;        // shape: material id or -1 (medium boundary)
    int32_t medium;          // shape: medium id or -1
    uint32_t animated;       // inside an AnimatedPrimitive (Primitive.cpp:77-96): translate(adir * t(time)) on top
    float shp[12];
    float adir[3];
    float atb[2];
    uint32_t tri_base;       // BLAS instance: offset of its BLAS in the per-triangle tables (DevBlasInfo::tri_base, copied here: one dependent load less per shaded hit)
    uint32_t pad[2];
};

// What shading needs to know about a triangle, one 16 B load at [instance.tri_base + prim]: its three vertices in the global
// vertex arrays and its mesh.  (Round 1 walked instance -> BLAS info -> tri_mesh -> mesh -> 3 indices -> vertices: five dependent
// loads before the first vertex arrives, in a kernel that is latency-bound at 24 warps per SM.)
struct ShadeTri { uint32_t i0, i1, i2, mesh; };

struct DevTexture {
    int32_t kind;            // pt_tex_kind
    float color[3];          // SOLID albedo
    float scale[3];          // colorScale
    int32_t w, h, ch;
    unsigned long long texels; // byte offset into the texel blob
    int32_t tex_a, tex_b;
    float inv_scale[2];      // CHECKER: 1/uvscale
};

struct DevMaterial {
    int32_t kind;            // pt_mat_kind
    int32_t tex, norm, rough, metal, alpha;
    float ri;
    float albedo[3];
    int32_t alpha_mode;
    float alpha_cutoff;
};

struct DevMedium {
    float sigma_a[3], sigma_s[3], sigma_t[3], Le[3];
    float g;
};

enum DevLightKind : int32_t {
    DL_UNIFORM_INF = 0, DL_GRADIENT_INF = 1, DL_TEXTURE_INF = 2, DL_DISTANT = 3, DL_POINT = 4, DL_AREA = 5,
    DL_SPHERE = 6, DL_QUAD = 7      // AreaLight over a SphereShape / QuadShape (geometry in S.inst[L.inst].shp)
};

struct DevLight {
    int32_t kind;
    float color[3];
    float color_b[3];
    float scale;
    float v[3];              // DISTANT dir / POINT position
    int32_t tex;             // TEXTURE_INF texture | AREA emissive texture
    float pmf_sample;        // LightSampler::PMF(light) of the light object the sampler holds
    float pmf_hit;           // AREA: LightSampler::PMF(interaction.AreaLight) — the inner, untransformed light
    uint32_t valid;          // AREA: 0 if Model::BuildBlas dropped it (Power <= FLT_EPSILON): not emissive at all
    uint32_t inst;           // AREA: instance id
    uint32_t mesh;           // AREA: scene mesh id
    uint32_t tri;            // AREA: mesh-local triangle index
    float area;              // AREA: TriangleShape::Area() in object space
    float total_weight;      // TEXTURE_INF: sum of the luminance table
    uint32_t one_sided;      // SPHERE / QUAD: AreaLight::oneSided (mesh lights are always two-sided, Model.hpp:50)
};

struct ShadeScene {
    const DevMesh* meshes;
    const DevBlasInfo* blas;
    const DevInstShade* inst;
    const float* pos;        // 3 floats per vertex
    const float* nrm;
    const float* uv;         // 2 floats per vertex
    const float* tan;        // 3 floats per vertex (zeros if mesh has none)
    const uint32_t* idx;
    const ShadeTri* stri;      // per BLAS-wide triangle (original order): global vertex ids + mesh id
    const int32_t* tri_emis;   // per BLAS-wide triangle: ordinal among the BLAS's emissive triangles, or -1
    const DevTexture* textures;
    const uint8_t* texels;
    const DevMaterial* materials;
    const DevMedium* media;
    const DevLight* lights;    // all lights (also the ones the sampler culled)
    const int32_t* sampler_lights; // indices of the lights the sampler can pick
    const float* sampler_cdf;  // running power sum (PowerLightSampler::Sample walks it)
    const int32_t* inf_lights; // indices of scene->infiniteLights
    const float* env_cdf;      // TEXTURE_INF: running luminance sum (env_w*env_h cells)
    int32_t n_lights, n_sampler_lights, n_inf_lights;
    int32_t sampler_kind;      // 0 uniform, 1 power
    float total_power;
    int32_t scene_medium;
    int32_t env_w, env_h;
};

} // namespace ptb
