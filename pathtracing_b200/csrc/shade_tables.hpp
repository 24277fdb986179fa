// shade_tables.hpp — host-side staging of everything ShadeScene points to (shade_types.h).
// Pure host code (no CUDA calls): shade_host.cu uploads it, tests/emul evaluates it on the CPU.
#pragma once
#include <vector>
#include "scene_host.hpp"

namespace ptb {

struct ShadeTables {
    std::vector<DevMesh> meshes;
    std::vector<DevBlasInfo> blas;
    std::vector<DevInstShade> inst;
    std::vector<float> pos, nrm, uv, tan;
    std::vector<uint32_t> idx;
    std::vector<ShadeTri> stri;
    std::vector<int32_t> tri_emis;
    std::vector<DevTexture> textures;
    std::vector<uint8_t> texels;
    std::vector<DevMaterial> materials;
    std::vector<DevMedium> media;
    std::vector<DevLight> lights;
    std::vector<int32_t> sampler_lights, inf_lights;
    std::vector<float> sampler_cdf, env_cdf;
    ShadeScene view {};     // pointers into the vectors above (host) for commit-time evaluation
    // geometry_on_device (set by the caller before the build): the per-vertex and per-triangle tables (pos, nrm, uv, tan, idx, stri,
    // tri_emis) are NOT staged here -- shade_host.cu assembles them on the device from the meshes' own arrays (a 10 M-triangle scene
    // spent 0.16 s concatenating them on the host and 0.2-0.7 s uploading the 420 MB copy from pageable memory).  What it needs:
    bool geometry_on_device = false;
    size_t n_vertices = 0, n_indices = 0, n_blas_tris = 0;
    struct TriRun { int mesh; uint32_t first, count; int32_t emis_first; };   // mesh's triangles at [first, first + count) of the per-triangle tables; emis_first: ordinal of its first triangle among the BLAS's emissive ones, or -1
    std::vector<TriRun> tri_runs;
};


// Builds every table and points T.view at them (host pointers).  inst_order: TLAS leaf order -> instance id.
int build_shade_tables_host(pt_scene* s, const std::vector<uint32_t>& inst_order, ShadeTables& T);

} // namespace ptb
