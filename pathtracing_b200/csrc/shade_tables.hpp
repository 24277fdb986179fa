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
};


// Builds every table and points T.view at them (host pointers).  inst_order: TLAS leaf order -> instance id.
int build_shade_tables_host(pt_scene* s, const std::vector<uint32_t>& inst_order, ShadeTables& T);

} // namespace ptb
