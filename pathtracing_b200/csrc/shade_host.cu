// shade_host.cu — uploads the shading tables built by shade_tables.cpp and fills pt_scene::shade
// with the device pointers.
#include <cuda_runtime.h>
#include <cstdlib>
#include <cstdio>
#include <chrono>

#include "shade_tables.hpp"

namespace ptb {

namespace {

thread_local pt_scene* g_scene = nullptr;      // the scene build_shade_tables is uploading for (commit is not re-entrant)

template <class T>
bool up(void** dst, const std::vector<T>& v, const char* what){
    *dst = nullptr;
    size_t bytes = v.size() * sizeof(T);
    if(bytes == 0)bytes = 16;
    if(!cuda_ok(cudaMalloc(dst, bytes), what))return false;
    if(!v.empty() && !cuda_ok(cudaMemcpy(*dst, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice), what))return false;
    note_upload(g_scene, dst, bytes);
    return true;
}

} // namespace

int build_shade_tables(pt_scene* s, const std::vector<uint32_t>& inst_order){
    ShadeTables T;
    int rc = build_shade_tables_host(s, inst_order, T);
    if(rc != PT_OK)return rc;
    DeviceBuffers& b = s->buf;
    g_scene = s;
    const auto tu0 = std::chrono::steady_clock::now();
    if(!up(&b.meshes, T.meshes, "meshes") || !up(&b.blas_info, T.blas, "blas info") || !up(&b.inst_shade, T.inst, "instances") ||
       !up(&b.vtx_pos, T.pos, "positions") || !up(&b.vtx_nrm, T.nrm, "normals") || !up(&b.vtx_uv, T.uv, "texcoords") ||
       !up(&b.vtx_tan, T.tan, "tangents") || !up(&b.indices, T.idx, "indices") || !up(&b.tri_mesh, T.stri, "shading triangles") ||
       !up(&b.tri_light, T.tri_emis, "tri_emis") || !up(&b.textures, T.textures, "textures") || !up(&b.texels, T.texels, "texels") ||
       !up(&b.materials, T.materials, "materials") || !up(&b.media, T.media, "media") || !up(&b.lights, T.lights, "lights") ||
       !up(&b.sampler_lights, T.sampler_lights, "sampler lights") || !up(&b.light_cdf, T.sampler_cdf, "light cdf") ||
       !up(&b.inf_lights, T.inf_lights, "infinite lights") || !up(&b.env_cdf, T.env_cdf, "env cdf"))return PT_ERR_CUDA;
    if(getenv("PTB_PROFILE"))fprintf(stderr, "[ptb] shade tables: upload %.1f ms\n", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tu0).count());
    ShadeScene& D = s->shade;      // pointer fields: bind_device_pointers (scene_host.cu)
    D.n_lights = (int32_t)T.lights.size(); D.n_sampler_lights = (int32_t)T.sampler_lights.size(); D.n_inf_lights = (int32_t)T.inf_lights.size();
    D.sampler_kind = T.view.sampler_kind; D.total_power = T.view.total_power;
    D.scene_medium = T.view.scene_medium;
    D.env_w = T.view.env_w; D.env_h = T.view.env_h;
    return PT_OK;
}

} // namespace ptb
