// obj_model.cpp — Model(path) / ResourceManager::CacheModel<BLAS4>(name, path) as the reference's model scenes use
// them (main.cpp:290,376-381,483), over the importers of the host API: Wavefront OBJ / MTL (ObjImport.hpp) and glTF 2.0
// .gltf / .glb (GltfImport.hpp), picked by the file extension.
//
//   ./obj_model file.obj                 -> one JSON line with the imported meshes (no device needed)
//   ./obj_model file.obj width height spp -> also renders it under a gradient sky + its own emitters (GPU)
#include <cstdio>
#include <cstdlib>
#include <memory>

#include "Model.hpp"
#include "Material.hpp"
#include "ResourceManager.hpp"
#include "Light.hpp"
#include "LightSampler.hpp"
#include "Sampler.hpp"
#include "Filter.hpp"
#include "Film.hpp"
#include "Camera.hpp"
#include "Scene.hpp"
#include "Integrators.hpp"

int main(int argc, char** argv){
    if(argc < 2){ fprintf(stderr, "usage: obj_model file.obj [width height spp]\n"); return 2; }
    std::shared_ptr<Model> model = ResourceManager::get_instance().CacheModel<BLAS4>("model", argv[1]);
    size_t tris = 0, verts = 0, emissive = 0, tangents = 0;
    glm::vec3 lo(1e30f), hi(-1e30f);
    for(const std::shared_ptr<Mesh>& m : model->GetMeshes()){
        tris += m->GetTriangleCount(); verts += m->GetVertexCount();
        emissive += m->GetEmissiveTexture() ? 1 : 0;
        tangents += m->GetTangents().empty() ? 0 : 1;
        for(const glm::vec3& p : m->GetVertices()){
            lo = glm::vec3(std::min(lo.x, p.x), std::min(lo.y, p.y), std::min(lo.z, p.z));
            hi = glm::vec3(std::max(hi.x, p.x), std::max(hi.y, p.y), std::max(hi.z, p.z));
        }
    }
    printf("{\"meshes\": %zu, \"triangles\": %zu, \"vertices\": %zu, \"emissive_meshes\": %zu, \"meshes_with_tangents\": %zu, \"lo\": [%g, %g, %g], \"hi\": [%g, %g, %g]",
           model->GetMeshes().size(), tris, verts, emissive, tangents, lo.x, lo.y, lo.z, hi.x, hi.y, hi.z);
    // per mesh: material class, medium, the first vertex as imported (position, normal, uv)
    printf(", \"per_mesh\": [");
    bool first = true;
    for(const std::shared_ptr<Mesh>& m : model->GetMeshes()){
        const char* kind = std::dynamic_pointer_cast<MicrofacetDiffuse>(m->GetMaterial()) ? "diffuse" : std::dynamic_pointer_cast<MicrofacetDielectric>(m->GetMaterial()) ? "dielectric"
                         : std::dynamic_pointer_cast<ThinDielectric>(m->GetMaterial()) ? "thin" : std::dynamic_pointer_cast<SpecularConductor>(m->GetMaterial()) ? "conductor" : "other";
        const glm::vec3 p = m->GetVertices()[0], n = m->GetNormals()[0];
        const glm::vec2 t = m->GetTexCoords()[0];
        printf("%s{\"material\": \"%s\", \"medium\": %d, \"emissive\": %d, \"triangles\": %u, \"p0\": [%g, %g, %g], \"n0\": [%g, %g, %g], \"uv0\": [%g, %g]}", first ? "" : ", ", kind,
               m->GetMedium() ? 1 : 0, m->GetEmissiveTexture() ? 1 : 0, m->GetTriangleCount(), p.x, p.y, p.z, n.x, n.y, n.z, t.x, t.y);
        first = false;
    }
    printf("]");
    if(argc < 5){ printf("}\n"); return 0; }

    const int width = atoi(argv[2]), height = atoi(argv[3]);
    const unsigned spp = (unsigned)atoi(argv[4]);
    auto scene = std::make_shared<Scene>();
    scene->Add(model);
    scene->infiniteLights.push_back(std::make_shared<UniformInfiniteLight>(glm::vec3(0.4f, 0.5f, 0.7f)));
    scene->BuildTlas<TLAS4>();
    const glm::vec3 c = (lo + hi) * 0.5f;
    const float r = glm::length(hi - lo);
    auto film = std::make_shared<Film>(glm::ivec2 { width, height }, std::make_shared<BoxFilter>());
    auto camera = std::make_shared<Camera>(c + glm::vec3(0.3f, 0.4f, 1.0f) * r, c, 0.9f, film);
    auto sampler = std::make_shared<UniformSampler>(spp);
    auto lightSampler = std::make_shared<PowerLightSampler>();
    lightSampler->Add(scene->GetLights());
    lightSampler->PreProcess(scene->BoundingBox());
    auto integrator = std::make_shared<PathIntegrator>(scene, camera, sampler, lightSampler, 6);
    integrator->SetAdaptive(false);
    integrator->Render();
    glm::dvec3 m(0.0);
    for(int y = 0; y < height; y++)for(int x = 0; x < width; x++)m += film->Pixel(x, y);
    m = m * (1.0 / ((double)width * height));
    film->WritePPM("obj_model");
    printf(", \"mean\": [%.6f, %.6f, %.6f], \"samples\": %llu}\n", m.x, m.y, m.z, (unsigned long long)integrator->LastStats().samples);
    return 0;
}
