// blas_list.cpp — a BLAS built from a hand-made GeometricPrimitive list (BVH.hpp:743 `BLAS4(const std::vector<GeometricPrimitive>&)`, with
// the recipe of Model::BuildBlas, Model.hpp:47-59: Mesh::GetShape / GetControlPtr, one AreaLight per emissive triangle), and nested
// wrappers (Primitive.hpp:37: a TransformedPrimitive takes any Primitive): TransformedPrimitive(TransformedPrimitive(blas)) and
// TransformedPrimitive(AnimatedPrimitive(sphere)).
//
//   ./blas_list                      -> compiles / builds the object graph, prints the sub-mesh split (no device needed)
//   ./blas_list width height spp     -> ray-casts one ray per list entry (primitive id == list position) and renders (GPU)
#include <cstdio>
#include <cstdlib>
#include <memory>

#include "Scene.hpp"
#include "Mesh.hpp"
#include "Shape.hpp"
#include "Material.hpp"
#include "Primitive.hpp"
#include "BVH.hpp"
#include "Light.hpp"
#include "LightSampler.hpp"
#include "Sampler.hpp"
#include "Filter.hpp"
#include "Film.hpp"
#include "Camera.hpp"
#include "Integrators.hpp"

int main(int argc, char** argv){
    // a strip of 6 quads (12 triangles) in the plane y = 0, x in [0, 6]
    std::vector<glm::vec3> P, N; std::vector<glm::vec2> UV; std::vector<uint32_t> I;
    for(int q = 0; q < 6; q++){
        const uint32_t b = (uint32_t)P.size();
        P.push_back({ (float)q, 0, 0 }); P.push_back({ (float)q + 1, 0, 0 }); P.push_back({ (float)q + 1, 0, 1 }); P.push_back({ (float)q, 0, 1 });
        for(int k = 0; k < 4; k++){ N.push_back({ 0, 1, 0 }); UV.push_back({ (float)(k == 1 || k == 2), (float)(k >= 2) }); }
        for(uint32_t k : { 0u, 2u, 1u, 0u, 3u, 2u })I.push_back(b + k);
    }
    auto grey = std::make_shared<MicrofacetDiffuse>(glm::vec3(0.7f));
    auto red = std::make_shared<MicrofacetDiffuse>(glm::vec3(0.8f, 0.1f, 0.1f));
    auto mirror = std::make_shared<SpecularConductor>(glm::vec3(0.9f));
    auto glow = std::make_shared<SolidColor>(glm::vec3(4.0f, 3.0f, 2.0f));
    auto mesh = std::make_shared<Mesh>(I, P, std::vector<glm::vec3>{}, N, UV, grey, nullptr, nullptr);

    // the recipe of Model::BuildBlas (Model.hpp:47-59), with a material of our own choosing per triangle
    std::vector<GeometricPrimitive> prims;
    for(uint32_t j = 0; j < mesh->GetTriangleCount(); j++){
        std::shared_ptr<Shape> shape = std::shared_ptr<Shape>(mesh->GetControlPtr(), mesh->GetShape(j));
        const bool emit = j == 4 || j == 5;
        std::shared_ptr<AreaLight> area = emit ? std::make_shared<AreaLight>(shape, glow) : nullptr;
        if(area){ area->PreProcess({}); if(area->Power() <= std::numeric_limits<float>::epsilon())area = nullptr; }
        prims.emplace_back(shape, j < 4 ? std::static_pointer_cast<Material>(grey) : j < 8 ? std::static_pointer_cast<Material>(red) : std::static_pointer_cast<Material>(mirror), area, nullptr);
    }
    auto blas = std::make_shared<BLAS4>(prims);
    printf("{\"list\": %zu, \"sub_meshes\": %zu", prims.size(), blas->GetMeshes().size());
    for(const auto& m : blas->GetMeshes())printf(", \"%u\": %d", m->GetTriangleCount(), m->GetEmissiveTexture() ? 1 : 0);

    // nested wrappers: scale by 2 inside a translation by (0, 0, -3)
    glm::mat4 S(1.0f), T(1.0f);
    S[0][0] = S[1][1] = S[2][2] = 2.0f;
    T[3][2] = -3.0f;
    auto nested = std::make_shared<TransformedPrimitive>(std::make_shared<TransformedPrimitive>(blas, S), T);
    // a transformed animated sphere: TransformedPrimitive(AnimatedPrimitive(...)) is rewritten to the chain the device understands
    auto ball = std::make_shared<GeometricPrimitive>(std::make_shared<SphereShape>(glm::vec3(0, 0.5f, 0), 0.5f), red);
    auto moving = std::make_shared<TransformedPrimitive>(std::make_shared<AnimatedPrimitive>(ball, glm::vec3(1, 0, 0), glm::vec2(0, 1)), T);

    auto scene = std::make_shared<Scene>();
    scene->Add(nested);
    scene->Add(moving);
    scene->infiniteLights.push_back(std::make_shared<UniformInfiniteLight>(glm::vec3(0.3f, 0.35f, 0.5f)));
    scene->BuildTlas<TLAS4>();
    if(argc < 4){ printf("}\n"); return 0; }

    // one ray per list entry, straight down onto the triangle's centroid (in world space: scaled by 2, moved by -3 in z)
    std::vector<pt_ray> rays;
    for(uint32_t j = 0; j < mesh->GetTriangleCount(); j++){
        glm::vec3 c(0.0f);
        for(int k = 0; k < 3; k++)c = c + P[I[3 * j + k]];
        c = c * (1.0f / 3.0f) * 2.0f + glm::vec3(0, 0, -3);
        rays.push_back({ c.x, 5.0f, c.z, 1e30f, 0, -1, 0, 0 });
    }
    std::vector<pt_hit> hits(rays.size());
    std::vector<uint32_t> inst(rays.size());
    scene->IntersectBatch(rays.data(), rays.size(), hits.data(), inst.data());
    int ok = 1;
    for(size_t j = 0; j < rays.size(); j++)ok = ok && hits[j].prim == j && inst[j] == 0 && std::fabs(hits[j].t - 5.0f) < 1e-4f;
    printf(", \"ids_are_list_positions\": %d", ok);

    const int width = atoi(argv[1]), height = atoi(argv[2]);
    const unsigned spp = (unsigned)atoi(argv[3]);
    auto film = std::make_shared<Film>(glm::ivec2 { width, height }, std::make_shared<BoxFilter>());
    auto camera = std::make_shared<Camera>(glm::vec3(6, 6, 6), glm::vec3(6, 0, -2), 0.9f, film);
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
    printf(", \"mean\": [%.6f, %.6f, %.6f], \"samples\": %llu}\n", m.x, m.y, m.z, (unsigned long long)integrator->LastStats().samples);
    return 0;
}
