// shapes.cpp — the scene of the reference's examples/example_1.cpp (analytic shapes only: a checker floor
// QuadShape, a diffuse SphereShape, a small QuadShape AreaLight, a material-less SphereShape holding a
// HomogeneusMedium, a uniform sky) against the drop-in host API, at a size given on the command line.
// The reference's own examples/example_1.cpp also compiles unmodified against these headers
// (tests/test_host_api.py); this twin exists so that the GPU box, which has no reference tree, can run it.
//
//   g++ -std=c++20 -Iinclude -Ipathtracing_b200/host -Ipathtracing_b200/host/glm_min examples/shapes.cpp
//     -Lpathtracing_b200 -lptb200 -Wl,-rpath,$PWD/pathtracing_b200 -o shapes
//   ./shapes [width height spp depth]   -> shapes_{simple,path,volpath}.png + one JSON line
#include <cstdio>
#include <cstdlib>
#include <memory>

#include "Material.hpp"
#include "Primitive.hpp"
#include "Shape.hpp"
#include "Light.hpp"
#include "LightSampler.hpp"
#include "Sampler.hpp"
#include "Filter.hpp"
#include "Film.hpp"
#include "Camera.hpp"
#include "Scene.hpp"
#include "ResourceManager.hpp"
#include "Integrators.hpp"

static glm::dvec3 FilmMean(const Film& film){
    const glm::ivec2 res = film.Resolution();
    glm::dvec3 acc(0.0);
    for(int y = 0; y < res.y; y++)for(int x = 0; x < res.x; x++)acc += film.Pixel(x, y);
    return acc * (1.0 / ((double)res.x * res.y));
}

int main(int argc, char** argv){
    const int width = argc > 2 ? atoi(argv[1]) : 480, height = argc > 2 ? atoi(argv[2]) : 270;
    const int spp = argc > 3 ? atoi(argv[3]) : 16;
    const uint32_t depth = argc > 4 ? (uint32_t)atoi(argv[4]) : 128;

    auto scene = std::make_shared<Scene>();
    ResourceManager& rm = ResourceManager::get_instance();
    std::shared_ptr<Texture> white = rm.GetTexture<SolidColor>("whiteTexture", glm::vec3(.9f));
    std::shared_ptr<Texture> green = rm.GetTexture<SolidColor>("greenTexture", glm::vec3(.2f, .3f, .1f));
    std::shared_ptr<Material> checker = std::make_shared<MicrofacetDiffuse>(std::make_shared<CheckerTexture>(white, green, glm::vec2(0.001f, 0.001f)));
    std::shared_ptr<Material> sphereMaterial = std::make_shared<MicrofacetDiffuse>(green);

    std::shared_ptr<Shape> floor = std::make_shared<QuadShape>(glm::vec3(-100, -0.3f, -100), glm::vec3(1000, 0, 0), glm::vec3(0, 0, 1000));
    std::shared_ptr<Shape> sphere = std::make_shared<SphereShape>(glm::vec3(0, 0.1f, -1.2f), 0.5f);
    std::shared_ptr<Shape> mediumSphere = std::make_shared<SphereShape>(glm::vec3(1, 0, -1), 0.5f);
    std::shared_ptr<Shape> lightShape = std::make_shared<QuadShape>(glm::vec3(-1, -0.28f, -1), glm::vec3(0.2f, 0, -0.2f), glm::vec3(0, 0.2f, 0));
    std::shared_ptr<AreaLight> areaLight = std::make_shared<AreaLight>(lightShape, glm::vec3(1, 0, 0) * 600.0f, false);

    std::shared_ptr<Medium> fog = std::make_shared<HomogeneusMedium>(glm::vec3(0.01f, 0.9f, 0.9f), glm::vec3(1.0f, 0.1f, 0.1f),
                                                                       std::make_shared<HenyeyGreenstein>(0.8f), 5.0f);
    scene->Add(std::make_shared<GeometricPrimitive>(floor, checker, nullptr, nullptr));
    scene->Add(std::make_shared<GeometricPrimitive>(sphere, sphereMaterial, nullptr));
    scene->Add(std::make_shared<GeometricPrimitive>(areaLight->getShape(), std::make_shared<MicrofacetDiffuse>(glm::vec3(0)), areaLight, nullptr));
    scene->Add(std::make_shared<GeometricPrimitive>(mediumSphere, nullptr, nullptr, fog));
    scene->infiniteLights.push_back(std::make_shared<UniformInfiniteLight>(glm::vec3(0.45f, 0.65f, 1)));
    scene->BuildTlas<TLAS4>();

    auto film = std::make_shared<Film>(glm::ivec2 { width, height }, std::make_shared<MitchellFilter>());
    auto camera = std::make_shared<Camera>(glm::vec3(0.3f, 0.4f, 1), glm::vec3(0, 0, 0), 1.7f, film);
    auto sampler = std::make_shared<UniformSampler>(spp);
    auto lightSampler = std::make_shared<UniformLightSampler>();
    lightSampler->Add(scene->GetLights());
    lightSampler->PreProcess(scene->BoundingBox());

    std::shared_ptr<TileIntegrator> integrators[3] = {
        std::make_shared<SimplePathIntegrator>(scene, camera, sampler, depth),
        std::make_shared<PathIntegrator>(scene, camera, sampler, lightSampler, depth),
        std::make_shared<VolPathIntegrator>(scene, camera, sampler, lightSampler, depth) };
    const char* names[3] = { "shapes_simple", "shapes_path", "shapes_volpath" };
    glm::dvec3 mean[3];
    unsigned long long samples[3], shadow[3];
    for(int k = 0; k < 3; k++){
        integrators[k]->SetAdaptive(false);
        integrators[k]->Render();
        mean[k] = FilmMean(*film);
        samples[k] = integrators[k]->LastStats().samples; shadow[k] = integrators[k]->LastStats().shadow_rays;
        film->WritePNG(names[k]);
        film->Clear();
    }
    // one ray down onto the sphere: SphereShape::Intersect through the GPU
    SurfaceInteraction si;
    const bool hit = scene->Intersect(Ray(glm::vec3(0, 2.0f, -1.2f), glm::vec3(0, -1, 0)), si);
    printf("{\"samples\": %llu, \"simple_mean\": [%.6g, %.6g, %.6g], \"path_mean\": [%.6g, %.6g, %.6g], \"volpath_mean\": [%.6g, %.6g, %.6g], "
           "\"path_shadow_rays\": %llu, \"volpath_shadow_rays\": %llu, \"hit\": %d, \"hit_t\": %.7g, \"hit_instance\": %u}\n",
           samples[1], mean[0].x, mean[0].y, mean[0].z, mean[1].x, mean[1].y, mean[1].z, mean[2].x, mean[2].y, mean[2].z,
           shadow[1], shadow[2], hit ? 1 : 0, si.t, si.instance);
    rm.releaseTextureCache();
    rm.releaseMeshCache();
    rm.releaseModelCache();
    return 0;
}
