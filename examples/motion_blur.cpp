// motion_blur.cpp — the moving sphere of the reference's main.cpp (lines 187-190: an AnimatedPrimitive around a
// sphere GeometricPrimitive, 0.2 up over the shutter interval) next to a TransformedPrimitive over a sphere
// (an ellipsoid) and a tilted quad light inside a TransformedPrimitive, against the drop-in host API.
//
//   g++ -std=c++20 -Iinclude -Ipathtracing_b200/host -Ipathtracing_b200/host/glm_min examples/motion_blur.cpp
//     -Lpathtracing_b200 -lptb200 -Wl,-rpath,$PWD/pathtracing_b200 -o motion_blur
//   ./motion_blur [width height spp depth]   -> motion_blur.png + one JSON line
#include <cstdio>
#include <cstdlib>
#include <memory>

#include <glm/gtc/matrix_transform.hpp>

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
#include "Integrators.hpp"

int main(int argc, char** argv){
    const int width = argc > 2 ? atoi(argv[1]) : 480, height = argc > 2 ? atoi(argv[2]) : 270;
    const int spp = argc > 3 ? atoi(argv[3]) : 64;
    const uint32_t depth = argc > 4 ? (uint32_t)atoi(argv[4]) : 16;

    auto scene = std::make_shared<Scene>();
    auto ground = std::make_shared<MicrofacetDiffuse>(std::make_shared<CheckerTexture>(
        std::make_shared<SolidColor>(glm::vec3(.8f)), std::make_shared<SolidColor>(glm::vec3(.2f, .3f, .1f)), glm::vec2(0.002f, 0.002f)));
    scene->Add(std::make_shared<GeometricPrimitive>(std::make_shared<QuadShape>(glm::vec3(-100, -0.3f, -100), glm::vec3(1000, 0, 0), glm::vec3(0, 0, 1000)), ground, nullptr, nullptr));

    // main.cpp:187-190
    auto tm = std::make_shared<GeometricPrimitive>(std::make_shared<SphereShape>(glm::vec3(0, 0.1f, -1.2f), 0.5f),
                                                   std::make_shared<MicrofacetDiffuse>(glm::vec3(0.1f, 0.2f, 0.5f)), nullptr);
    scene->Add(std::make_shared<AnimatedPrimitive>(tm, glm::vec3 { 0, 0.2f, 0 }, glm::vec2 { 0, 1 }));

    // an ellipsoid: a unit sphere under translate * rotate * scale
    glm::mat4 m = glm::translate(glm::mat4(1), glm::vec3(1.2f, 0.1f, -0.8f));
    m = glm::rotate(m, 0.6f, glm::vec3(0, 0, 1));
    m = glm::scale(m, glm::vec3(0.6f, 0.25f, 0.4f));
    auto ball = std::make_shared<GeometricPrimitive>(std::make_shared<SphereShape>(glm::vec3(0), 1.0f), std::make_shared<SpecularConductor>(glm::vec3(0.8f, 0.6f, 0.2f)), nullptr);
    scene->Add(std::make_shared<TransformedPrimitive>(ball, m));

    // a quad light tilted by a TransformedPrimitive (its AreaLight becomes a TransformedLight)
    std::shared_ptr<Shape> lightShape = std::make_shared<QuadShape>(glm::vec3(-0.4f, 0, -0.4f), glm::vec3(0.8f, 0, 0), glm::vec3(0, 0, 0.8f));
    auto areaLight = std::make_shared<AreaLight>(lightShape, glm::vec3(12.0f), false);
    auto lamp = std::make_shared<GeometricPrimitive>(lightShape, std::make_shared<MicrofacetDiffuse>(glm::vec3(0)), areaLight, nullptr);
    scene->Add(std::make_shared<TransformedPrimitive>(lamp, glm::rotate(glm::translate(glm::mat4(1), glm::vec3(-0.6f, 1.8f, -0.6f)), 0.5f, glm::vec3(1, 0, 0))));

    scene->infiniteLights.push_back(std::make_shared<UniformInfiniteLight>(glm::vec3(0.25f, 0.35f, 0.5f)));
    scene->BuildTlas<TLAS4>();

    auto film = std::make_shared<Film>(glm::ivec2 { width, height }, std::make_shared<MitchellFilter>());
    auto camera = std::make_shared<Camera>(glm::vec3(0.3f, 0.6f, 1.6f), glm::vec3(0.2f, 0.1f, -1), 1.2f, film);
    auto sampler = std::make_shared<UniformSampler>(spp);
    auto lightSampler = std::make_shared<PowerLightSampler>();
    lightSampler->Add(scene->GetLights());
    lightSampler->PreProcess(scene->BoundingBox());

    PathIntegrator integrator(scene, camera, sampler, lightSampler, depth);
    integrator.SetAdaptive(false);
    integrator.Render();
    film->WritePNG("motion_blur");

    // the sphere's top is at y = 0.6 when the shutter opens and at y = 0.8 when it closes
    SurfaceInteraction s0, s1;
    const bool h0 = scene->Intersect(Ray(glm::vec3(0, 3.0f, -1.2f), glm::vec3(0, -1, 0), 0.0f), s0);
    const bool h1 = scene->Intersect(Ray(glm::vec3(0, 3.0f, -1.2f), glm::vec3(0, -1, 0), 1.0f), s1);
    // a ray along the ellipsoid's long axis leaves it 0.6 from its centre
    SurfaceInteraction s2;
    const glm::vec3 axis(std::cos(0.6f), std::sin(0.6f), 0.0f);
    const bool h2 = scene->Intersect(Ray(glm::vec3(1.2f, 0.1f, -0.8f) + 3.0f * axis, -axis, 0.0f), s2);
    printf("{\"samples\": %llu, \"shadow_rays\": %llu, \"hit0\": %d, \"t0\": %.7g, \"hit1\": %d, \"t1\": %.7g, \"hit2\": %d, \"t2\": %.7g, \"inst2\": %u}\n",
           (unsigned long long)integrator.LastStats().samples, (unsigned long long)integrator.LastStats().shadow_rays,
           h0 ? 1 : 0, s0.t, h1 ? 1 : 0, s1.t, h2 ? 1 : 0, s2.t, s2.instance);
    return 0;
}
