// cornell_box.cpp — the structure of the reference's examples/example_1.cpp (scene -> textures ->
// materials -> meshes -> primitives -> TLAS -> film -> camera -> sampler -> integrators -> render ->
// write) written against the drop-in host API, rendering BASELINE config C1: a Cornell box of triangle
// meshes with a two-triangle area light, NEE + MIS, depth 8.
//
//   g++ -std=c++20 -Iinclude -Ipathtracing_b200/host -Ipathtracing_b200/host/glm_min examples/cornell_box.cpp
//     -Lpathtracing_b200 -lptb200 -Wl,-rpath,$PWD/pathtracing_b200 -o cornell_box
//   ./cornell_box [width height spp]      -> cornell_path.ppm + one JSON line with the mean radiance
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <vector>

#include "Material.hpp"
#include "Primitive.hpp"
#include "Model.hpp"
#include "Light.hpp"
#include "LightSampler.hpp"
#include "Sampler.hpp"
#include "Filter.hpp"
#include "Film.hpp"
#include "Camera.hpp"
#include "Scene.hpp"
#include "Integrators.hpp"
#include <glm/gtc/matrix_transform.hpp>

// a parallelogram as two triangles: origin, edge u, edge v; normal = cross(u, v)
static std::shared_ptr<Mesh> Quad(const glm::vec3& o, const glm::vec3& u, const glm::vec3& v, const std::shared_ptr<Material>& mat,
                                  const std::shared_ptr<Texture>& emissive = nullptr){
    const glm::vec3 n = glm::normalize(glm::cross(u, v));
    return std::make_shared<Mesh>(std::vector<uint32_t> { 0, 1, 2, 0, 2, 3 }, std::vector<glm::vec3> { o, o + u, o + u + v, o + v },
                                  std::vector<glm::vec3> {}, std::vector<glm::vec3> { n, n, n, n },
                                  std::vector<glm::vec2> { { 0, 0 }, { 1, 0 }, { 1, 1 }, { 0, 1 } }, mat, emissive, nullptr);
}

// an axis-aligned box [lo,hi] as six outward-facing quads
static void Box(std::vector<std::shared_ptr<Mesh>>& out, const glm::vec3& lo, const glm::vec3& hi, const std::shared_ptr<Material>& mat){
    const glm::vec3 d = hi - lo, X(d.x, 0, 0), Y(0, d.y, 0), Z(0, 0, d.z);
    out.push_back(Quad(lo, Y, X, mat));                 // -z
    out.push_back(Quad(lo + Z, X, Y, mat));             // +z
    out.push_back(Quad(lo, Z, Y, mat));                 // -x
    out.push_back(Quad(lo + X, Y, Z, mat));             // +x
    out.push_back(Quad(lo, X, Z, mat));                 // -y
    out.push_back(Quad(lo + Y, Z, X, mat));             // +y
}

int main(int argc, char** argv){
    const int width = argc > 2 ? atoi(argv[1]) : 256, height = argc > 2 ? atoi(argv[2]) : 256;
    const unsigned spp = argc > 3 ? (unsigned)atoi(argv[3]) : 16;

    // 1) Scene
    auto scene = std::make_shared<Scene>();

    // 2) Textures and 3) materials
    std::shared_ptr<Texture> whiteTex = std::make_shared<SolidColor>(glm::vec3(.73f));
    std::shared_ptr<Material> white = std::make_shared<MicrofacetDiffuse>(whiteTex);
    std::shared_ptr<Material> red = std::make_shared<MicrofacetDiffuse>(glm::vec3(.65f, .05f, .05f));
    std::shared_ptr<Material> green = std::make_shared<MicrofacetDiffuse>(glm::vec3(.12f, .45f, .15f));
    std::shared_ptr<Material> black = std::make_shared<MicrofacetDiffuse>(glm::vec3(0));
    std::shared_ptr<Texture> emission = std::make_shared<SolidColor>(glm::vec3(1.0f, 0.85f, 0.6f), glm::vec3(15.0f));

    // 4) Meshes (the reference's Model imports them from a file; here they are generated)
    std::vector<std::shared_ptr<Mesh>> room;
    room.push_back(Quad({ 0, 0, 0 }, { 0, 0, 1 }, { 1, 0, 0 }, white));      // floor
    room.push_back(Quad({ 0, 1, 0 }, { 1, 0, 0 }, { 0, 0, 1 }, white));      // ceiling
    room.push_back(Quad({ 0, 0, 0 }, { 1, 0, 0 }, { 0, 1, 0 }, white));      // back wall
    room.push_back(Quad({ 0, 0, 0 }, { 0, 1, 0 }, { 0, 0, 1 }, red));        // left
    room.push_back(Quad({ 1, 0, 0 }, { 0, 0, 1 }, { 0, 1, 0 }, green));      // right
    room.push_back(Quad({ 0.35f, 0.998f, 0.35f }, { 0.3f, 0, 0 }, { 0, 0, 0.3f }, black, emission));   // area light (emissive mesh)
    std::vector<std::shared_ptr<Mesh>> block;
    Box(block, { -0.15f, 0.0f, -0.15f }, { 0.15f, 0.6f, 0.15f }, white);

    // 5) Primitives: one BLAS for the room, one for the block, instanced twice through TransformedPrimitive
    auto roomModel = std::make_shared<Model>(room);
    roomModel->BuildBlas<BLAS4>();
    auto blockModel = std::make_shared<Model>(block);
    blockModel->BuildBlas<BLAS4>();
    glm::mat4 tall = glm::rotate(glm::translate(glm::mat4(1.0f), glm::vec3(0.3f, 0.0f, 0.6f)), 0.35f, glm::vec3(0, 1, 0));
    glm::mat4 small = glm::scale(glm::rotate(glm::translate(glm::mat4(1.0f), glm::vec3(0.68f, 0.0f, 0.25f)), -0.3f, glm::vec3(0, 1, 0)), glm::vec3(1.0f, 0.5f, 1.0f));

    // 6) Add primitives to the scene, 8) build the TLAS
    scene->Add(roomModel);
    scene->Add(std::make_shared<TransformedPrimitive>(blockModel, tall));
    scene->Add(std::make_shared<TransformedPrimitive>(blockModel, small));
    scene->BuildTlas<TLAS4>();

    // 9) Film + filter, 10) camera (fov in radians)
    auto film = std::make_shared<Film>(glm::ivec2 { width, height }, std::make_shared<BoxFilter>());
    auto camera = std::make_shared<Camera>(glm::vec3(0.5f, 0.5f, 2.4f), glm::vec3(0.5f, 0.5f, 0.0f), 2.0f * std::atan(0.5f / 1.4f), film);

    // 11) Sampler + 12) light sampler + integrators
    auto sampler = std::make_shared<UniformSampler>(spp);
    auto lightSampler = std::make_shared<PowerLightSampler>();
    lightSampler->Add(scene->GetLights());
    lightSampler->PreProcess(scene->BoundingBox());
    auto pathIntegrator = std::make_shared<PathIntegrator>(scene, camera, sampler, lightSampler, 8);
    auto simpleIntegrator = std::make_shared<SimplePathIntegrator>(scene, camera, sampler, 8);
    pathIntegrator->SetAdaptive(false);       // exactly spp samples per pixel, so the two means below are comparable
    simpleIntegrator->SetAdaptive(false);

    // 13) Render
    auto mean = [&](){
        glm::dvec3 m(0.0);
        for(int y = 0; y < height; y++)for(int x = 0; x < width; x++)m += film->Pixel(x, y);
        return m * (1.0 / ((double)width * height));
    };
    pathIntegrator->Render();
    film->WritePPM("cornell_path");
    const glm::dvec3 mp = mean();
    const pt_render_stats st = pathIntegrator->LastStats();
    film->Clear();
    simpleIntegrator->Render();
    const glm::dvec3 ms = mean();

    // a ray down the middle of the box hits the back wall at z = 0
    SurfaceInteraction si;
    const bool hit = scene->Intersect(Ray(glm::vec3(0.5f, 0.9f, 2.4f), glm::vec3(0, 0, -1)), si);

    printf("{\"path_mean\": [%.6f, %.6f, %.6f], \"simple_mean\": [%.6f, %.6f, %.6f], \"samples\": %llu, \"gpu_ms\": %.3f, \"hit\": %d, \"hit_t\": %.6f, \"hit_instance\": %u}\n",
           mp.x, mp.y, mp.z, ms.x, ms.y, ms.z, (unsigned long long)st.samples, st.gpu_ms, hit ? 1 : 0, si.t, si.instance);
    return 0;
}
