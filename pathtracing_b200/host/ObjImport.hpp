// ObjImport.hpp — Wavefront OBJ / MTL import behind Model(path) and ResourceManager::CacheModel<BLASx>(name, path)
// (SURVEY §8f rank 4).  The reference imports through assimp (Model.cpp:14-357), an un-vendored submodule that
// is not part of this path; this is a from-scratch reader of the one text format its scenes can be exported
// to, applying the same post-processing the reference asks assimp for (Model.cpp:36-52): triangulation,
// joined identical vertices, one mesh per material, flipped V, smooth normals where the file has none,
// per-vertex tangents where it has texture coordinates — and the same material mapping as
// Model::SetupMaterial (Model.cpp:123-276) for what an .mtl can express.  Host I/O only: no kernel work.
//
// Images: `map_Kd` / `map_Ke` / `map_Bump` / `map_d` are read from binary PPM (P6) files; other formats are
// the job of stb_image in the reference (Texture.cpp:3-12) and are skipped with a warning here.
// Extension: a material with `Ke` but no `map_Ke` emits with SolidColor(Ke) (the reference only honours an
// emissive texture, Model.cpp:343).
#pragma once
// included at the end of PathTracing.hpp

#include <map>
#include <sstream>
#include <tuple>

namespace ptb_host {

struct ObjMaterial {
    std::string name;
    glm::vec3 Kd { 0.6f, 0.6f, 0.6f }, Ks { 0.0f, 0.0f, 0.0f }, Ke { 0.0f, 0.0f, 0.0f };
    float Ni = 1.0f, d = 1.0f, Pr = -1.0f, Pm = -1.0f;      // assimp's OBJ defaults: ior 1, opacity 1
    std::string map_Kd, map_Ke, map_Bump, map_d;
};

inline std::string obj_dir(const std::string& path){
    const size_t k = path.find_last_of("/\\");
    return k == std::string::npos ? std::string() : path.substr(0, k + 1);
}

// binary PPM (P6, maxval 255) -> RGB8 texels; empty on failure
inline std::vector<uint8_t> read_ppm(const std::string& file, int& w, int& h){
    std::ifstream in(file, std::ios::binary);
    std::vector<uint8_t> px;
    std::string magic;
    int maxv = 0;
    auto next_int = [&](int& v){
        while(true){
            int c = in.peek();
            if(c == '#'){ std::string skip; std::getline(in, skip); }
            else if(c == ' ' || c == '\n' || c == '\r' || c == '\t')in.get();
            else break;
        }
        in >> v;
    };
    in >> magic;
    if(!in || magic != "P6")return px;
    next_int(w); next_int(h); next_int(maxv);
    in.get();
    if(!in || w <= 0 || h <= 0 || maxv != 255)return px;
    px.resize((size_t)w * h * 3);
    in.read((char*)px.data(), (std::streamsize)px.size());
    if(!in)px.clear();
    return px;
}

inline std::shared_ptr<Texture> obj_image(const std::string& dir, const std::string& name){
    if(name.empty())return nullptr;
    int w = 0, h = 0;
    std::vector<uint8_t> px = read_ppm(dir + name, w, h);
    if(px.empty()){
        std::fprintf(stderr, "[pathtracing_b200] OBJ import: texture '%s' skipped (only binary PPM is read here)\n", name.c_str());
        return nullptr;
    }
    return std::make_shared<ImageTexture>(std::move(px), w, h, 3);
}

inline std::map<std::string, ObjMaterial> read_mtl(const std::string& file){
    std::map<std::string, ObjMaterial> out;
    std::ifstream in(file);
    if(!in){ std::fprintf(stderr, "[pathtracing_b200] OBJ import: cannot open material library '%s'\n", file.c_str()); return out; }
    std::string line;
    ObjMaterial* cur = nullptr;
    auto last_token = [](std::istringstream& ss){ std::string t, last; while(ss >> t)last = t; return last; };   // map_* options precede the file name
    while(std::getline(in, line)){
        std::istringstream ss(line);
        std::string key;
        if(!(ss >> key) || key[0] == '#')continue;
        if(key == "newmtl"){ std::string n; ss >> n; cur = &out[n]; cur->name = n; continue; }
        if(!cur)continue;
        if(key == "Kd")ss >> cur->Kd.x >> cur->Kd.y >> cur->Kd.z;
        else if(key == "Ks")ss >> cur->Ks.x >> cur->Ks.y >> cur->Ks.z;
        else if(key == "Ke")ss >> cur->Ke.x >> cur->Ke.y >> cur->Ke.z;
        else if(key == "Ni")ss >> cur->Ni;
        else if(key == "d")ss >> cur->d;
        else if(key == "Tr"){ float tr = 0; ss >> tr; cur->d = 1.0f - tr; }
        else if(key == "Pr")ss >> cur->Pr;
        else if(key == "Pm")ss >> cur->Pm;
        else if(key == "map_Kd")cur->map_Kd = last_token(ss);
        else if(key == "map_Ke")cur->map_Ke = last_token(ss);
        else if(key == "map_Bump" || key == "map_bump" || key == "bump" || key == "norm")cur->map_Bump = last_token(ss);
        else if(key == "map_d")cur->map_d = last_token(ss);
    }
    return out;
}

// Model::SetupMaterial (Model.cpp:123-276) for the properties assimp's OBJ importer provides: no transmission
// factor; a metallic factor when the .mtl has `Pm`; otherwise the Kd / Ks / Ni / d chain of :222-271.
inline std::shared_ptr<Material> obj_material(const ObjMaterial& m, const std::string& dir, std::shared_ptr<Texture>& emissive){
    std::shared_ptr<Texture> albedo = obj_image(dir, m.map_Kd);
    const bool has_diffuse_tex = albedo != nullptr;
    if(!albedo)albedo = std::make_shared<SolidColor>(glm::vec3(1.0f));                 // base colour default (Model.cpp:139,166)
    std::shared_ptr<Texture> normal = obj_image(dir, m.map_Bump);
    std::shared_ptr<Texture> alpha = obj_image(dir, m.map_d);
    emissive = obj_image(dir, m.map_Ke);
    if(!emissive && (m.Ke.x > 0.0f || m.Ke.y > 0.0f || m.Ke.z > 0.0f))emissive = std::make_shared<SolidColor>(m.Ke);
    if(m.Pm >= 0.0f)return std::make_shared<MicrofacetDiffuse>(albedo, normal, nullptr, nullptr, alpha);   // :185-189
    if(has_diffuse_tex)return std::make_shared<MicrofacetDiffuse>(albedo, normal, nullptr, nullptr, alpha); // :194-202
    const float kdLum = (m.Kd.x + m.Kd.y + m.Kd.z) / 3.0f, ksLum = (m.Ks.x + m.Ks.y + m.Ks.z) / 3.0f;
    if(m.d < 0.99f)return std::make_shared<MicrofacetDielectric>(1.5f, 0.0f, glm::vec3(1.0f));              // :241-243
    if(ksLum > 0.1f && ksLum >= 0.4f)                                                                      // :244-249
        return std::make_shared<SpecularConductor>(glm::vec3((float)linear_to_sRGB(m.Ks.x), (float)linear_to_sRGB(m.Ks.y), (float)linear_to_sRGB(m.Ks.z)));
    if(kdLum < ksLum)return std::make_shared<MicrofacetDielectric>(1.33f, 0.0f, glm::vec3(0.98f, 1.0f, 1.0f));
    if(ksLum > 0.1f)return std::make_shared<MicrofacetDielectric>(1.5f, 0.0f, glm::vec3(1.0f));
    if(kdLum > 0.1f && ksLum < 0.03f)return std::make_shared<MicrofacetDielectric>(1.33f, 0.0f, m.Ks + m.Kd); // :255-261
    return std::make_shared<MicrofacetDiffuse>(std::make_shared<SolidColor>(m.Kd));                        // :263-267
}

// one mesh per material, vertices joined on identical (v, vt, vn) triples
struct ObjGroup {
    std::vector<uint32_t> indices;
    std::vector<glm::vec3> pos, nrm;
    std::vector<glm::vec2> uv;
    std::map<std::tuple<int, int, int>, uint32_t> seen;
    bool has_uv = false;
};

inline std::vector<std::shared_ptr<Mesh>> load_obj(const std::string& path){
    std::ifstream in(path);
    if(!in)throw std::runtime_error("Model: cannot open '" + path + "'");
    const std::string dir = obj_dir(path);
    std::vector<glm::vec3> V, N;
    std::vector<glm::vec2> T;
    std::map<std::string, ObjMaterial> mtl;
    std::vector<std::string> order;                 // materials in order of first use
    std::map<std::string, ObjGroup> groups;
    std::string current = "";
    std::string line;
    struct Corner { int v, t, n; };
    auto fix = [](int i, size_t n){ return i > 0 ? i - 1 : i < 0 ? (int)n + i : -1; };   // 1-based, negative = relative
    while(std::getline(in, line)){
        std::istringstream ss(line);
        std::string key;
        if(!(ss >> key) || key[0] == '#')continue;
        if(key == "v"){ glm::vec3 p(0.0f); ss >> p.x >> p.y >> p.z; V.push_back(p); }
        else if(key == "vn"){ glm::vec3 p(0.0f); ss >> p.x >> p.y >> p.z; N.push_back(p); }
        else if(key == "vt"){ glm::vec2 p(0.0f); ss >> p.x >> p.y; T.push_back(glm::vec2(p.x, 1.0f - p.y)); }   // aiProcess_FlipUVs
        else if(key == "mtllib"){ std::string f; ss >> f; for(auto& kv : read_mtl(dir + f))mtl[kv.first] = kv.second; }
        else if(key == "usemtl"){ ss >> current; }
        else if(key == "f"){
            std::vector<Corner> poly;
            std::string tok;
            while(ss >> tok){
                Corner c { 0, 0, 0 };
                const size_t a = tok.find('/');
                c.v = std::atoi(tok.substr(0, a).c_str());
                if(a != std::string::npos){
                    const size_t b = tok.find('/', a + 1);
                    const std::string t = tok.substr(a + 1, b == std::string::npos ? std::string::npos : b - a - 1);
                    if(!t.empty())c.t = std::atoi(t.c_str());
                    if(b != std::string::npos)c.n = std::atoi(tok.substr(b + 1).c_str());
                }
                c.v = fix(c.v, V.size()); c.t = fix(c.t, T.size()); c.n = fix(c.n, N.size());
                if(c.v < 0 || c.v >= (int)V.size())throw std::runtime_error("Model: vertex index out of range in '" + path + "'");
                if(c.t >= (int)T.size())c.t = -1;
                if(c.n >= (int)N.size())c.n = -1;
                poly.push_back(c);
            }
            if(poly.size() < 3)continue;                // aiProcess_SortByPType: points / lines are not triangles
            if(!groups.count(current))order.push_back(current);
            ObjGroup& g = groups[current];
            auto vertex = [&](const Corner& c) -> uint32_t{
                const auto key3 = std::make_tuple(c.v, c.t, c.n);
                auto it = g.seen.find(key3);
                if(it != g.seen.end())return it->second;
                const uint32_t id = (uint32_t)g.pos.size();
                g.pos.push_back(V[c.v]);
                g.nrm.push_back(c.n >= 0 ? N[c.n] : glm::vec3(0.0f));
                g.uv.push_back(c.t >= 0 ? T[c.t] : glm::vec2(0.0f));
                g.has_uv = g.has_uv || c.t >= 0;
                g.seen.emplace(key3, id);
                return id;
            };
            for(size_t k = 1; k + 1 < poly.size(); k++){   // fan triangulation
                g.indices.push_back(vertex(poly[0])); g.indices.push_back(vertex(poly[k])); g.indices.push_back(vertex(poly[k + 1]));
            }
        }
    }
    std::vector<std::shared_ptr<Mesh>> meshes;
    for(const std::string& name : order){
        ObjGroup& g = groups[name];
        const size_t nv = g.pos.size();
        // aiProcess_GenSmoothNormals: vertices without a normal get the area-weighted mean of their faces' normals
        std::vector<glm::vec3> acc(nv, glm::vec3(0.0f));
        std::vector<glm::vec3> tan(g.has_uv ? nv : 0, glm::vec3(0.0f));
        for(size_t f = 0; f + 2 < g.indices.size(); f += 3){
            const uint32_t i0 = g.indices[f], i1 = g.indices[f + 1], i2 = g.indices[f + 2];
            const glm::vec3 e1 = g.pos[i1] - g.pos[i0], e2 = g.pos[i2] - g.pos[i0];
            const glm::vec3 fn = glm::cross(e1, e2);
            acc[i0] = acc[i0] + fn; acc[i1] = acc[i1] + fn; acc[i2] = acc[i2] + fn;
            if(g.has_uv){                               // aiProcess_CalcTangentSpace
                const glm::vec2 d1 = g.uv[i1] - g.uv[i0], d2 = g.uv[i2] - g.uv[i0];
                const float det = d1.x * d2.y - d2.x * d1.y;
                if(std::fabs(det) > 1e-20f){
                    const glm::vec3 t = (e1 * d2.y - e2 * d1.y) * (1.0f / det);
                    tan[i0] = tan[i0] + t; tan[i1] = tan[i1] + t; tan[i2] = tan[i2] + t;
                }
            }
        }
        for(size_t i = 0; i < nv; i++){
            glm::vec3& n = g.nrm[i];
            if(glm::dot(n, n) == 0.0f)n = acc[i];
            const float l2 = glm::dot(n, n);
            n = l2 > 0.0f ? n * (1.0f / std::sqrt(l2)) : glm::vec3(0.0f, 1.0f, 0.0f);
            if(g.has_uv){
                glm::vec3& t = tan[i];
                t = t - n * glm::dot(n, t);
                const float t2 = glm::dot(t, t);
                t = t2 > 0.0f ? t * (1.0f / std::sqrt(t2)) : glm::vec3(0.0f);
            }
        }
        std::shared_ptr<Texture> emissive;
        std::shared_ptr<Material> mat;
        auto it = mtl.find(name);
        if(it != mtl.end())mat = obj_material(it->second, dir, emissive);
        else mat = std::make_shared<MicrofacetDiffuse>(std::make_shared<SolidColor>(glm::vec3(0.65f, 0.05f, 0.05f)));   // Model.cpp:354-355
        meshes.push_back(std::make_shared<Mesh>(g.indices, g.pos, tan, g.nrm, g.uv, mat, emissive, nullptr));
    }
    if(meshes.empty())std::fprintf(stderr, "[pathtracing_b200] OBJ import: '%s' holds no faces\n", path.c_str());
    return meshes;
}

} // namespace ptb_host
