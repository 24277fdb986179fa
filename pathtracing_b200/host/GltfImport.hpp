// GltfImport.hpp — glTF 2.0 (.gltf + .bin / data URIs, and .glb) import behind Model(path).
// The reference imports through assimp (Model.cpp:22-71), which is not part of this path; this is a from-scratch reader of
// the format the reference's `opacity()` / `transmission()` / `knight()` / helmet scenes use (main.cpp:475-830), producing
// what the reference's post-processing flags ask assimp for (Model.cpp:44-58): triangles only, node transforms baked into the
// vertices (aiProcess_PreTransformVertices), flipped V (aiProcess_FlipUVs), smooth normals where the file has none
// (aiProcess_GenSmoothNormals), tangents where it has texture coordinates but no TANGENT attribute (aiProcess_CalcTangentSpace),
// one Mesh per primitive -- and the material mapping of Model::GetTextures / SetupMaterial / process_mesh (Model.cpp:85-357) for
// the glTF material model: base colour (factor as the texture's colour scale, sRGB -> linear), normal map, the metallic-roughness
// texture as both roughness (.g) and metallic (.b) source, emissive colour / texture x KHR_materials_emissive_strength, alphaMode /
// alphaCutoff, KHR_materials_transmission (-> MicrofacetDielectric, or ThinDielectric when KHR_materials_volume says thickness 0),
// KHR_materials_ior, KHR_materials_volume attenuation (-> HomogeneusMedium inside the mesh).  Host I/O only: no kernel work.
// Images: PNG (also embedded in a .glb / data URI), PPM, BMP, TGA through ImageIO.hpp; a JPEG texture is skipped with a warning.
#pragma once
// included at the end of PathTracing.hpp

#include <cctype>
#include <map>

namespace ptb_host {

// ---- a small JSON value + recursive-descent parser (RFC 8259; numbers as double) ------------------------------------------
struct Json {
    enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
    bool b = false; double num = 0; std::string str;
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;
    const Json* get(const std::string& k) const{ for(const auto& kv : obj)if(kv.first == k)return &kv.second; return nullptr; }
    const Json& at(const std::string& k) const{ const Json* p = get(k); if(!p)throw std::runtime_error("glTF: missing key '" + k + "'"); return *p; }
    double number(const std::string& k, double dflt) const{ const Json* p = get(k); return p && p->kind == Num ? p->num : dflt; }
    int integer(const std::string& k, int dflt) const{ return (int)number(k, dflt); }
    std::string string(const std::string& k, const std::string& dflt = "") const{ const Json* p = get(k); return p && p->kind == Str ? p->str : dflt; }
    size_t size() const{ return kind == Arr ? arr.size() : 0; }
};

struct JsonParser {
    const char* p; const char* end;
    void ws(){ while(p < end && (*p == ' ' || *p == '\n' || *p == '\r' || *p == '\t'))p++; }
    [[noreturn]] void fail(const char* what){ throw std::runtime_error(std::string("glTF JSON: ") + what); }
    Json value(){
        ws();
        if(p >= end)fail("unexpected end");
        Json j;
        if(*p == '{'){
            p++; j.kind = Json::Obj; ws();
            if(p < end && *p == '}'){ p++; return j; }
            for(;;){
                ws();
                if(p >= end || *p != '"')fail("expected a key");
                std::string k = string();
                ws();
                if(p >= end || *p != ':')fail("expected ':'");
                p++;
                j.obj.emplace_back(std::move(k), value());
                ws();
                if(p < end && *p == ','){ p++; continue; }
                if(p < end && *p == '}'){ p++; return j; }
                fail("expected ',' or '}'");
            }
        }
        if(*p == '['){
            p++; j.kind = Json::Arr; ws();
            if(p < end && *p == ']'){ p++; return j; }
            for(;;){
                j.arr.push_back(value());
                ws();
                if(p < end && *p == ','){ p++; continue; }
                if(p < end && *p == ']'){ p++; return j; }
                fail("expected ',' or ']'");
            }
        }
        if(*p == '"'){ j.kind = Json::Str; j.str = string(); return j; }
        if(end - p >= 4 && !std::strncmp(p, "true", 4)){ p += 4; j.kind = Json::Bool; j.b = true; return j; }
        if(end - p >= 5 && !std::strncmp(p, "false", 5)){ p += 5; j.kind = Json::Bool; return j; }
        if(end - p >= 4 && !std::strncmp(p, "null", 4)){ p += 4; return j; }
        char* e = nullptr;
        j.num = std::strtod(p, &e);
        if(e == p)fail("bad value");
        p = e; j.kind = Json::Num;
        return j;
    }
    std::string string(){
        std::string s;
        p++;
        while(p < end && *p != '"'){
            if(*p == '\\' && p + 1 < end){
                p++;
                switch(*p){
                case 'n': s.push_back('\n'); break; case 't': s.push_back('\t'); break; case 'r': s.push_back('\r'); break;
                case 'b': s.push_back('\b'); break; case 'f': s.push_back('\f'); break;
                case 'u': {
                    if(end - p < 5)fail("bad \\u escape");
                    unsigned cp = (unsigned)std::strtoul(std::string(p + 1, p + 5).c_str(), nullptr, 16);
                    p += 4;
                    if(cp < 0x80)s.push_back((char)cp);
                    else if(cp < 0x800){ s.push_back((char)(0xC0 | cp >> 6)); s.push_back((char)(0x80 | (cp & 63))); }
                    else{ s.push_back((char)(0xE0 | cp >> 12)); s.push_back((char)(0x80 | ((cp >> 6) & 63))); s.push_back((char)(0x80 | (cp & 63))); }
                    break;
                }
                default: s.push_back(*p);
                }
                p++;
            } else s.push_back(*p++);
        }
        if(p >= end)fail("unterminated string");
        p++;
        return s;
    }
};

inline std::vector<uint8_t> base64_decode(const std::string& s, size_t from){
    std::vector<uint8_t> out;
    uint32_t acc = 0; int bits = 0;
    for(size_t i = from; i < s.size(); i++){
        const char c = s[i];
        int v;
        if(c >= 'A' && c <= 'Z')v = c - 'A'; else if(c >= 'a' && c <= 'z')v = c - 'a' + 26; else if(c >= '0' && c <= '9')v = c - '0' + 52;
        else if(c == '+' || c == '-')v = 62; else if(c == '/' || c == '_')v = 63; else continue;
        acc = acc << 6 | (uint32_t)v; bits += 6;
        if(bits >= 8){ bits -= 8; out.push_back((uint8_t)(acc >> bits)); }
    }
    return out;
}

struct GltfDoc {
    Json root;
    std::string dir;
    std::vector<std::vector<uint8_t>> buffers;

    const uint8_t* view(int accessor, int& count, int& comp_type, int& ncomp, size_t& stride, bool& normalized) const{
        const Json& a = root.at("accessors").arr.at((size_t)accessor);
        count = a.integer("count", 0); comp_type = a.integer("componentType", 5126);
        const std::string t = a.string("type", "SCALAR");
        ncomp = t == "SCALAR" ? 1 : t == "VEC2" ? 2 : t == "VEC3" ? 3 : t == "VEC4" ? 4 : t == "MAT4" ? 16 : 0;
        const Json* nz = a.get("normalized");
        normalized = nz && nz->kind == Json::Bool && nz->b;
        if(a.get("sparse"))throw std::runtime_error("glTF: sparse accessors are not supported");
        const int bv = a.integer("bufferView", -1);
        if(bv < 0 || ncomp == 0)throw std::runtime_error("glTF: accessor without a buffer view");
        const Json& v = root.at("bufferViews").arr.at((size_t)bv);
        const std::vector<uint8_t>& buf = buffers.at((size_t)v.integer("buffer", 0));
        const size_t csize = comp_type == 5120 || comp_type == 5121 ? 1 : comp_type == 5122 || comp_type == 5123 ? 2 : 4;
        stride = (size_t)v.integer("byteStride", 0);
        if(stride == 0)stride = csize * (size_t)ncomp;
        const size_t off = (size_t)v.number("byteOffset", 0) + (size_t)a.number("byteOffset", 0);
        if(count > 0 && off + stride * (size_t)(count - 1) + csize * (size_t)ncomp > buf.size())throw std::runtime_error("glTF: accessor reads beyond its buffer");
        return buf.data() + off;
    }
    std::vector<float> floats(int accessor, int want_comp) const{
        int count, ct, nc; size_t stride; bool norm;
        const uint8_t* p = view(accessor, count, ct, nc, stride, norm);
        if(nc < want_comp)throw std::runtime_error("glTF: attribute has too few components");
        std::vector<float> out((size_t)count * want_comp);
        for(int i = 0; i < count; i++)for(int c = 0; c < want_comp; c++){
            const uint8_t* q = p + stride * (size_t)i;
            float v;
            switch(ct){
            case 5126: std::memcpy(&v, q + 4 * c, 4); break;
            case 5121: v = norm ? q[c] / 255.0f : (float)q[c]; break;
            case 5120: v = norm ? std::max(-1.0f, (int8_t)q[c] / 127.0f) : (float)(int8_t)q[c]; break;
            case 5123: { uint16_t u; std::memcpy(&u, q + 2 * c, 2); v = norm ? u / 65535.0f : (float)u; break; }
            case 5122: { int16_t u; std::memcpy(&u, q + 2 * c, 2); v = norm ? std::max(-1.0f, u / 32767.0f) : (float)u; break; }
            default: throw std::runtime_error("glTF: unsupported attribute component type");
            }
            out[(size_t)i * want_comp + c] = v;
        }
        return out;
    }
    std::vector<uint32_t> indices(int accessor) const{
        int count, ct, nc; size_t stride; bool norm;
        const uint8_t* p = view(accessor, count, ct, nc, stride, norm);
        std::vector<uint32_t> out((size_t)count);
        for(int i = 0; i < count; i++){
            const uint8_t* q = p + stride * (size_t)i;
            if(ct == 5121)out[(size_t)i] = q[0];
            else if(ct == 5123){ uint16_t u; std::memcpy(&u, q, 2); out[(size_t)i] = u; }
            else if(ct == 5125){ uint32_t u; std::memcpy(&u, q, 4); out[(size_t)i] = u; }
            else throw std::runtime_error("glTF: unsupported index component type");
        }
        return out;
    }

    // textures[index] -> images[source]; gamma: the reference loads base colour and emissive images with the sRGB LUT (Model.cpp:112)
    std::shared_ptr<Texture> texture(const Json* info, bool gamma, const glm::vec3& scale){
        if(!info)return nullptr;
        const int ti = info->integer("index", -1);
        const Json* texs = root.get("textures");
        if(ti < 0 || !texs || (size_t)ti >= texs->size())return nullptr;
        const int src = texs->arr[(size_t)ti].integer("source", -1);
        const Json* imgs = root.get("images");
        if(src < 0 || !imgs || (size_t)src >= imgs->size())return nullptr;
        const Json& im = imgs->arr[(size_t)src];
        try{
            ptb_io::ImageU8 px;
            const std::string uri = im.string("uri");
            if(!uri.empty() && uri.compare(0, 5, "data:") == 0){
                const size_t comma = uri.find(',');
                px = decode_memory(base64_decode(uri, comma == std::string::npos ? 0 : comma + 1));
            } else if(!uri.empty())px = ptb_io::load_u8(dir + uri);
            else{
                const Json& v = root.at("bufferViews").arr.at((size_t)im.integer("bufferView", -1));
                const std::vector<uint8_t>& buf = buffers.at((size_t)v.integer("buffer", 0));
                const size_t off = (size_t)v.number("byteOffset", 0), len = (size_t)v.number("byteLength", 0);
                if(off + len > buf.size())throw std::runtime_error("image view beyond its buffer");
                px = decode_memory(std::vector<uint8_t>(buf.begin() + (long)off, buf.begin() + (long)(off + len)));
            }
            if(gamma){
                uint8_t lut[256];
                for(int i = 0; i < 256; i++){
                    const double s = i / 255.0, lin = s <= 0.04045 ? s / 12.92 : std::pow((s + 0.055) / 1.055, 2.4);
                    lut[i] = (uint8_t)std::lround(std::min(1.0, std::max(0.0, lin)) * 255.0);
                }
                for(size_t i = 0; i < px.data.size(); i++)if((int)(i % (size_t)px.channels) < std::min(px.channels, 3))px.data[i] = lut[px.data[i]];
            }
            return std::make_shared<ImageTexture>(std::move(px.data), px.width, px.height, px.channels, scale);
        } catch(const std::exception& e){
            std::fprintf(stderr, "[pathtracing_b200] glTF import: texture %d skipped (%s)\n", ti, e.what());
            return nullptr;
        }
    }
    static ptb_io::ImageU8 decode_memory(const std::vector<uint8_t>& f){
        if(f.size() >= 8 && !std::memcmp(f.data(), "\x89PNG\r\n\x1a\n", 8))return ptb_io::decode_png(f);
        if(f.size() >= 2 && f[0] == 'P' && (f[1] == '5' || f[1] == '6'))return ptb_io::decode_pnm(f);
        if(f.size() >= 2 && f[0] == 'B' && f[1] == 'M')return ptb_io::decode_bmp(f);
        throw std::runtime_error("embedded image is not PNG / PPM / BMP (JPEG decoding is not implemented)");
    }
};

inline glm::mat4 gltf_node_matrix(const Json& n){
    glm::mat4 M(1.0f);
    if(const Json* m = n.get("matrix")){
        if(m->size() == 16)for(int c = 0; c < 4; c++)for(int r = 0; r < 4; r++)M[c][r] = (float)m->arr[(size_t)(4 * c + r)].num;     // column-major, like glm
        return M;
    }
    glm::vec3 t(0.0f), s(1.0f);
    float q[4] = { 0, 0, 0, 1 };
    if(const Json* a = n.get("translation"))if(a->size() == 3)t = glm::vec3((float)a->arr[0].num, (float)a->arr[1].num, (float)a->arr[2].num);
    if(const Json* a = n.get("scale"))if(a->size() == 3)s = glm::vec3((float)a->arr[0].num, (float)a->arr[1].num, (float)a->arr[2].num);
    if(const Json* a = n.get("rotation"))if(a->size() == 4)for(int k = 0; k < 4; k++)q[k] = (float)a->arr[(size_t)k].num;
    const float x = q[0], y = q[1], z = q[2], w = q[3];
    const float R[3][3] = { { 1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w) },
                            { 2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w) },
                            { 2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y) } };
    for(int c = 0; c < 3; c++)for(int r = 0; r < 3; r++)M[c][r] = R[r][c] * s[c];
    M[3][0] = t.x; M[3][1] = t.y; M[3][2] = t.z;
    return M;
}

// Model::GetTextures + SetupMaterial + the medium of process_mesh (Model.cpp:85-357) for a glTF material
inline std::shared_ptr<Material> gltf_material(GltfDoc& doc, const Json& m, std::shared_ptr<Texture>& emissive, std::shared_ptr<Medium>& medium){
    const Json* pbr = m.get("pbrMetallicRoughness");
    glm::vec3 base(1.0f);
    if(pbr)if(const Json* f = pbr->get("baseColorFactor"))if(f->size() >= 3)base = glm::vec3((float)f->arr[0].num, (float)f->arr[1].num, (float)f->arr[2].num);
    const Json* ext = m.get("extensions");
    auto extension = [&](const char* name) -> const Json*{ return ext ? ext->get(name) : nullptr; };
    float intensity = 1.0f;
    if(const Json* e = extension("KHR_materials_emissive_strength"))intensity = (float)e->number("emissiveStrength", 1.0);
    glm::vec3 ecol(0.0f);
    if(const Json* f = m.get("emissiveFactor"))if(f->size() == 3)ecol = glm::vec3((float)f->arr[0].num, (float)f->arr[1].num, (float)f->arr[2].num);
    emissive = nullptr;
    if(!(ecol == glm::vec3(0.0f)) && intensity != 0.0f)emissive = std::make_shared<SolidColor>(ecol, glm::vec3(intensity));        // Model.cpp:103
    if(std::shared_ptr<Texture> t = doc.texture(m.get("emissiveTexture"), true, glm::vec3(intensity)))emissive = t;                // :106-116 overrides it
    std::shared_ptr<Texture> albedo = doc.texture(pbr ? pbr->get("baseColorTexture") : nullptr, true, base);
    if(!albedo)albedo = std::make_shared<SolidColor>(base);
    std::shared_ptr<Texture> normal = doc.texture(m.get("normalTexture"), false, glm::vec3(1.0f));
    std::shared_ptr<Texture> orm = doc.texture(pbr ? pbr->get("metallicRoughnessTexture") : nullptr, false, glm::vec3(1.0f));    // roughness .g, metallic .b
    AlphaTester tester {};
    const std::string mode = m.string("alphaMode", "OPAQUE");                                                                     // assimp always sets it for glTF
    if(mode == "OPAQUE")tester.mode = AlphaMode::Opaque;
    else if(mode == "BLEND")tester.mode = AlphaMode::Blend;
    else if(mode == "MASK"){ tester.mode = AlphaMode::Mask; tester.cutoff = (float)m.number("alphaCutoff", 0.5); }
    float ior = 1.5f;
    if(const Json* e = extension("KHR_materials_ior"))ior = (float)e->number("ior", 1.5);
    medium = nullptr;
    const Json* vol = extension("KHR_materials_volume");
    const float thickness = vol ? (float)vol->number("thicknessFactor", 0.0) : 0.0f;
    if(vol && thickness > 0.0f){                                                                                                  // Model.cpp:343-351
        glm::vec3 col(1.0f);
        if(const Json* c = vol->get("attenuationColor"))if(c->size() == 3)col = glm::vec3((float)c->arr[0].num, (float)c->arr[1].num, (float)c->arr[2].num);
        const float dist = (float)vol->number("attenuationDistance", std::numeric_limits<double>::infinity());
        const glm::vec3 sigma_a(-std::log(col.x) / dist, -std::log(col.y) / dist, -std::log(col.z) / dist);
        medium = std::make_shared<HomogeneusMedium>(sigma_a, glm::vec3(0.0f), std::make_shared<HenyeyGreenstein>(0.0f), 1.0f);
    }
    if(extension("KHR_materials_transmission")){                                                                                  // :176-186
        if(vol && thickness == 0.0f)return std::make_shared<ThinDielectric>(ior, albedo);
        auto d = std::make_shared<MicrofacetDielectric>(ior, albedo, normal, orm);
        d->setAlphaTester(tester);
        return d;
    }
    auto d = std::make_shared<MicrofacetDiffuse>(albedo, normal, orm, orm, nullptr);                                              // :187-190 (glTF always carries a metallic factor)
    d->setAlphaTester(tester);
    return d;
}

inline std::vector<std::shared_ptr<Mesh>> load_gltf(const std::string& path){
    GltfDoc doc;
    doc.dir = obj_dir(path);
    std::vector<uint8_t> file = ptb_io::read_file(path);
    std::vector<uint8_t> glb_bin;
    std::string json;
    if(file.size() >= 20 && !std::memcmp(file.data(), "glTF", 4)){
        auto le32 = [&](size_t o){ return (uint32_t)file[o] | file[o + 1] << 8 | file[o + 2] << 16 | (uint32_t)file[o + 3] << 24; };
        size_t o = 12;
        while(o + 8 <= file.size()){
            const uint32_t len = le32(o), type = le32(o + 4);
            if(o + 8 + len > file.size())throw std::runtime_error("glb: truncated chunk");
            if(type == 0x4E4F534A)json.assign((const char*)&file[o + 8], len);
            else if(type == 0x004E4942)glb_bin.assign(file.begin() + (long)(o + 8), file.begin() + (long)(o + 8 + len));
            o += 8 + ((len + 3) & ~3u);
        }
    } else json.assign((const char*)file.data(), file.size());
    JsonParser jp { json.data(), json.data() + json.size() };
    doc.root = jp.value();
    if(const Json* bufs = doc.root.get("buffers"))for(size_t i = 0; i < bufs->size(); i++){
        const std::string uri = bufs->arr[i].string("uri");
        if(uri.empty())doc.buffers.push_back(glb_bin);
        else if(uri.compare(0, 5, "data:") == 0){ const size_t c = uri.find(','); doc.buffers.push_back(base64_decode(uri, c == std::string::npos ? 0 : c + 1)); }
        else doc.buffers.push_back(ptb_io::read_file(doc.dir + uri));
    }
    std::vector<std::shared_ptr<Mesh>> meshes;
    const Json* nodes = doc.root.get("nodes");
    const Json* jm = doc.root.get("meshes");
    const Json* mats = doc.root.get("materials");
    if(!nodes || !jm)return meshes;
    // one Material (and emissive texture, medium) per glTF material, shared by its primitives (aiProcess_RemoveRedundantMaterials)
    struct Mat { std::shared_ptr<Material> m; std::shared_ptr<Texture> e; std::shared_ptr<Medium> med; };
    std::map<int, Mat> made;
    auto material = [&](int idx) -> Mat&{
        auto it = made.find(idx);
        if(it != made.end())return it->second;
        Mat x;
        if(mats && idx >= 0 && (size_t)idx < mats->size())x.m = gltf_material(doc, mats->arr[(size_t)idx], x.e, x.med);
        else x.m = std::make_shared<MicrofacetDiffuse>(std::make_shared<SolidColor>(glm::vec3(0.65f, 0.05f, 0.05f)));              // Model.cpp:354-355
        return made.emplace(idx, std::move(x)).first->second;
    };
    std::function<void(int, const glm::mat4&, int)> walk = [&](int ni, const glm::mat4& parent, int depth){
        if(ni < 0 || (size_t)ni >= nodes->size() || depth > 256)return;
        const Json& n = nodes->arr[(size_t)ni];
        const glm::mat4 M = parent * gltf_node_matrix(n);
        const int mi = n.integer("mesh", -1);
        if(mi >= 0 && (size_t)mi < jm->size()){
            const Json* prims = jm->arr[(size_t)mi].get("primitives");
            for(size_t pi = 0; prims && pi < prims->size(); pi++){
                const Json& pr = prims->arr[pi];
                if(pr.integer("mode", 4) != 4)continue;                                   // aiProcess_SortByPType: triangles only
                const Json* at = pr.get("attributes");
                if(!at || !at->get("POSITION"))continue;
                const std::vector<float> P = doc.floats(at->integer("POSITION", -1), 3);
                const size_t nv = P.size() / 3;
                std::vector<float> N, T, U;
                if(at->get("NORMAL"))N = doc.floats(at->integer("NORMAL", -1), 3);
                if(at->get("TANGENT"))T = doc.floats(at->integer("TANGENT", -1), 3);
                if(at->get("TEXCOORD_0"))U = doc.floats(at->integer("TEXCOORD_0", -1), 2);
                std::vector<uint32_t> idx;
                if(pr.get("indices"))idx = doc.indices(pr.integer("indices", -1));
                else{ idx.resize(nv); for(size_t i = 0; i < nv; i++)idx[i] = (uint32_t)i; }
                idx.resize(idx.size() / 3 * 3);
                for(uint32_t i : idx)if(i >= nv)throw std::runtime_error("glTF: index beyond the vertex count");
                // aiProcess_PreTransformVertices: positions by M, normals by the inverse transpose, tangents by the linear part
                float L[3][3], NM[3][3];                                    // L[r][c]: linear part of M; NM = inverse transpose (cofactors / det)
                for(int c = 0; c < 3; c++)for(int r = 0; r < 3; r++)L[r][c] = M[c][r];
                const float det = L[0][0] * (L[1][1] * L[2][2] - L[1][2] * L[2][1]) - L[0][1] * (L[1][0] * L[2][2] - L[1][2] * L[2][0]) + L[0][2] * (L[1][0] * L[2][1] - L[1][1] * L[2][0]);
                const float id = det != 0.0f ? 1.0f / det : 0.0f;
                for(int r = 0; r < 3; r++)for(int c = 0; c < 3; c++){
                    const int r1 = (r + 1) % 3, r2 = (r + 2) % 3, c1 = (c + 1) % 3, c2 = (c + 2) % 3;
                    NM[r][c] = (L[r1][c1] * L[r2][c2] - L[r1][c2] * L[r2][c1]) * id;
                }
                auto mul3 = [](const float A[3][3], float x, float y, float z){ return glm::vec3(A[0][0] * x + A[0][1] * y + A[0][2] * z, A[1][0] * x + A[1][1] * y + A[1][2] * z, A[2][0] * x + A[2][1] * y + A[2][2] * z); };
                std::vector<glm::vec3> pos(nv), nrm(nv, glm::vec3(0.0f)), tan(T.empty() ? 0 : nv);
                std::vector<glm::vec2> uv(nv, glm::vec2(0.0f));
                for(size_t i = 0; i < nv; i++){
                    const glm::vec4 q = M * glm::vec4(P[3 * i], P[3 * i + 1], P[3 * i + 2], 1.0f);
                    pos[i] = glm::vec3(q.x, q.y, q.z);
                    if(!N.empty())nrm[i] = mul3(NM, N[3 * i], N[3 * i + 1], N[3 * i + 2]);
                    if(!T.empty())tan[i] = mul3(L, T[3 * i], T[3 * i + 1], T[3 * i + 2]);
                    if(!U.empty())uv[i] = glm::vec2(U[2 * i], 1.0f - U[2 * i + 1]);        // aiProcess_FlipUVs
                }
                const bool gen_tan = T.empty() && !U.empty();
                if(gen_tan)tan.assign(nv, glm::vec3(0.0f));
                std::vector<glm::vec3> acc(nv, glm::vec3(0.0f));
                for(size_t f = 0; f + 2 < idx.size(); f += 3){
                    const uint32_t i0 = idx[f], i1 = idx[f + 1], i2 = idx[f + 2];
                    const glm::vec3 e1 = pos[i1] - pos[i0], e2 = pos[i2] - pos[i0], fn = glm::cross(e1, e2);
                    acc[i0] = acc[i0] + fn; acc[i1] = acc[i1] + fn; acc[i2] = acc[i2] + fn;
                    if(gen_tan){
                        const glm::vec2 d1 = uv[i1] - uv[i0], d2 = uv[i2] - uv[i0];
                        const float det = d1.x * d2.y - d2.x * d1.y;
                        if(std::fabs(det) > 1e-20f){
                            const glm::vec3 t = (e1 * d2.y - e2 * d1.y) * (1.0f / det);
                            tan[i0] = tan[i0] + t; tan[i1] = tan[i1] + t; tan[i2] = tan[i2] + t;
                        }
                    }
                }
                for(size_t i = 0; i < nv; i++){
                    glm::vec3& nn = nrm[i];
                    if(glm::dot(nn, nn) == 0.0f)nn = acc[i];                               // aiProcess_GenSmoothNormals
                    const float l2 = glm::dot(nn, nn);
                    nn = l2 > 0.0f ? nn * (1.0f / std::sqrt(l2)) : glm::vec3(0.0f, 1.0f, 0.0f);
                    if(!tan.empty()){
                        glm::vec3& t = tan[i];
                        t = t - nn * glm::dot(nn, t);
                        const float t2 = glm::dot(t, t);
                        t = t2 > 0.0f ? t * (1.0f / std::sqrt(t2)) : glm::vec3(0.0f);
                    }
                }
                Mat& mat = material(pr.integer("material", -1));
                meshes.push_back(std::make_shared<Mesh>(idx, pos, tan, nrm, uv, mat.m, mat.e, mat.med));
            }
        }
        if(const Json* ch = n.get("children"))for(size_t k = 0; k < ch->size(); k++)walk((int)ch->arr[k].num, M, depth + 1);
    };
    const Json* scenes = doc.root.get("scenes");
    const int si = doc.root.integer("scene", 0);
    if(scenes && (size_t)si < scenes->size()){
        if(const Json* roots = scenes->arr[(size_t)si].get("nodes"))for(size_t k = 0; k < roots->size(); k++)walk((int)roots->arr[k].num, glm::mat4(1.0f), 0);
    } else{
        // no scene: every node that is nobody's child is a root
        std::vector<bool> child(nodes->size(), false);
        for(const Json& n : nodes->arr)if(const Json* ch = n.get("children"))for(const Json& c : ch->arr)if((size_t)c.num < child.size())child[(size_t)c.num] = true;
        for(size_t k = 0; k < nodes->size(); k++)if(!child[k])walk((int)k, glm::mat4(1.0f), 0);
    }
    if(meshes.empty())std::fprintf(stderr, "[pathtracing_b200] glTF import: '%s' holds no triangles\n", path.c_str());
    return meshes;
}

inline std::vector<std::shared_ptr<Mesh>> load_model(const std::string& path){
    const size_t dot = path.rfind('.');
    std::string ext = dot == std::string::npos ? "" : path.substr(dot + 1);
    for(char& c : ext)c = (char)std::tolower(c);
    if(ext == "gltf" || ext == "glb")return load_gltf(path);
    if(ext == "obj")return load_obj(path);
    throw std::runtime_error("Model: unsupported model format '" + ext + "' (Wavefront OBJ and glTF 2.0 are read here; the reference's other formats go through assimp)");
}

} // namespace ptb_host
