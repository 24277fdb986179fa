// scene_build.cpp — host-only half of the scene object: validation of the POD
// description and packing of BLAS/TLAS into the device layout (no CUDA calls,
// so the same code is exercised by the CPU-side emulation tests).
#include <chrono>
#include <cmath>
#include <cstring>
#include <thread>

#include "scene_host.hpp"

namespace ptb {

static thread_local std::string g_err;
void set_error(const std::string& msg){ g_err = msg; }
const char* get_error(){ return g_err.c_str(); }

int scene_add_texture(pt_scene* s, const pt_texture_desc* d){
    if(!d){ set_error("null desc"); return PT_ERR_INVALID; }
    HostTexture t;
    t.d = *d;
    const int ntex = (int)s->textures.size();
    switch(d->kind){
    case PT_TEX_SOLID: case PT_TEX_UV: case PT_TEX_NORMAL: break;
    case PT_TEX_IMAGE_U8: case PT_TEX_IMAGE_F32: {
        if(d->width <= 0 || d->height <= 0 || d->channels < 1 || d->channels > 4 || !d->pixels){ set_error("bad image texture"); return PT_ERR_INVALID; }
        const size_t elems = (size_t)d->width * d->height * d->channels;
        const size_t bytes = elems * (d->kind == PT_TEX_IMAGE_U8 ? 1 : 4);
        t.pixels.assign((const uint8_t*)d->pixels, (const uint8_t*)d->pixels + bytes);
        t.d.pixels = nullptr;
        break;
    }
    case PT_TEX_CHECKER:
        if(d->tex_a < 0 || d->tex_a >= ntex || d->tex_b < 0 || d->tex_b >= ntex){ set_error("checker child texture id out of range"); return PT_ERR_INVALID; }
        break;
    default: set_error("unknown texture kind"); return PT_ERR_INVALID;
    }
    s->textures.push_back(std::move(t));
    return ntex;
}

int scene_add_material(pt_scene* s, const pt_material_desc* d){
    if(!d){ set_error("null desc"); return PT_ERR_INVALID; }
    const int ntex = (int)s->textures.size();
    auto ok = [&](int id, bool required){ return (id >= 0 && id < ntex) || (!required && id < 0); };
    switch(d->kind){
    case PT_MAT_DIFFUSE: case PT_MAT_DIELECTRIC:
        if(!ok(d->tex, true) || !ok(d->norm, false) || !ok(d->rough, false) || !ok(d->metal, false) || !ok(d->alpha, false)){ set_error("material texture id out of range"); return PT_ERR_INVALID; }
        break;
    case PT_MAT_THIN_DIELECTRIC:
        if(!ok(d->tex, false)){ set_error("material texture id out of range"); return PT_ERR_INVALID; }
        break;
    case PT_MAT_CONDUCTOR: break;
    default: set_error("unknown material kind"); return PT_ERR_INVALID;
    }
    s->materials.push_back(*d);
    return (int)s->materials.size() - 1;
}

int scene_add_medium(pt_scene* s, const pt_medium_desc* d){
    if(!d){ set_error("null desc"); return PT_ERR_INVALID; }
    s->media.push_back(*d);
    return (int)s->media.size() - 1;
}

int scene_add_mesh(pt_scene* s, const pt_mesh_desc* d){
    if(!d || !d->positions || !d->indices || !d->normals || !d->texcoords){ set_error("mesh needs positions, normals, texcoords and indices"); return PT_ERR_INVALID; }
    if(d->material >= (int)s->materials.size() || d->emissive_tex >= (int)s->textures.size() || d->medium >= (int)s->media.size()){ set_error("mesh material/texture/medium id out of range"); return PT_ERR_INVALID; }
    for(size_t i = 0; i < 3 * (size_t)d->n_triangles; i++)
        if(d->indices[i] >= d->n_vertices){ set_error("mesh index out of range"); return PT_ERR_INVALID; }
    HostMesh m;
    m.d = *d;
    m.pos.assign(d->positions, d->positions + 3 * (size_t)d->n_vertices);
    m.nrm.assign(d->normals, d->normals + 3 * (size_t)d->n_vertices);
    m.uv.assign(d->texcoords, d->texcoords + 2 * (size_t)d->n_vertices);
    if(d->tangents)m.tan.assign(d->tangents, d->tangents + 3 * (size_t)d->n_vertices);
    m.idx.assign(d->indices, d->indices + 3 * (size_t)d->n_triangles);
    m.d.positions = m.d.normals = m.d.texcoords = m.d.tangents = nullptr; m.d.indices = nullptr;
    s->meshes.push_back(std::move(m));
    return (int)s->meshes.size() - 1;
}

int scene_add_blas(pt_scene* s, const int32_t* mesh_ids, int n){
    if(!mesh_ids || n <= 0){ set_error("a BLAS needs at least one mesh"); return PT_ERR_INVALID; }
    HostBlas b;
    for(int i = 0; i < n; i++){
        if(mesh_ids[i] < 0 || mesh_ids[i] >= (int)s->meshes.size()){ set_error("mesh id out of range"); return PT_ERR_INVALID; }
        // a Mesh belongs to one Model / BLAS (Model.hpp:41-60); share geometry through instances instead
        for(const HostBlas& o : s->blas)for(int m : o.meshes)if(m == mesh_ids[i]){ set_error("mesh already belongs to another BLAS"); return PT_ERR_INVALID; }
        for(int m : b.meshes)if(m == mesh_ids[i]){ set_error("mesh listed twice in one BLAS"); return PT_ERR_INVALID; }
        b.meshes.push_back(mesh_ids[i]);
    }
    s->blas.push_back(std::move(b));
    return (int)s->blas.size() - 1;
}

int scene_add_instance(pt_scene* s, int blas_id, const float* xf){
    if(blas_id < 0 || blas_id >= (int)s->blas.size()){ set_error("blas id out of range"); return PT_ERR_INVALID; }
    HostInstance i;
    i.blas = blas_id;
    i.has_xf = xf != nullptr;
    memset(i.xf, 0, sizeof i.xf);
    if(xf)memcpy(i.xf, xf, sizeof i.xf);
    else i.xf[0] = i.xf[5] = i.xf[10] = i.xf[15] = 1.0f;
    memset(&i.shape, 0, sizeof i.shape);
    s->instances.push_back(i);
    return (int)s->instances.size() - 1;
}

int scene_add_shape(pt_scene* s, const pt_shape_desc* d, const float* xf){
    if(!d || (d->kind != PT_SHAPE_SPHERE && d->kind != PT_SHAPE_QUAD)){ set_error("bad shape"); return PT_ERR_INVALID; }
    if(d->material >= (int)s->materials.size() || d->emissive_tex >= (int)s->textures.size() || d->medium >= (int)s->media.size()){ set_error("shape material/texture/medium id out of range"); return PT_ERR_INVALID; }
    if(d->kind == PT_SHAPE_SPHERE && !(d->radius > 0.0f)){ set_error("sphere radius must be positive"); return PT_ERR_INVALID; }
    HostInstance i;
    i.blas = -1; i.has_xf = xf != nullptr;       // xf: TransformedPrimitive(GeometricPrimitive(shape), M)
    memset(i.xf, 0, sizeof i.xf);
    if(xf)memcpy(i.xf, xf, sizeof i.xf);
    else i.xf[0] = i.xf[5] = i.xf[10] = i.xf[15] = 1.0f;
    i.shape = *d;
    s->instances.push_back(i);
    return (int)s->instances.size() - 1;
}

// AnimatedPrimitive(primitive, direction, timeBounds) around an already added top-level primitive
// (Primitive.hpp:52-66; main.cpp:189 wraps a sphere GeometricPrimitive).
int scene_animate(pt_scene* s, int prim_id, const float* dir, const float* tb){
    if(prim_id < 0 || prim_id >= (int)s->instances.size()){ set_error("primitive id out of range"); return PT_ERR_INVALID; }
    if(!dir || !tb){ set_error("null direction / timeBounds"); return PT_ERR_INVALID; }
    HostInstance& i = s->instances[prim_id];
    if(i.animated){ set_error("nested AnimatedPrimitive wrappers are not supported"); return PT_ERR_INVALID; }
    i.animated = true;
    for(int k = 0; k < 3; k++)i.adir[k] = dir[k];
    i.atb[0] = tb[0]; i.atb[1] = tb[1];
    return PT_OK;
}

int scene_add_light(pt_scene* s, const pt_light_desc* d){
    if(!d || d->kind < 0 || d->kind > PT_LIGHT_POINT){ set_error("bad light"); return PT_ERR_INVALID; }
    if(d->kind == PT_LIGHT_TEXTURE_INFINITE && (d->tex < 0 || d->tex >= (int)s->textures.size())){ set_error("light texture id out of range"); return PT_ERR_INVALID; }
    if(d->kind == PT_LIGHT_GRADIENT_INFINITE && d->tex >= (int)s->textures.size()){ set_error("light texture id out of range"); return PT_ERR_INVALID; }
    s->lights.push_back(*d);
    return (int)s->lights.size() - 1;
}

// BLAS builds (Model::BuildBlas, Model.hpp:41-60), TLAS build (Scene::BuildTlas,
// Scene.hpp:20-24) and packing into one node array / one triangle array.
// Texture::alpha() is 1 unless an image has 4 channels (Texture.cpp:46-62,72-88); a checker forwards to
// its children (Texture.cpp:90-94)
static bool texture_alpha_below_one(const pt_scene* s, int tex, int depth){
    if(tex < 0 || depth > 8)return false;
    const pt_texture_desc& td = s->textures[tex].d;
    if(td.kind == PT_TEX_IMAGE_U8 || td.kind == PT_TEX_IMAGE_F32)return td.channels == 4;
    if(td.kind == PT_TEX_CHECKER)return texture_alpha_below_one(s, td.tex_a, depth + 1) || texture_alpha_below_one(s, td.tex_b, depth + 1);
    return false;
}

// TRI_ALPHA_* flags of a material: HasAlpha() (Material.hpp:330-332) and whether Alpha() can ever reject
static uint32_t material_alpha_flags(const pt_scene* s, int mat){
    if(mat < 0)return 0u;
    const pt_material_desc& md = s->materials[mat];
    const bool tester = md.kind == PT_MAT_DIFFUSE || md.kind == PT_MAT_DIELECTRIC;
    if(!tester || md.alpha_mode == PT_ALPHA_OPAQUE)return 0u;
    bool can_be_transparent = md.alpha >= 0;
    if(!can_be_transparent && md.tex >= 0)can_be_transparent = texture_alpha_below_one(s, md.tex, 0);
    return TRI_ALPHA_PATH | (can_be_transparent ? TRI_ALPHA_TEX : 0u);
}

uint32_t mesh_tri_flags(const pt_scene* s, int mesh){ return material_alpha_flags(s, s->meshes[mesh].d.material); }

// SphereShape ctor (Shape.hpp:23-27) / QuadShape ctor (Shape.hpp:120-130): AABB::Expand over the same points
static Box shape_box(const pt_shape_desc& d){
    Box b = box_empty();
    if(d.kind == PT_SHAPE_SPHERE){
        const float lo[3] = { d.p[0] - d.radius, d.p[1] - d.radius, d.p[2] - d.radius }, hi[3] = { d.p[0] + d.radius, d.p[1] + d.radius, d.p[2] + d.radius };
        box_grow(b, lo); box_grow(b, hi);
    } else{
        float a[3], c[3], e[3];
        for(int k = 0; k < 3; k++){ a[k] = (d.p[k] + d.u[k]) + d.v[k]; c[k] = d.p[k] + d.u[k]; e[k] = d.p[k] + d.v[k]; }
        box_grow(b, d.p); box_grow(b, a); box_grow(b, c); box_grow(b, e);
    }
    return b;
}

// The shape's constants in the 96 B instance record (pt_types.h): inv[0..11] and blo/bhi0 are reused.
//  sphere: inv[0..2] center, inv[3] radius
//  quad  : inv[0..2] Q, inv[3..5] u, inv[6..8] v, inv[9..11] normal, blo[0..2] w = n / dot(n, n), bhi0 = D = dot(normal, Q)
static void pack_shape(const pt_shape_desc& d, Instance& I){
    if(d.kind == PT_SHAPE_SPHERE){
        I.flags = INST_SPHERE;
        for(int k = 0; k < 3; k++)I.inv[k] = d.p[k];
        I.inv[3] = d.radius;
        return;
    }
    I.flags = INST_QUAD;
    const float* u = d.u; const float* v = d.v;
    const float n[3] = { u[1] * v[2] - v[1] * u[2], u[2] * v[0] - v[2] * u[0], u[0] * v[1] - v[0] * u[1] };   // glm::cross
    const float nn = n[0] * n[0] + n[1] * n[1] + n[2] * n[2];
    const float inv_len = 1.0f / std::sqrt(nn);                                                                 // glm::normalize = v * inversesqrt(dot)
    for(int k = 0; k < 3; k++){
        I.inv[k] = d.p[k]; I.inv[3 + k] = u[k]; I.inv[6 + k] = v[k];
        I.inv[9 + k] = n[k] * inv_len;
        I.blo[k] = n[k] / nn;
    }
    I.bhi0 = I.inv[9] * d.p[0] + I.inv[10] * d.p[1] + I.inv[11] * d.p[2];
}

void (*g_dev_free)(void*) = nullptr;
void DevBlasPiece::release(){
    if(g_dev_free){ if(d_nodes)g_dev_free(d_nodes); if(d_tris)g_dev_free(d_tris); }
    d_nodes = d_tris = nullptr;
}

// the per-triangle loops of a 10 M-triangle commit (boxes, edge records, the gather into leaf order, ref rebasing) took
// 1.3 s on one thread: split them over the host threads
template <class F>
static void parallel_for(size_t n, F fn){
    unsigned nt = std::thread::hardware_concurrency();
    if(nt == 0)nt = 1;
    if(nt > 16)nt = 16;
    if(n < (size_t)1 << 16 || nt == 1){ fn((size_t)0, n); return; }
    std::vector<std::thread> th;
    for(unsigned t = 0; t < nt; t++)th.emplace_back([=]{ fn(n * t / nt, n * (t + 1) / nt); });
    for(std::thread& t : th)t.join();
}

int pack_traversal(pt_scene* s, int width, PackedTraversal& P, BvhBuilder builder, void* user, BlasPacker packer){
    if(width != 2 && width != 4 && width != 8){ set_error("bvh_width must be 2, 4 or 8"); return PT_ERR_INVALID; }
    if(s->instances.empty()){ set_error("scene has no instances"); return PT_ERR_INVALID; }
    std::vector<Bvh2Node>& n2 = P.n2;
    std::vector<Bvh4Node>& n4 = P.n4;
    std::vector<Tri>& tris = P.tris;
    // scene-wide counts so far: host vectors hold only what was packed on the host (P.node_segs / P.tri_segs say where it goes)
    size_t g_nodes = 0, g_tris = 0;
    auto host_nodes = [&]() -> size_t{ return width == 2 ? n2.size() : width == 4 ? n4.size() : P.n8.size(); };

    // ---- BLASes ---------------------------------------------------------------------
    for(HostBlas& hb : s->blas){
        hb.n_tris = 0;
        hb.tri_mesh_base.clear();
        for(int mi : hb.meshes){ hb.tri_mesh_base.push_back(hb.n_tris); hb.n_tris += s->meshes[mi].d.n_triangles; }
        if(hb.n_tris == 0){ set_error("empty BLAS"); return PT_ERR_INVALID; }
        if((uint64_t)g_tris + hb.n_tris >= 0x7FFFFFFFull){ set_error("too many triangles"); return PT_ERR_INVALID; }
        if(packer){
            // the whole BLAS on the device: boxes, binary build, collapse, triangle records (scene_pack_gpu.cu)
            for(int mi : hb.meshes){
                const HostMesh& m = s->meshes[mi];
                if(m.d.material < 0)s->has_boundary = true;
                if(material_alpha_flags(s, m.d.material) & TRI_ALPHA_TEX)s->has_alpha_tex = true;
            }
            DevBlasPiece piece;
            Box bounds = box_empty();
            const int prc = packer(s, hb, width, piece, bounds, user);
            if(prc < 0)return prc;
            if(prc == PT_OK){
                piece.node_base = (uint32_t)g_nodes; piece.tri_base = (uint32_t)g_tris;
                hb.tri_base = piece.tri_base; hb.bounds = bounds; hb.root = piece.node_base;     // the root cluster is node 0 of the piece
                g_nodes += piece.n_nodes; g_tris += piece.n_tris;
                P.dev_pieces.push_back(std::move(piece));
                continue;
            }
        }
        // per-triangle data in original (BLAS-wide) order
        std::vector<Box> boxes(hb.n_tris);
        std::vector<Tri> src(hb.n_tris);
        uint32_t g = 0;
        for(size_t k = 0; k < hb.meshes.size(); k++){
            const int mi = hb.meshes[k];
            const HostMesh& m = s->meshes[mi];
            const uint32_t flags = material_alpha_flags(s, m.d.material);
            if(m.d.material < 0)s->has_boundary = true;
            if(flags & TRI_ALPHA_TEX)s->has_alpha_tex = true;
            const uint32_t g0 = g;
            parallel_for(m.d.n_triangles, [&, g0, mi, flags](size_t lo, size_t hi){
                for(size_t t = lo; t < hi; t++){
                    const uint32_t gg = g0 + (uint32_t)t;
                    const uint32_t i0 = m.idx[3 * t], i1 = m.idx[3 * t + 1], i2 = m.idx[3 * t + 2];
                    const float* p0 = &m.pos[3 * (size_t)i0]; const float* p1 = &m.pos[3 * (size_t)i1]; const float* p2 = &m.pos[3 * (size_t)i2];
                    Box bx = box_empty();
                    box_grow(bx, p0); box_grow(bx, p1); box_grow(bx, p2);   // TriangleShape::BoundingBox, Shape.cpp:270-276
                    boxes[gg] = bx;
                    Tri& T = src[gg];
                    for(int a = 0; a < 3; a++){ T.v0[a] = p0[a]; T.e1[a] = p1[a] - p0[a]; T.e2[a] = p2[a] - p0[a]; }
                    T.prim = gg; T.mesh = (uint32_t)mi; T.flags = flags;
                }
            });
            g += m.d.n_triangles;
        }
        const auto tb0 = std::chrono::steady_clock::now();
        BinaryBvh bb;
        if(builder){ const int brc = builder(boxes.data(), hb.n_tris, bb, user); if(brc != PT_OK)return brc; }
        else bb = build_binary_bvh(boxes.data(), hb.n_tris);
        const auto tb1 = std::chrono::steady_clock::now();
        FlatBvh fb = flatten_bvh(bb, width);
        const auto tb2 = std::chrono::steady_clock::now();
        s->build_stats.bvh2_s += std::chrono::duration<double>(tb1 - tb0).count();
        s->build_stats.flatten_s += std::chrono::duration<double>(tb2 - tb1).count();
        hb.tri_base = (uint32_t)g_tris;
        hb.bounds = fb.bounds;
        const uint32_t node_base = (uint32_t)g_nodes;
        const size_t h_nodes = host_nodes(), h_tris = tris.size();
        const size_t fb_nodes = width == 2 ? fb.n2.size() : width == 4 ? fb.n4.size() : fb.n8.size();
        P.node_segs.push_back({ h_nodes, fb_nodes, g_nodes });
        P.tri_segs.push_back({ h_tris, hb.n_tris, g_tris });
        g_nodes += fb_nodes; g_tris += hb.n_tris;
        auto rebase = [&](uint32_t ref) -> uint32_t{
            if(ref == REF_EMPTY)return ref;
            if(ref & REF_LEAF)return REF_LEAF | ((ref & ~REF_LEAF) + hb.tri_base);
            return ref + node_base;
        };
        hb.root = rebase(fb.root);
        if(width == 2){
            n2.resize(h_nodes + fb.n2.size());
            parallel_for(fb.n2.size(), [&](size_t lo, size_t hi){ for(size_t q = lo; q < hi; q++){ Bvh2Node nd = fb.n2[q]; nd.ref0 = rebase(nd.ref0); nd.ref1 = rebase(nd.ref1); n2[h_nodes + q] = nd; } });
        } else if(width == 4){
            n4.resize(h_nodes + fb.n4.size());
            parallel_for(fb.n4.size(), [&](size_t lo, size_t hi){ for(size_t q = lo; q < hi; q++){ Bvh4Node nd = fb.n4[q]; for(int k = 0; k < 4; k++)nd.ref[k] = rebase(nd.ref[k]); n4[h_nodes + q] = nd; } });
        } else{
            P.n8.resize(h_nodes + fb.n8.size());
            parallel_for(fb.n8.size(), [&](size_t lo, size_t hi){ for(size_t q = lo; q < hi; q++){ Bvh8Node nd = fb.n8[q]; for(int k = 0; k < 8; k++)nd.ref[k] = rebase(nd.ref[k]); P.n8[h_nodes + q] = nd; } });
        }
        tris.resize(h_tris + hb.n_tris);
        parallel_for(hb.n_tris, [&](size_t lo, size_t hi){
            for(size_t i = lo; i < hi; i++){
                Tri T = src[fb.order[i]];
                if(fb.leaf_last[i])T.flags |= TRI_LAST;
                tris[h_tris + i] = T;
            }
        });
    }

    // ---- TLAS -----------------------------------------------------------------------
    const uint32_t ni = (uint32_t)s->instances.size();
    std::vector<Box> iboxes(ni);
    for(uint32_t i = 0; i < ni; i++){
        const HostInstance& hi = s->instances[i];
        if(hi.blas < 0){
            if(material_alpha_flags(s, hi.shape.material) & TRI_ALPHA_TEX)s->has_alpha_tex = true;
            if(hi.shape.material < 0)s->has_boundary = true;
            iboxes[i] = shape_box(hi.shape);
            if(hi.has_xf)iboxes[i] = transform_box(hi.xf, iboxes[i]);    // TransformedPrimitive::BoundingBox, Primitive.cpp:32-40
        } else{
            if(hi.blas >= (int)s->blas.size()){ set_error("instance refers to an unknown BLAS"); return PT_ERR_INVALID; }
            const Box& bb = s->blas[hi.blas].bounds;
            iboxes[i] = hi.has_xf ? transform_box(hi.xf, bb) : bb;
        }
        if(hi.animated){
            // AnimatedPrimitive::BoundingBox (Primitive.cpp:77-81): the child's box and the same box under translate(dir)
            float tr[16] = { 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, hi.adir[0], hi.adir[1], hi.adir[2], 1 };
            const Box moved = transform_box(tr, iboxes[i]);
            box_grow(iboxes[i], moved);
        }
    }
    const auto tt0 = std::chrono::steady_clock::now();
    BinaryBvh tb;
    if(builder){ const int brc = builder(iboxes.data(), ni, tb, user); if(brc != PT_OK)return brc; }
    else tb = build_binary_bvh(iboxes.data(), ni);
    const auto tt1 = std::chrono::steady_clock::now();
    FlatBvh tf = flatten_bvh(tb, width);
    s->build_stats.bvh2_s += std::chrono::duration<double>(tt1 - tt0).count();
    s->build_stats.flatten_s += std::chrono::duration<double>(std::chrono::steady_clock::now() - tt1).count();
    s->bounds = tf.bounds;
    const uint32_t tnode_base = (uint32_t)g_nodes;
    const size_t th_nodes = host_nodes();       // where the TLAS starts in the host vector
    {
        const size_t tf_nodes = width == 2 ? tf.n2.size() : width == 4 ? tf.n4.size() : tf.n8.size();
        P.node_segs.push_back({ th_nodes, tf_nodes, g_nodes });
        g_nodes += tf_nodes;
    }
    P.total_nodes = g_nodes; P.total_tris = g_tris;
    auto trebase = [&](uint32_t ref) -> uint32_t{
        if(ref == REF_EMPTY || (ref & REF_LEAF))return ref;
        return ref + tnode_base;
    };
    if(width == 2){
        for(Bvh2Node nd : tf.n2){ nd.ref0 = trebase(nd.ref0); nd.ref1 = trebase(nd.ref1); n2.push_back(nd); }
    } else if(width == 4){
        for(Bvh4Node nd : tf.n4){ for(int k = 0; k < 4; k++)nd.ref[k] = trebase(nd.ref[k]); n4.push_back(nd); }
    } else{
        for(Bvh8Node nd : tf.n8){ for(int k = 0; k < 8; k++)nd.ref[k] = trebase(nd.ref[k]); P.n8.push_back(nd); }
    }
    // One chain of 1-3 records per top-level primitive (pt_types.h): [ANIM]? [WRAP|XF]? (BLAS instance | shape)
    std::vector<Instance>& insts = P.insts;
    insts.clear();
    std::vector<uint32_t> rec_of(ni);
    for(uint32_t k = 0; k < ni; k++){
        const uint32_t id = tf.order[k];
        const HostInstance& hi = s->instances[id];
        const uint32_t head = (uint32_t)insts.size();
        rec_of[k] = head;
        float inv[16];
        if(hi.has_xf)mat4_inverse(hi.xf, inv);   // TransformedPrimitive ctor, Primitive.hpp:40
        else{ memset(inv, 0, sizeof inv); inv[0] = inv[5] = inv[10] = inv[15] = 1.0f; }
        if(hi.animated){
            Instance A;
            memset(&A, 0, sizeof A);
            for(int a = 0; a < 3; a++)A.inv[a] = hi.adir[a];
            A.inv[3] = hi.atb[0]; A.inv[4] = hi.atb[1];
            A.flags = INST_ANIM; A.id = id;
            insts.push_back(A);
        }
        if(hi.blas < 0 && hi.has_xf){
            Instance W;
            memset(&W, 0, sizeof W);
            for(int c = 0; c < 4; c++)for(int r = 0; r < 3; r++)W.inv[c * 3 + r] = inv[c * 4 + r];
            W.flags = INST_WRAP | INST_XF; W.id = id;
            insts.push_back(W);
        }
        Instance I;
        memset(&I, 0, sizeof I);
        if(hi.blas < 0){
            pack_shape(hi.shape, I);
            I.id = id;
            if(material_alpha_flags(s, hi.shape.material) & TRI_ALPHA_TEX){ I.flags |= INST_ALPHA; I.root = (uint32_t)hi.shape.material; }
        } else{
            const HostBlas& hb = s->blas[hi.blas];
            for(int c = 0; c < 4; c++)for(int r = 0; r < 3; r++)I.inv[c * 3 + r] = inv[c * 4 + r];
            I.root = hb.root;
            I.flags = hi.has_xf ? INST_XF : 0u;
            I.id = id;
            I.blas = (uint32_t)hi.blas;
            for(int a = 0; a < 3; a++)I.blo[a] = hb.bounds.lo[a];
            I.bhi0 = hb.bounds.hi[0]; I.bhi12[0] = hb.bounds.hi[1]; I.bhi12[1] = hb.bounds.hi[2];
            if(width != 2 && !hi.animated){
                // BVH4 / BVH8: the record carries the instance's WORLD box instead (the BLAS root box is only read by the BVH2 kernels).  A TLAS
                // leaf holds up to two instances and its cluster box is their union: testing the ray against each one's own box before paying for
                // the transform (about 250 instructions with the reference's arithmetic) skips the one it misses.  Enlarged by 1e-4 of its extent
                // (+ 1e-6 of its magnitude) per side, so that a hit found in object space is never lost to the rounding of this test: an instance
                // is skipped only when the ray cannot hit anything in it, so the hits stay the reference's bit for bit (ties included).
                const Box& wb = iboxes[id];
                float lo[3], hh[3];
                for(int a = 0; a < 3; a++){
                    const float m = 1e-4f * (wb.hi[a] - wb.lo[a]) + 1e-6f * std::fmax(std::fabs(wb.lo[a]), std::fabs(wb.hi[a])) + 1e-30f;
                    lo[a] = wb.lo[a] - m; hh[a] = wb.hi[a] + m;
                }
                for(int a = 0; a < 3; a++)I.blo[a] = lo[a];
                I.bhi0 = hh[0]; I.bhi12[0] = hh[1]; I.bhi12[1] = hh[2];
                I.flags |= INST_WBOX;
            }
        }
        insts.push_back(I);
        const uint32_t extra = (uint32_t)insts.size() - head - 1;
        if(extra || hi.blas < 0)s->has_wrappers = true;
        insts[head].flags |= (tf.leaf_last[k] ? INST_LAST : 0u) | (extra << INST_EXTRA_SHIFT);
    }
    // TLAS leaf refs: leaf position -> head record
    auto leaf_fix = [&](uint32_t ref) -> uint32_t{
        if(ref == REF_EMPTY || !(ref & REF_LEAF))return ref;
        return REF_LEAF | rec_of[ref & ~REF_LEAF];
    };
    if(width == 2){ for(size_t q = th_nodes; q < n2.size(); q++){ n2[q].ref0 = leaf_fix(n2[q].ref0); n2[q].ref1 = leaf_fix(n2[q].ref1); } }
    else if(width == 4){ for(size_t q = th_nodes; q < n4.size(); q++)for(int k = 0; k < 4; k++)n4[q].ref[k] = leaf_fix(n4[q].ref[k]); }
    else{ for(size_t q = th_nodes; q < P.n8.size(); q++)for(int k = 0; k < 8; k++)P.n8[q].ref[k] = leaf_fix(P.n8[q].ref[k]); }

    if(ni == 1 && !s->instances[0].has_xf && !s->instances[0].animated && s->instances[0].blas >= 0){ P.single_root = s->blas[s->instances[0].blas].root; P.single_id = 0; }
    P.tlas_root = leaf_fix(trebase(tf.root));
    P.bounds = tf.bounds;
    P.inst_order = tf.order;
    return PT_OK;
}

} // namespace ptb
