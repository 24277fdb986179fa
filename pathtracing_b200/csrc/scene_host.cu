// scene_host.cu — scene commit: host BVH build, packing into the device layout,
// one upload.  Replaces Model::BuildBlas (Model.hpp:41-60) + Scene::BuildTlas
// (Scene.hpp:20-24) + the light-sampler wiring of examples/example_1.cpp:99-101.
#include <cuda_runtime.h>
#include <cmath>
#include <cstring>
#include <thread>

#include "scene_host.hpp"

namespace ptb {

static thread_local std::string g_err;
void set_error(const std::string& msg){ g_err = msg; }
const char* get_error(){ return g_err.c_str(); }
bool cuda_ok(int err, const char* what){
    if(err == (int)cudaSuccess)return true;
    set_error(std::string(what) + ": " + cudaGetErrorString((cudaError_t)err));
    return false;
}

template <class T>
static bool upload(void** dst, const std::vector<T>& v, const char* what){
    *dst = nullptr;
    size_t bytes = v.size() * sizeof(T);
    if(bytes == 0)bytes = 16;   // keep pointers valid
    if(!cuda_ok(cudaMalloc(dst, bytes), what))return false;
    if(!v.empty() && !cuda_ok(cudaMemcpy(*dst, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice), what))return false;
    return true;
}

int build_shade_tables(pt_scene* s, const std::vector<uint32_t>& inst_order);   // shade_host.cu

void scene_free_device(pt_scene* s){
    DeviceBuffers& b = s->buf;
    void** ptrs[] = { &b.n2, &b.n4, &b.tris, &b.inst, &b.meshes, &b.blas_info, &b.inst_shade, &b.vtx_pos, &b.vtx_nrm, &b.vtx_uv, &b.vtx_tan,
        &b.indices, &b.tri_light, &b.textures, &b.texels, &b.materials, &b.media, &b.lights, &b.light_cdf, &b.env_cdf, &b.film, &b.work,
        &b.stage_rays, &b.stage_hits, &b.stage_inst, &b.stage_occ, &b.stats, &b.tri_mesh, &b.inst_light_base, &b.sampler_lights, &b.inf_lights };
    for(void** p : ptrs){ if(*p){ cudaFree(*p); *p = nullptr; } }
    b.work_bytes = 0; b.stage_cap = 0;
}

int scene_commit(pt_scene* s, int width, int light_sampler){
    if(s->committed){ set_error("scene already committed"); return PT_ERR_STATE; }
    if(width != 2 && width != 4){ set_error("bvh_width must be 2 or 4"); return PT_ERR_INVALID; }
    if(s->instances.empty()){ set_error("scene has no instances"); return PT_ERR_INVALID; }
    if(!cuda_ok(cudaSetDevice(s->device), "cudaSetDevice"))return PT_ERR_CUDA;
    s->width = width;
    s->light_sampler = light_sampler;

    std::vector<Bvh2Node> n2;
    std::vector<Bvh4Node> n4;
    std::vector<Tri> tris;

    // ---- BLASes ---------------------------------------------------------------------
    for(HostBlas& hb : s->blas){
        hb.n_tris = 0;
        hb.tri_mesh_base.clear();
        for(int mi : hb.meshes){ hb.tri_mesh_base.push_back(hb.n_tris); hb.n_tris += s->meshes[mi].d.n_triangles; }
        if(hb.n_tris == 0){ set_error("empty BLAS"); return PT_ERR_INVALID; }
        if((uint64_t)tris.size() + hb.n_tris >= 0x7FFFFFFFull){ set_error("too many triangles"); return PT_ERR_INVALID; }
        // per-triangle data in original (BLAS-wide) order
        std::vector<Box> boxes(hb.n_tris);
        std::vector<Tri> src(hb.n_tris);
        uint32_t g = 0;
        for(size_t k = 0; k < hb.meshes.size(); k++){
            const int mi = hb.meshes[k];
            const HostMesh& m = s->meshes[mi];
            uint32_t flags = 0;
            if(m.d.material >= 0){
                const pt_material_desc& md = s->materials[m.d.material];
                const bool tester = md.kind == PT_MAT_DIFFUSE || md.kind == PT_MAT_DIELECTRIC;
                if(tester && md.alpha_mode != PT_ALPHA_OPAQUE){
                    flags |= TRI_ALPHA_PATH;     // HasAlpha(), Material.hpp:330-332
                    bool can_be_transparent = md.alpha >= 0;
                    if(!can_be_transparent && md.tex >= 0){
                        const pt_texture_desc& td = s->textures[md.tex].d;
                        // Texture::alpha() is 1 unless an image has 4 channels (Texture.cpp:48-62)
                        can_be_transparent = (td.kind == PT_TEX_IMAGE_U8 || td.kind == PT_TEX_IMAGE_F32 || td.kind == PT_TEX_CHECKER) && (td.kind == PT_TEX_CHECKER || td.channels == 4);
                    }
                    if(can_be_transparent)flags |= TRI_ALPHA_TEX;
                }
            }
            for(uint32_t t = 0; t < m.d.n_triangles; t++, g++){
                const uint32_t i0 = m.idx[3 * (size_t)t], i1 = m.idx[3 * (size_t)t + 1], i2 = m.idx[3 * (size_t)t + 2];
                const float* p0 = &m.pos[3 * (size_t)i0]; const float* p1 = &m.pos[3 * (size_t)i1]; const float* p2 = &m.pos[3 * (size_t)i2];
                Box bx = box_empty();
                box_grow(bx, p0); box_grow(bx, p1); box_grow(bx, p2);   // TriangleShape::BoundingBox, Shape.cpp:270-276
                boxes[g] = bx;
                Tri& T = src[g];
                for(int a = 0; a < 3; a++){ T.v0[a] = p0[a]; T.e1[a] = p1[a] - p0[a]; T.e2[a] = p2[a] - p0[a]; }
                T.prim = g; T.mesh = (uint32_t)mi; T.flags = flags;
            }
        }
        BinaryBvh bb = build_binary_bvh(boxes.data(), hb.n_tris);
        FlatBvh fb = flatten_bvh(bb, width);
        hb.tri_base = (uint32_t)tris.size();
        hb.bounds = fb.bounds;
        const uint32_t node_base = width == 2 ? (uint32_t)n2.size() : (uint32_t)n4.size();
        auto rebase = [&](uint32_t ref) -> uint32_t{
            if(ref == REF_EMPTY)return ref;
            if(ref & REF_LEAF)return REF_LEAF | ((ref & ~REF_LEAF) + hb.tri_base);
            return ref + node_base;
        };
        hb.root = rebase(fb.root);
        if(width == 2){
            for(Bvh2Node nd : fb.n2){ nd.ref0 = rebase(nd.ref0); nd.ref1 = rebase(nd.ref1); n2.push_back(nd); }
        } else{
            for(Bvh4Node nd : fb.n4){ for(int k = 0; k < 4; k++)nd.ref[k] = rebase(nd.ref[k]); n4.push_back(nd); }
        }
        for(uint32_t i = 0; i < hb.n_tris; i++){
            Tri T = src[fb.order[i]];
            if(fb.leaf_last[i])T.flags |= TRI_LAST;
            tris.push_back(T);
        }
    }

    // ---- TLAS -----------------------------------------------------------------------
    const uint32_t ni = (uint32_t)s->instances.size();
    std::vector<Box> iboxes(ni);
    for(uint32_t i = 0; i < ni; i++){
        const HostInstance& hi = s->instances[i];
        if(hi.blas < 0 || hi.blas >= (int)s->blas.size()){ set_error("instance refers to an unknown BLAS"); return PT_ERR_INVALID; }
        const Box& bb = s->blas[hi.blas].bounds;
        iboxes[i] = hi.has_xf ? transform_box(hi.xf, bb) : bb;
    }
    BinaryBvh tb = build_binary_bvh(iboxes.data(), ni);
    FlatBvh tf = flatten_bvh(tb, width);
    s->bounds = tf.bounds;
    const uint32_t tnode_base = width == 2 ? (uint32_t)n2.size() : (uint32_t)n4.size();
    auto trebase = [&](uint32_t ref) -> uint32_t{
        if(ref == REF_EMPTY || (ref & REF_LEAF))return ref;
        return ref + tnode_base;
    };
    if(width == 2){
        for(Bvh2Node nd : tf.n2){ nd.ref0 = trebase(nd.ref0); nd.ref1 = trebase(nd.ref1); n2.push_back(nd); }
    } else{
        for(Bvh4Node nd : tf.n4){ for(int k = 0; k < 4; k++)nd.ref[k] = trebase(nd.ref[k]); n4.push_back(nd); }
    }
    std::vector<Instance> insts(ni);
    for(uint32_t k = 0; k < ni; k++){
        const uint32_t id = tf.order[k];
        const HostInstance& hi = s->instances[id];
        const HostBlas& hb = s->blas[hi.blas];
        Instance& I = insts[k];
        memset(&I, 0, sizeof I);
        float inv[16];
        if(hi.has_xf)mat4_inverse(hi.xf, inv);   // TransformedPrimitive ctor, Primitive.hpp:40
        else{ memset(inv, 0, sizeof inv); inv[0] = inv[5] = inv[10] = inv[15] = 1.0f; }
        for(int c = 0; c < 4; c++)for(int r = 0; r < 3; r++)I.inv[c * 3 + r] = inv[c * 4 + r];
        I.root = hb.root;
        I.flags = (hi.has_xf ? INST_XF : 0u) | (tf.leaf_last[k] ? INST_LAST : 0u);
        I.id = id;
        I.blas = (uint32_t)hi.blas;
        for(int a = 0; a < 3; a++)I.blo[a] = hb.bounds.lo[a];
        I.bhi0 = hb.bounds.hi[0]; I.bhi12[0] = hb.bounds.hi[1]; I.bhi12[1] = hb.bounds.hi[2];
    }

    // ---- upload ---------------------------------------------------------------------
    DeviceBuffers& b = s->buf;
    if(!upload(&b.n2, n2, "upload bvh2 nodes") || !upload(&b.n4, n4, "upload bvh4 clusters") ||
       !upload(&b.tris, tris, "upload triangles") || !upload(&b.inst, insts, "upload instances")){
        scene_free_device(s);
        return PT_ERR_CUDA;
    }
    s->dev.n2 = (const Bvh2Node*)b.n2; s->dev.n4 = (const Bvh4Node*)b.n4;
    s->dev.tris = (const Tri*)b.tris; s->dev.inst = (const Instance*)b.inst;
    s->dev.tlas_root = trebase(tf.root);
    s->dev.width = width;
    for(int a = 0; a < 3; a++){ s->dev.tlo[a] = tf.bounds.lo[a]; s->dev.thi[a] = tf.bounds.hi[a]; }
    s->total_tris = tris.size();
    s->total_nodes = width == 2 ? n2.size() : n4.size();

    int rc = build_shade_tables(s, tf.order);
    if(rc != PT_OK){ scene_free_device(s); return rc; }

    cudaStream_t st, st2;
    if(!cuda_ok(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking), "cudaStreamCreate") ||
       !cuda_ok(cudaStreamCreateWithFlags(&st2, cudaStreamNonBlocking), "cudaStreamCreate")){ scene_free_device(s); return PT_ERR_CUDA; }
    s->stream = st; s->stream2 = st2;
    if(!cuda_ok(cudaMalloc(&b.stats, 64), "cudaMalloc stats")){ scene_free_device(s); return PT_ERR_CUDA; }
    s->committed = true;
    return PT_OK;
}

} // namespace ptb
