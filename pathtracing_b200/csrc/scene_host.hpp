// scene_host.hpp — host-side scene object behind the C-ABI (pt_scene).
// Collects the POD description (pt_b200.h), builds BLAS/TLAS on the host
// (bvh_build.hpp), packs everything into the device layout (pt_types.h,
// shade_types.h) and uploads it once at commit.  The scene is immutable after
// commit, like the reference's after Scene::BuildTlas (Scene.hpp:20-24).
#pragma once
#include <cstdint>
#include <string>
#include <utility>
#include <vector>

#include "../../include/pt_b200.h"
#include "bvh_build.hpp"
#include "pt_types.h"
#include "shade_types.h"

namespace ptb {

struct HostTexture {
    pt_texture_desc d;
    std::vector<uint8_t> pixels;   // owned copy of image texels (u8 or f32 bytes)
};

struct HostMesh {
    pt_mesh_desc d;
    std::vector<float> pos, nrm, uv, tan;
    std::vector<uint32_t> idx;
};

struct HostBlas {
    std::vector<int> meshes;
    // filled at commit
    uint32_t n_tris = 0;
    uint32_t tri_base = 0;        // offset of this BLAS's triangles in the global leaf-order array
    uint32_t root = REF_EMPTY;    // rebased root ref
    Box bounds = box_empty();
    std::vector<uint32_t> tri_mesh_base; // per mesh: first BLAS-wide triangle id
};

struct HostInstance {        // one top-level primitive (Scene::Add order)
    int blas;                // >= 0: a BLAS, optionally transformed; -1: an analytic shape
    bool has_xf;
    float xf[16];
    pt_shape_desc shape;     // blas < 0
    bool animated = false;   // wrapped in an AnimatedPrimitive (pt_scene_animate)
    float adir[3] = { 0, 0, 0 };
    float atb[2] = { 0, 1 };
};

struct DeviceBuffers {
    void* n2 = nullptr; void* n4 = nullptr; void* n8 = nullptr; void* tris = nullptr; void* inst = nullptr;
    void* n4q = nullptr; void* tris64 = nullptr;      // PT_LAYOUT_FAST
    // shading tables
    void* meshes = nullptr; void* blas_info = nullptr; void* inst_shade = nullptr;
    void* vtx_pos = nullptr; void* vtx_nrm = nullptr; void* vtx_uv = nullptr; void* vtx_tan = nullptr; void* indices = nullptr;
    void* tri_light = nullptr; void* tri_mesh = nullptr; void* inst_light_base = nullptr; void* sampler_lights = nullptr; void* inf_lights = nullptr;
    void* textures = nullptr; void* texels = nullptr; void* materials = nullptr; void* media = nullptr;
    void* lights = nullptr; void* light_cdf = nullptr; void* env_cdf = nullptr;
    void* film = nullptr;
    void* film_sq = nullptr;      // per-pixel sum of squared sample radiance, double[3*W*H]
    void* work = nullptr;         // wavefront state arena
    size_t work_bytes = 0;
    void* h_counters = nullptr;   // pinned host memory: the watch slots of the wavefront loop (copies of its device control block)
    void* wave_events[8] = {};    // one event per watch slot
    void* tile_list = nullptr; size_t tile_list_bytes = 0;   // tile-sharded frames: this rank's pixels
    void* adapt = nullptr;        // adaptive loop: per-pixel estimator sums + pixel lists
    size_t adapt_bytes = 0;
    // staging for the host-buffer ray-cast entry points
    void* stage_rays = nullptr; void* stage_hits = nullptr; void* stage_inst = nullptr; void* stage_occ = nullptr;
    void* stage_rays24 = nullptr; void* stage_hits8 = nullptr; size_t stage24_cap = 0;   // pt_trace_closest_compact
    size_t stage_cap = 0;
    void* stream_aux = nullptr; size_t stream_aux_bytes = 0;   // device: landed watermark, abort flag, per-granule counters
    void* h_stream = nullptr; size_t h_stream_bytes = 0;       // pinned: watermark values + polled copy of stream_aux
    void* stats = nullptr;
    void* cursors = nullptr;      // ring of ray-dispenser cursors for the persistent kernels
};

constexpr unsigned kCursorRing = 256;

// bvh_build_gpu.cu: accumulated over the BLAS / TLAS builds of one commit
struct GpuBuildTimes {
    float upload_ms = 0, build_ms = 0, download_ms = 0;
    int levels = 0;              // level-synchronous rounds of the large phase
    uint32_t subtrees = 0;       // shared-memory subtree CTAs
    uint64_t items = 0;
};

} // namespace ptb

struct pt_scene {
    int device = 0;
    bool committed = false;
    int width = 0;
    int light_sampler = 1;
    std::vector<ptb::HostTexture> textures;
    std::vector<pt_material_desc> materials;
    std::vector<pt_medium_desc> media;
    std::vector<ptb::HostMesh> meshes;
    std::vector<ptb::HostBlas> blas;
    std::vector<ptb::HostInstance> instances;
    std::vector<pt_light_desc> lights;
    int scene_medium = -1;

    ptb::Box bounds = ptb::box_empty();
    ptb::DevScene dev {};
    ptb::ShadeScene shade {};
    ptb::DeviceBuffers buf;
    int film_w = 0, film_h = 0;
    uint64_t total_tris = 0, total_nodes = 0, n_inst_records = 0;
    uint64_t n_shade_vertices = 0, n_shade_indices = 0, n_shade_tris = 0;     // sizes of the shading geometry tables
    void* stream = nullptr;       // internal stream for the host-buffer entry points
    void* stream2 = nullptr;      // H2D stage of the host-buffer pipeline
    void* stream3 = nullptr;      // D2H stage
    void* stream4 = nullptr;      // second trace stream of the host-buffer pipeline (odd chunks): chunk tails overlap
    std::vector<void*> events;    // cudaEvent_t pool of the pipeline (no timing)
    unsigned cursor_next = 0;
    bool streamed_probed = false; // probe_async_launch ran for this scene
    bool streamed_off = false;    // the streamed host-buffer ray cast gave up once (capi.cu): do not try again
    bool has_wrappers = false;    // some top-level primitive is an analytic shape or a wrapper chain (INST_SPHERE / QUAD / ANIM / WRAP): kWrap kernel variants
    bool has_alpha_tex = false;   // some triangle carries TRI_ALPHA_TEX: traversal must evaluate Material::Alpha
    bool has_boundary = false;    // some primitive has no material (a medium boundary): VolPath shadow rays walk through it
    uint64_t alpha_salt = 0;      // advances per launch (AlphaTester Blend draws)
    unsigned long long overflow_seen = 0;  // device overflow counter (buf.stats + 64) as the last checked call left it
    pt_allreduce_fn allreduce = nullptr;   // sums a device buffer over the ranks that shard this scene's sample passes
    void* allreduce_user = nullptr;
    int builder = 1;              // pt_builder: where the binary SAH build runs (default: on the GPU, same tree; PTB_BUILDER=host in the environment or pt_scene_set_builder)
    bool counting = false;        // pt_scene_set_counting: pt_render launches the counting build of the traversal kernels
    int layout = 0;               // pt_layout: exact 128 B clusters, or the quantised 64 B ones (width 4)
    // every device table of the committed scene as (byte offset of its pointer inside `buf`, bytes): what
    // pt_scene_replicate copies to another GPU (multi.cu) without running the build again
    std::vector<std::pair<size_t, size_t>> uploads;
    // multi-GPU (multi.cu): this scene's rank in the communicator that shards its sample passes
    void* comm = nullptr;         // ncclComm_t, or null
    int comm_rank = 0, comm_world = 1;
    bool comm_owned = false;      // created by pt_dist_init (destroyed with the scene / pt_dist_finalize)
    int tile_rank = 0, tile_world = 1;   // set by multi.cu around a frame that is sharded by tiles instead of passes (render.cu)
    pt_build_stats build_stats {};
};

namespace ptb {

// A BLAS packed on the device (scene_pack_gpu.cu): its clusters and triangle records sit in device buffers of their own, refs local
// to the BLAS (inner refs index d_nodes, leaf refs index d_tris); pack_traversal gives it its place in the scene-wide arrays.
struct DevBlasPiece {
    void* d_nodes = nullptr;      // Bvh2Node / Bvh4Node / Bvh8Node[n_nodes] (the commit's width), DFS preorder (the host flatten's numbering)
    void* d_tris = nullptr;       // Tri[n_tris], leaf order
    uint32_t n_nodes = 0, n_tris = 0;
    uint32_t node_base = 0, tri_base = 0;     // where it goes in the scene-wide arrays
    DevBlasPiece() = default;
    DevBlasPiece(DevBlasPiece&& o) noexcept { *this = std::move(o); }
    DevBlasPiece& operator=(DevBlasPiece&& o) noexcept {
        release(); d_nodes = o.d_nodes; d_tris = o.d_tris; n_nodes = o.n_nodes; n_tris = o.n_tris; node_base = o.node_base; tri_base = o.tri_base;
        o.d_nodes = o.d_tris = nullptr; return *this;
    }
    DevBlasPiece(const DevBlasPiece&) = delete;
    DevBlasPiece& operator=(const DevBlasPiece&) = delete;
    ~DevBlasPiece(){ release(); }
    void release();               // through g_dev_free: scene_build.cpp is also linked into host-only test code, which never holds a piece
};
extern void (*g_dev_free)(void*);     // set by scene_pack_gpu.cu to a cudaFree wrapper
// a run of the host vectors below and where it goes in the scene-wide arrays (everything, at offset 0, when no BLAS was packed on the device)
struct HostSeg { size_t host_off, count, global_base; };

struct PackedTraversal {
    std::vector<Bvh2Node> n2;
    std::vector<Bvh4Node> n4;
    std::vector<Bvh8Node> n8;
    std::vector<Tri> tris;
    // scene-wide layout: node array = host segments + device pieces, likewise the triangles.  Refs are scene-wide indices.
    std::vector<HostSeg> node_segs, tri_segs;
    std::vector<DevBlasPiece> dev_pieces;
    size_t total_nodes = 0, total_tris = 0;
    std::vector<Instance> insts;      // TLAS leaf order
    std::vector<uint32_t> inst_order; // leaf order -> instance id
    uint32_t tlas_root = REF_EMPTY;
    uint32_t single_root = REF_EMPTY, single_id = 0;
    Box bounds = box_empty();
};

// scene_build.cpp (host only)
int scene_add_texture(pt_scene* s, const pt_texture_desc* d);
int scene_add_material(pt_scene* s, const pt_material_desc* d);
int scene_add_medium(pt_scene* s, const pt_medium_desc* d);
int scene_add_mesh(pt_scene* s, const pt_mesh_desc* d);
int scene_add_blas(pt_scene* s, const int32_t* mesh_ids, int n);
int scene_add_instance(pt_scene* s, int blas_id, const float* xf);
// Streamed ray cast: what the persistent kernel shares with the host while the rays are still being copied in
struct StreamCtl {
    const unsigned long long* landed = nullptr;   // rays [0, *landed) have arrived (written on the H2D stream after each chunk)
    unsigned* done = nullptr;                     // per granule of 2^shift rays: how many are retired
    unsigned* abort = nullptr;                    // raised by the kernel when a wait timed out, or by the host
    unsigned shift = 18;
};
// A launch whose batch size and queue are decided on the device (the wavefront loop of render.cu): n = *n_dev rays,
// read from rays + *sel_dev * sel_stride (float4 units).  All null / zero: the host-side arguments are used.
struct DynBatch {
    const uint32_t* n_dev = nullptr;
    const uint32_t* sel_dev = nullptr;
    size_t sel_stride = 0;
};
// Any-hit launches of the wavefront loop: a shadow ray that turns out unoccluded adds its contribution to its path's
// radiance when it retires (the k_connect pass of round 1, fused into the kernel that produces the answer).
// A slot has at most one shadow ray per bounce, so the read-modify-write needs no atomics.
struct ConnectAcc {
    const uint32_t* slot = nullptr;     // per shadow ray: the path slot
    const void* contrib = nullptr;      // per shadow ray: rgb contribution (float4)
    void* ps = nullptr;                 // path state (float4 units); radiance of slot i at ps[i * ps_slot + ps_off]
    uint32_t ps_slot = 0, ps_off = 0;
};
bool probe_async_launch(pt_scene* sc, void* d_words, void* compute_stream, void* copy_stream);
int launch_trace_streamed(pt_scene* sc, bool any, const void* d_rays, size_t n, void* d_hits, void* d_inst, void* d_occ, const StreamCtl& ctl, void* stream, bool compact = false);
int launch_expand_rays24(const void* d_rays24, void* d_rays, size_t n, void* stream);
int launch_pack_hits8(const void* d_hits, void* d_hits8, size_t n, void* stream);

int scene_add_shape(pt_scene* s, const pt_shape_desc* d, const float* xf = nullptr);
int scene_animate(pt_scene* s, int prim_id, const float* dir, const float* tb);
int scene_add_light(pt_scene* s, const pt_light_desc* d);
// builder: binary build used for every BLAS and the TLAS (null = build_binary_bvh on the host); returns PT_OK or an error
using BvhBuilder = int (*)(const Box* boxes, uint32_t n, BinaryBvh& out, void* user);
// packs one BLAS on the device: PT_OK and `out` filled (bounds = its root box), or 1 = not taken (the host path packs it), or an error
using BlasPacker = int (*)(pt_scene* s, const HostBlas& hb, int width, DevBlasPiece& out, Box& bounds, void* user);
// PT_LAYOUT_FAST: quantised copy of the BVH4 clusters (pt_types.h Bvh4QNode); every decoded box contains the exact one
void encode_bvh4q(const std::vector<Bvh4Node>& n4, std::vector<Bvh4QNode>& out);
int pack_traversal(pt_scene* s, int width, PackedTraversal& out, BvhBuilder builder = nullptr, void* user = nullptr, BlasPacker packer = nullptr);
uint32_t mesh_tri_flags(const pt_scene* s, int mesh);      // TRI_ALPHA_* flags of the mesh's material (scene_build.cpp)
// pageable host memory to the device through pinned staging buffers (shade_host.cu); synchronous
bool h2d_staged(int device, void* dst, const void* src, size_t bytes, const char* what);
// scene_pack_gpu.cu
int pack_blas_on_device(pt_scene* s, const HostBlas& hb, int width, DevBlasPiece& out, Box& bounds, void* user);
int place_blas_piece(const DevBlasPiece& p, int width, void* d_nodes, void* d_tris);
int encode_fast_range(const void* d_n4, void* d_n4q, uint32_t node_base, uint32_t n_nodes, const void* d_tris, void* d_tris64, uint32_t tri_base, uint32_t n_tris);
int build_binary_bvh_gpu(const Box* boxes, uint32_t n, uint32_t leaf_size, BinaryBvh& out, GpuBuildTimes* times);   // bvh_build_gpu.cu
// a binary tree left on the device by the GPU builder: BuildNode[n_nodes] (node 0 = root, unused pool slots zero) and the item order
struct DeviceBvh2 {
    void* nodes = nullptr; void* order = nullptr;
    uint32_t n_nodes = 0, n = 0;
    DeviceBvh2() = default;
    DeviceBvh2(const DeviceBvh2&) = delete;
    DeviceBvh2& operator=(const DeviceBvh2&) = delete;
    ~DeviceBvh2();
};
int build_binary_bvh_gpu_ex(const Box* boxes, const Box* d_boxes_in, uint32_t n, uint32_t leaf_size, BinaryBvh* out, DeviceBvh2* keep, GpuBuildTimes* times);
// BinaryBvh with nodes renumbered in DFS preorder (left subtree first), the host builder's own numbering
// below 256 Ki items; makes trees from different builders comparable array by array
BinaryBvh canonical_bvh(const BinaryBvh& b);

void set_error(const std::string& msg);
const char* get_error();
bool cuda_ok(int err, const char* what);     // records the message on failure

int scene_commit(pt_scene* s, int width, int light_sampler);
void note_upload(pt_scene* s, void** slot, size_t bytes);     // slot points into s->buf
void bind_device_pointers(pt_scene* s);                      // s->dev / s->shade pointer fields from s->buf
int scene_replicate(pt_scene* src, int device, pt_scene** out);   // multi.cu
void scene_free_device(pt_scene* s);

// trace.cu
int launch_trace_closest(pt_scene* s, const void* d_rays, size_t n, void* d_hits, void* d_inst, void* stream);
int launch_trace_any(pt_scene* s, const void* d_rays, size_t n, void* d_occ, void* stream);
int launch_trace_stats(pt_scene* s, const void* d_rays, size_t n, void* d_counters /* 6 x u64 */, void* stream);
// wavefront form: full persistent grid, batch described on the device, the caller owns (and clears) the cursor
int launch_trace_closest_dyn(pt_scene* s, const void* d_rays, const DynBatch& dyn, void* d_hits, void* d_inst, unsigned long long* cursor, void* stream);
int launch_trace_any_dyn(pt_scene* s, const void* d_rays, const DynBatch& dyn, void* d_occ, const ConnectAcc& acc, unsigned long long* cursor, void* stream);
// traversal-stack overflows counted on this scene's device since commit (synchronises the device)
int scene_overflow_count(pt_scene* s, unsigned long long* out);

// render.cu
int render_passes(pt_scene* s, const pt_camera_desc* cam, const pt_render_params* rp, int clear, pt_render_stats* stats);

} // namespace ptb
