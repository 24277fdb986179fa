// scene_pack_gpu.cu — one BLAS packed entirely on the device: triangle boxes, the binned-SAH binary build
// (bvh_build_gpu.cu), the collapse into BVH2 child-pair nodes / BVH4 / BVH8 clusters (BVH.hpp:461-473, 788-1017, 1312-1372; per-cluster
// logic shared with the host flatten through bvh_flatten4.hpp / bvh_flatten28.hpp) and the triangle records in leaf order.  What the host path of scene_build.cpp
// pack_traversal does for a BLAS, with the same decisions and the same numbering (DFS preorder, children in lane
// order), so the arrays are the ones the host would have produced; the binary tree is never downloaded and the
// host never touches a per-triangle array (a 10 M-triangle commit spent 0.33 s in that download, 0.2 s flattening
// and about a second forming, gathering and first-touching 1.2 GB of host records).
//
// The collapse is top-down, the numbering needs subtree sizes: three level-synchronous sweeps over the clusters in
// breadth-first order -- expand (which binary nodes become lanes; children appended to the next level), count
// (bottom-up: clusters below each cluster), assign (top-down: preorder index = parent's + 1 + sizes of the earlier
// siblings' subtrees) -- then one pass writes every cluster to its final place.
#include <cuda_runtime.h>
#include <math_constants.h>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "scene_host.hpp"
#include "bvh_flatten4.hpp"
#include "bvh_flatten28.hpp"
#include "bvh4q_encode.hpp"

namespace ptb {

namespace {

struct DevMem {
    void* p = nullptr;
    ~DevMem(){ if(p)cudaFree(p); }
    bool alloc(size_t bytes, const char* what){ return cuda_ok(cudaMalloc(&p, bytes ? bytes : 16), what); }
    template <class T> T* as(){ return (T*)p; }
    void* take(){ void* q = p; p = nullptr; return q; }
};

// one mesh of the BLAS on the device (its positions and indices, uploaded for the duration of the pack)
struct MeshRef {
    const float* pos; const uint32_t* idx;
    uint32_t first, count;      // BLAS-wide id of its first triangle, number of triangles
    uint32_t mesh, flags;       // scene mesh id, TRI_ALPHA_* flags of its material
};
struct MeshTable { const MeshRef* m; int n; };   // the BLAS's meshes, by first triangle id, in device memory

__device__ __forceinline__ float fmin2(float a, float b){ return b < a ? b : a; }   // glm::min, as bvh_build.cpp
__device__ __forceinline__ float fmax2(float a, float b){ return a < b ? b : a; }

// TriangleShape::BoundingBox (Shape.cpp:270-276)
__global__ void k_tri_boxes(const MeshRef m, Box* __restrict__ boxes){
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if(t >= m.count)return;
    const uint32_t i0 = m.idx[3 * (size_t)t], i1 = m.idx[3 * (size_t)t + 1], i2 = m.idx[3 * (size_t)t + 2];
    Box b;
    for(int a = 0; a < 3; a++){
        const float p0 = m.pos[3 * (size_t)i0 + a], p1 = m.pos[3 * (size_t)i1 + a], p2 = m.pos[3 * (size_t)i2 + a];
        float lo = CUDART_INF_F, hi = -CUDART_INF_F;
        lo = fmin2(lo, p0); hi = fmax2(hi, p0);
        lo = fmin2(lo, p1); hi = fmax2(hi, p1);
        lo = fmin2(lo, p2); hi = fmax2(hi, p2);
        b.lo[a] = lo; b.hi[a] = hi;
    }
    boxes[m.first + t] = b;
}

// A cluster while its final index is not known yet: refs of inner lanes hold the breadth-first index of the child
struct Work { uint32_t root; uint32_t subtree; uint32_t final_index; uint32_t pad; };

// The three widths behind one set of sweeps: the node type, its refs, and the cluster of one binary node (the shared
// host / device functions of bvh_flatten4.hpp and bvh_flatten28.hpp).  make() leaves every ref REF_EMPTY and says which
// binary node sits in which lane.
template <int W> struct Wide;
template <> struct Wide<2> {
    using Node = Bvh2Node;
    static constexpr int kLanes = 2;
    __device__ static uint32_t& ref(Node& o, int k){ return k ? o.ref1 : o.ref0; }
    __device__ static void make(const BuildNode* nodes, uint32_t root, Node& o, uint32_t lane_node[8], bool lane_used[8]){
        fill_cluster2(nodes, root, o, lane_node);
        lane_used[0] = lane_used[1] = true;
    }
};
template <> struct Wide<4> {
    using Node = Bvh4Node;
    static constexpr int kLanes = 4;
    __device__ static uint32_t& ref(Node& o, int k){ return o.ref[k]; }
    __device__ static void make(const BuildNode* nodes, uint32_t root, Node& o, uint32_t lane_node[8], bool lane_used[8]){
        ClusterPlan p;
        plan_cluster4(nodes, root, p);
        fill_cluster4(nodes, p, o, lane_node, lane_used);
    }
};
template <> struct Wide<8> {
    using Node = Bvh8Node;
    static constexpr int kLanes = 8;
    __device__ static uint32_t& ref(Node& o, int k){ return o.ref[k]; }
    __device__ static void make(const BuildNode* nodes, uint32_t root, Node& o, uint32_t lane_node[8], bool lane_used[8]){
        const int end = fill_cluster8(nodes, root, o, lane_node);
        for(int k = 0; k < 8; k++)lane_used[k] = k < end;
    }
};

// expand one level: bfs[lo .. hi) -> their clusters, children appended at bfs[hi + ...)
template <int W>
__global__ void k_expand(const BuildNode* __restrict__ nodes, Work* __restrict__ work, typename Wide<W>::Node* __restrict__ tmp, uint32_t lo, uint32_t hi,
                         uint32_t* __restrict__ next_count, uint32_t* __restrict__ leaf_last, uint32_t cap){
    const uint32_t i = lo + blockIdx.x * blockDim.x + threadIdx.x;
    if(i >= hi)return;
    typename Wide<W>::Node o;
    uint32_t lane_node[8];
    bool lane_used[8];
    Wide<W>::make(nodes, work[i].root, o, lane_node, lane_used);
    uint32_t inner = 0;
    for(int k = 0; k < Wide<W>::kLanes; k++)if(lane_used[k] && nodes[lane_node[k]].count == 0)inner++;
    uint32_t at = inner ? hi + atomicAdd(next_count, inner) : 0u;
    for(int k = 0; k < Wide<W>::kLanes; k++){
        if(!lane_used[k])continue;
        const BuildNode& c = nodes[lane_node[k]];
        if(c.count != 0){
            Wide<W>::ref(o, k) = REF_LEAF | c.right;        // first item of the leaf, BLAS-local
            leaf_last[c.right + c.count - 1] = 1u;
        } else{
            if(at < cap){ work[at].root = lane_node[k]; work[at].subtree = 0; work[at].final_index = 0; work[at].pad = 0; }
            Wide<W>::ref(o, k) = at++;
        }
    }
    tmp[i] = o;
}

// clusters in the subtree of each cluster of one level (its children's counts are final: deeper levels ran first)
template <int W>
__global__ void k_count(Work* __restrict__ work, typename Wide<W>::Node* __restrict__ tmp, uint32_t lo, uint32_t hi){
    const uint32_t i = lo + blockIdx.x * blockDim.x + threadIdx.x;
    if(i >= hi)return;
    uint32_t c = 1;
    for(int k = 0; k < Wide<W>::kLanes; k++){ const uint32_t r = Wide<W>::ref(tmp[i], k); if(!(r & REF_LEAF))c += work[r].subtree; }
    work[i].subtree = c;
}

// preorder index of the children of each cluster of one level
template <int W>
__global__ void k_assign(Work* __restrict__ work, typename Wide<W>::Node* __restrict__ tmp, uint32_t lo, uint32_t hi){
    const uint32_t i = lo + blockIdx.x * blockDim.x + threadIdx.x;
    if(i >= hi)return;
    uint32_t run = work[i].final_index + 1;
    for(int k = 0; k < Wide<W>::kLanes; k++){
        const uint32_t r = Wide<W>::ref(tmp[i], k);
        if(r & REF_LEAF)continue;                           // leaves and REF_EMPTY
        work[r].final_index = run;
        run += work[r].subtree;
    }
}

template <int W>
__global__ void k_emit(const Work* __restrict__ work, const typename Wide<W>::Node* __restrict__ tmp, typename Wide<W>::Node* __restrict__ out, uint32_t n){
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if(i >= n)return;
    typename Wide<W>::Node o = tmp[i];
    for(int k = 0; k < Wide<W>::kLanes; k++){ uint32_t& r = Wide<W>::ref(o, k); if(!(r & REF_LEAF))r = work[r].final_index; }
    out[work[i].final_index] = o;
}

// triangle records in leaf order: v0 and the two edges as glm::intersectRayTriangle forms them (one subtraction each)
__global__ void k_gather_tris(const MeshTable T, const uint32_t* __restrict__ order, const uint32_t* __restrict__ leaf_last, uint32_t n, Tri* __restrict__ out){
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if(i >= n)return;
    const uint32_t g = order[i];
    int k = 0, hi = T.n - 1;                         // the last mesh whose first triangle id is <= g
    while(k < hi){ const int mid = (k + hi + 1) >> 1; if(T.m[mid].first <= g)k = mid; else hi = mid - 1; }
    const MeshRef m = T.m[k];
    const uint32_t t = g - m.first;
    const uint32_t i0 = m.idx[3 * (size_t)t], i1 = m.idx[3 * (size_t)t + 1], i2 = m.idx[3 * (size_t)t + 2];
    Tri r;
    for(int a = 0; a < 3; a++){
        const float p0 = m.pos[3 * (size_t)i0 + a], p1 = m.pos[3 * (size_t)i1 + a], p2 = m.pos[3 * (size_t)i2 + a];
        r.v0[a] = p0; r.e1[a] = __fsub_rn(p1, p0); r.e2[a] = __fsub_rn(p2, p0);
    }
    r.prim = g; r.mesh = m.mesh; r.flags = m.flags | (leaf_last[i] ? TRI_LAST : 0u);
    out[i] = r;
}

// ---- assembling the scene-wide arrays (scene_host.cu) ----------------------------------------------------------------
template <int W>
__global__ void k_place_nodes(const typename Wide<W>::Node* __restrict__ in, typename Wide<W>::Node* __restrict__ out, uint32_t n, uint32_t node_base, uint32_t tri_base){
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if(i >= n)return;
    typename Wide<W>::Node o = in[i];
    for(int k = 0; k < Wide<W>::kLanes; k++){
        uint32_t& r = Wide<W>::ref(o, k);
        if(r == REF_EMPTY)continue;
        r = (r & REF_LEAF) ? (REF_LEAF | ((r & ~REF_LEAF) + tri_base)) : r + node_base;
    }
    out[i] = o;
}
__global__ void k_encode4q(const Bvh4Node* __restrict__ in, Bvh4QNode* __restrict__ out, uint32_t n){
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if(i >= n)return;
    Bvh4QNode q;
    encode_node4q(in[i], q);
    out[i] = q;
}
__global__ void k_pad_tris(const Tri* __restrict__ in, Tri64* __restrict__ out, uint32_t n){
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if(i >= n)return;
    Tri64 t;
    t.t = in[i];
    t.pad[0] = t.pad[1] = t.pad[2] = t.pad[3] = 0u;
    out[i] = t;
}

void dev_free(void* p){ cudaFree(p); }

inline unsigned blocks(size_t n){ return (unsigned)((n + 255) / 256); }

// collapse: breadth-first expand, then sizes, then preorder indices; then the clusters at their final places and the triangle records
template <int W>
static int collapse_on_device(const BuildNode* d_nodes, uint32_t n, const MeshTable& T, const uint32_t* d_order, cudaStream_t st,
                              uint32_t& n_clusters_out, void*& d_out_nodes, DevMem& d_tris){
    const uint32_t cap = n;                      // clusters <= inner binary nodes < items
    DevMem d_work, d_tmp, d_last, d_cnt;
    if(!d_work.alloc(sizeof(Work) * (size_t)cap, "cudaMalloc collapse work") || !d_tmp.alloc(sizeof(typename Wide<W>::Node) * (size_t)cap, "cudaMalloc clusters") ||
       !d_last.alloc(4 * (size_t)n, "cudaMalloc leaf marks") || !d_cnt.alloc(4, "cudaMalloc counter"))return PT_ERR_CUDA;
    if(!cuda_ok(cudaMemsetAsync(d_last.p, 0, 4 * (size_t)n, st), "leaf marks clear"))return PT_ERR_CUDA;
    const Work w0 = { 0u, 0u, 0u, 0u };
    if(!cuda_ok(cudaMemcpyAsync(d_work.p, &w0, sizeof w0, cudaMemcpyHostToDevice, st), "H2D root cluster"))return PT_ERR_CUDA;
    std::vector<uint32_t> level_lo;              // level l = [level_lo[l], level_lo[l + 1])
    level_lo.push_back(0);
    uint32_t lo = 0, hi = 1;
    while(hi > lo){
        level_lo.push_back(hi);
        if(!cuda_ok(cudaMemsetAsync(d_cnt.p, 0, 4, st), "counter clear"))return PT_ERR_CUDA;
        k_expand<W><<<blocks(hi - lo), 256, 0, st>>>(d_nodes, d_work.as<Work>(), d_tmp.as<typename Wide<W>::Node>(), lo, hi, d_cnt.as<uint32_t>(), d_last.as<uint32_t>(), cap);
        uint32_t next = 0;
        if(!cuda_ok(cudaGetLastError(), "k_expand") || !cuda_ok(cudaMemcpyAsync(&next, d_cnt.p, 4, cudaMemcpyDeviceToHost, st), "D2H level size") ||
           !cuda_ok(cudaStreamSynchronize(st), "collapse level"))return PT_ERR_CUDA;
        if((uint64_t)hi + next > cap){ set_error("gpu collapse: more clusters than items"); return PT_ERR_CUDA; }
        lo = hi; hi += next;
    }
    const uint32_t n_clusters = hi;
    const int levels = (int)level_lo.size() - 1;
    for(int l = levels - 1; l >= 0; l--){
        const uint32_t a = level_lo[l], b = level_lo[l + 1];
        k_count<W><<<blocks(b - a), 256, 0, st>>>(d_work.as<Work>(), d_tmp.as<typename Wide<W>::Node>(), a, b);
    }
    for(int l = 0; l < levels; l++){
        const uint32_t a = level_lo[l], b = level_lo[l + 1];
        k_assign<W><<<blocks(b - a), 256, 0, st>>>(d_work.as<Work>(), d_tmp.as<typename Wide<W>::Node>(), a, b);
    }
    DevMem d_out;
    if(!d_out.alloc(sizeof(typename Wide<W>::Node) * (size_t)n_clusters, "cudaMalloc BLAS clusters") || !d_tris.alloc(sizeof(Tri) * (size_t)n, "cudaMalloc BLAS triangles"))return PT_ERR_CUDA;
    k_emit<W><<<blocks(n_clusters), 256, 0, st>>>(d_work.as<Work>(), d_tmp.as<typename Wide<W>::Node>(), d_out.as<typename Wide<W>::Node>(), n_clusters);
    k_gather_tris<<<blocks(n), 256, 0, st>>>(T, d_order, d_last.as<uint32_t>(), n, d_tris.as<Tri>());
    if(!cuda_ok(cudaGetLastError(), "collapse kernels") || !cuda_ok(cudaStreamSynchronize(st), "gpu pack"))return PT_ERR_CUDA;
    n_clusters_out = n_clusters;
    d_out_nodes = d_out.take();
    return PT_OK;
}

} // namespace

// BlasPacker (scene_host.hpp).  user: GpuBuildTimes* of the commit.
int pack_blas_on_device(pt_scene* s, const HostBlas& hb, int width, DevBlasPiece& out, Box& bounds, void* user){
    static const uint32_t gpu_min = []{ const char* e = getenv("PTB_GPU_BUILD_MIN"); return e ? (uint32_t)atoll(e) : 4096u; }();
    static const bool off = []{ const char* e = getenv("PTB_GPU_PACK"); return e && atoi(e) == 0; }();
    if(off || hb.n_tris < gpu_min || hb.n_tris < 8)return 1;
    g_dev_free = dev_free;
    GpuBuildTimes* times = (GpuBuildTimes*)user;
    const uint32_t n = hb.n_tris;
    cudaStream_t st = 0;

    // meshes up, boxes
    std::vector<MeshRef> refs;
    std::vector<DevMem> mesh_mem(2 * hb.meshes.size());
    DevMem d_boxes;
    if(!d_boxes.alloc(sizeof(Box) * (size_t)n, "cudaMalloc boxes"))return PT_ERR_CUDA;
    for(size_t k = 0; k < hb.meshes.size(); k++){
        const int mi = hb.meshes[k];
        const HostMesh& m = s->meshes[mi];
        DevMem& dp = mesh_mem[2 * k];
        DevMem& di = mesh_mem[2 * k + 1];
        if(!dp.alloc(m.pos.size() * sizeof(float), "cudaMalloc positions") || !di.alloc(m.idx.size() * sizeof(uint32_t), "cudaMalloc indices") ||
           !h2d_staged(s->device, dp.p, m.pos.data(), m.pos.size() * sizeof(float), "H2D positions") ||
           !h2d_staged(s->device, di.p, m.idx.data(), m.idx.size() * sizeof(uint32_t), "H2D indices"))return PT_ERR_CUDA;
        MeshRef r;
        r.pos = dp.as<float>(); r.idx = di.as<uint32_t>();
        r.first = hb.tri_mesh_base[k]; r.count = m.d.n_triangles; r.mesh = (uint32_t)mi; r.flags = mesh_tri_flags(s, mi);
        if(r.count == 0)continue;                   // an empty mesh owns no triangle id
        k_tri_boxes<<<blocks(r.count), 256, 0, st>>>(r, d_boxes.as<Box>());
        refs.push_back(r);
    }
    DevMem d_refs;
    if(!cuda_ok(cudaGetLastError(), "k_tri_boxes") || !d_refs.alloc(sizeof(MeshRef) * refs.size(), "cudaMalloc mesh table") ||
       !cuda_ok(cudaMemcpyAsync(d_refs.p, refs.data(), sizeof(MeshRef) * refs.size(), cudaMemcpyHostToDevice, st), "H2D mesh table"))return PT_ERR_CUDA;
    const MeshTable T = { d_refs.as<MeshRef>(), (int)refs.size() };

    // binary tree, left on the device
    DeviceBvh2 tree;
    int rc = build_binary_bvh_gpu_ex(nullptr, d_boxes.as<Box>(), n, 2, nullptr, &tree, times);
    if(rc != PT_OK)return rc;
    cudaFree(d_boxes.take());
    const BuildNode* d_nodes = (const BuildNode*)tree.nodes;
    BuildNode root;
    if(!cuda_ok(cudaMemcpy(&root, d_nodes, sizeof root, cudaMemcpyDeviceToHost), "D2H root"))return PT_ERR_CUDA;
    if(root.count != 0)return 1;                 // a single leaf: nothing to collapse (cannot happen above gpu_min)
    bounds = root.box;

    const auto t0 = std::chrono::steady_clock::now();
    uint32_t n_clusters = 0;
    void* d_out_nodes = nullptr;
    DevMem d_tris;
    rc = width == 2 ? collapse_on_device<2>(d_nodes, n, T, (const uint32_t*)tree.order, st, n_clusters, d_out_nodes, d_tris)
       : width == 4 ? collapse_on_device<4>(d_nodes, n, T, (const uint32_t*)tree.order, st, n_clusters, d_out_nodes, d_tris)
                    : collapse_on_device<8>(d_nodes, n, T, (const uint32_t*)tree.order, st, n_clusters, d_out_nodes, d_tris);
    if(rc != PT_OK)return rc;
    s->build_stats.flatten_s += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    out.d_nodes = d_out_nodes; out.d_tris = d_tris.take();
    out.n_nodes = n_clusters; out.n_tris = n;
    return PT_OK;
}

// the piece's clusters into the scene-wide array, refs rebased (scene_host.cu)
int place_blas_piece(const DevBlasPiece& p, int width, void* d_nodes, void* d_tris){
    if(width == 2)k_place_nodes<2><<<blocks(p.n_nodes), 256>>>((const Bvh2Node*)p.d_nodes, (Bvh2Node*)d_nodes + p.node_base, p.n_nodes, p.node_base, p.tri_base);
    else if(width == 4)k_place_nodes<4><<<blocks(p.n_nodes), 256>>>((const Bvh4Node*)p.d_nodes, (Bvh4Node*)d_nodes + p.node_base, p.n_nodes, p.node_base, p.tri_base);
    else k_place_nodes<8><<<blocks(p.n_nodes), 256>>>((const Bvh8Node*)p.d_nodes, (Bvh8Node*)d_nodes + p.node_base, p.n_nodes, p.node_base, p.tri_base);
    if(!cuda_ok(cudaGetLastError(), "k_place_nodes") ||
       !cuda_ok(cudaMemcpy((Tri*)d_tris + p.tri_base, p.d_tris, sizeof(Tri) * (size_t)p.n_tris, cudaMemcpyDeviceToDevice), "D2D triangles"))return PT_ERR_CUDA;
    return PT_OK;
}

// PT_LAYOUT_FAST for a range that was packed on the device: quantised clusters and padded triangles from the exact ones
int encode_fast_range(const void* d_n4, void* d_n4q, uint32_t node_base, uint32_t n_nodes, const void* d_tris, void* d_tris64, uint32_t tri_base, uint32_t n_tris){
    if(n_nodes)k_encode4q<<<blocks(n_nodes), 256>>>((const Bvh4Node*)d_n4 + node_base, (Bvh4QNode*)d_n4q + node_base, n_nodes);
    if(n_tris)k_pad_tris<<<blocks(n_tris), 256>>>((const Tri*)d_tris + tri_base, (Tri64*)d_tris64 + tri_base, n_tris);
    return cuda_ok(cudaGetLastError(), "fast layout kernels") ? PT_OK : PT_ERR_CUDA;
}

} // namespace ptb
