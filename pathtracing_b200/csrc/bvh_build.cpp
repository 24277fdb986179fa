// bvh_build.cpp — see bvh_build.hpp.  Compiled with -ffp-contract=off: the SAH
// costs decide the topology and must not depend on FMA contraction.
#include "bvh_build.hpp"
#include "bvh_flatten4.hpp"
#include "bvh_flatten28.hpp"
#include "bvh4q_encode.hpp"
#include "scene_host.hpp"
#include <algorithm>
#include <cstring>

#include <atomic>
#include <cfloat>
#include <cmath>
#include <limits>
#include <thread>

namespace ptb {

static inline float fmin2(float a, float b){ return b < a ? b : a; }   // glm::min
static inline float fmax2(float a, float b){ return a < b ? b : a; }   // glm::max

Box box_empty(){
    Box b;
    const float inf = std::numeric_limits<float>::infinity();
    for(int k = 0; k < 3; k++){ b.lo[k] = inf; b.hi[k] = -inf; }
    return b;
}
void box_grow(Box& b, const Box& o){
    for(int k = 0; k < 3; k++){ b.lo[k] = fmin2(b.lo[k], o.lo[k]); b.hi[k] = fmax2(b.hi[k], o.hi[k]); }
}
void box_grow(Box& b, const float p[3]){
    for(int k = 0; k < 3; k++){ b.lo[k] = fmin2(b.lo[k], p[k]); b.hi[k] = fmax2(b.hi[k], p[k]); }
}
float box_half_area(const Box& b){
    const float ex = b.hi[0] - b.lo[0], ey = b.hi[1] - b.lo[1], ez = b.hi[2] - b.lo[2];
    return ex * ey + ey * ez + ez * ex;
}

namespace {

struct Item {
    uint32_t index;
    Box box;
    float c[3];   // centroid = 0.5f * (hi + lo)  (BVH.hpp:89)
};

struct BinAcc {
    Box box = box_empty();
    uint32_t n = 0;
};

struct Builder {
    std::vector<Item>& items;
    std::vector<BuildNode>& nodes;
    std::atomic<uint32_t> next { 0 };
    uint32_t leaf_size;

    Builder(std::vector<Item>& it, std::vector<BuildNode>& nd, uint32_t ls) : items(it), nodes(nd), leaf_size(ls){}

    // Hoare-style in-place partition with the visiting order of a
    // bidirectional std::partition, so item order inside leaves (and with it
    // the winner among exactly tied hits) is the reference's.
    uint32_t split_items(uint32_t first, uint32_t last, int axis, float pos){
        while(true){
            while(true){
                if(first == last)return first;
                if(items[first].c[axis] <= pos)++first; else break;
            }
            --last;
            while(true){
                if(first == last)return first;
                if(!(items[last].c[axis] <= pos))--last; else break;
            }
            std::swap(items[first], items[last]);
            ++first;
        }
    }

    uint32_t build(uint32_t first, uint32_t last){
        const uint32_t me = next.fetch_add(1, std::memory_order_relaxed);
        BuildNode& node = nodes[me];
        const uint32_t span = last - first;

        float cmax[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
        float cmin[3] = { FLT_MAX, FLT_MAX, FLT_MAX };
        node.box = box_empty();
        for(uint32_t i = first; i < last; i++){
            box_grow(node.box, items[i].box);
            for(int k = 0; k < 3; k++){
                cmax[k] = fmax2(cmax[k], items[i].c[k]);
                cmin[k] = fmin2(cmin[k], items[i].c[k]);
            }
        }
        node.right = first;
        node.count = span;
        if(span <= leaf_size)return me;

        int best_axis = 0;
        float best_pos = 0.0f;
        float best_cost = std::numeric_limits<float>::infinity();
        const uint32_t nbins = span >= 1024 ? 32u : span >= 64 ? 16u : 8u;
        BinAcc bins[32];
        float right_area[31];

        for(int axis = 0; axis < 3; axis++){
            const float mx = cmax[axis], mn = cmin[axis];
            if(std::fabs(mx - mn) <= FLT_EPSILON)continue;
            for(uint32_t b = 0; b < nbins; b++)bins[b] = BinAcc();
            float scale = nbins / (mx - mn);
            for(uint32_t i = first; i < last; i++){
                int id = (int)((items[i].c[axis] - mn) * scale);
                if(id > (int)nbins - 1)id = (int)nbins - 1;
                ++bins[id].n;
                box_grow(bins[id].box, items[i].box);
            }
            Box rbox = box_empty(), lbox = box_empty();
            for(uint32_t b = nbins - 1; b >= 1; b--){
                box_grow(rbox, bins[b].box);
                right_area[b - 1] = box_half_area(rbox);
            }
            scale = (mx - mn) / nbins;
            uint32_t lsum = 0;
            for(uint32_t b = 0; b + 1 < nbins; b++){
                lsum += bins[b].n;
                box_grow(lbox, bins[b].box);
                // an empty side has area = inf*inf+... and count 0 -> NaN cost,
                // which never compares < best (same as the reference)
                const float cost = lsum * box_half_area(lbox) + (span - lsum) * right_area[b];
                if(cost < best_cost){
                    best_axis = axis;
                    best_pos = mn + (b + 1) * scale;
                    best_cost = cost;
                }
            }
        }
        const float parent_cost = box_half_area(node.box) * span;
        if(best_cost >= parent_cost)return me;

        const uint32_t mid = split_items(first, last, best_axis, best_pos);
        if(mid == first || mid == last)return me;

        uint32_t l, r;
        if(span > 256u * 1024u){
            std::thread t([&]{ l = build(first, mid); });
            r = build(mid, last);
            t.join();
        } else{
            l = build(first, mid);
            r = build(mid, last);
        }
        // `node` stays valid: the vector is pre-sized and never reallocates
        node.left = l;
        node.right = r;
        node.count = 0;
        node.axis = (uint32_t)best_axis;
        return me;
    }
};

} // namespace

BinaryBvh build_binary_bvh(const Box* boxes, uint32_t n, uint32_t leaf_size){
    BinaryBvh out;
    if(n == 0)return out;
    std::vector<Item> items(n);
    for(uint32_t i = 0; i < n; i++){
        items[i].index = i;
        items[i].box = boxes[i];
        for(int k = 0; k < 3; k++)items[i].c[k] = 0.5f * (boxes[i].hi[k] + boxes[i].lo[k]);
    }
    out.nodes.resize(2 * (size_t)n - 1);
    Builder b(items, out.nodes, leaf_size);
    b.build(0, n);
    out.nodes.resize(b.next.load());
    out.order.resize(n);
    for(uint32_t i = 0; i < n; i++)out.order[i] = items[i].index;
    return out;
}

// ---- flattening ---------------------------------------------------------------------

namespace {

inline uint32_t leaf_ref(uint32_t first){ return REF_LEAF | first; }

void mark_leaf(FlatBvh& f, const BuildNode& n){
    f.last_out[n.right + n.count - 1] = 1;
}

// Items below binary node ni: the build partitions `order` in place, so a subtree owns one contiguous range of
// it, from the first item of its leftmost leaf to the end of its rightmost one (O(depth); used for load balancing only).
uint32_t subtree_items(const BinaryBvh& b, uint32_t ni){
    uint32_t l = ni, r = ni;
    while(b.nodes[l].count == 0)l = b.nodes[l].left;
    while(b.nodes[r].count == 0)r = b.nodes[r].right;
    const uint32_t lo = std::min(b.nodes[l].right, b.nodes[r].right);
    const uint32_t hi = std::max(b.nodes[l].right + b.nodes[l].count, b.nodes[r].right + b.nodes[r].count);
    return hi - lo;
}

constexpr uint32_t REF_JOB = 0x40000000u;     // top pass only: placeholder for the root of a deferred subtree

} // namespace

// Threaded flatten: the top pass runs the usual recursion but hands every subtree of at most `threshold` items to a
// worker, which flattens it into arrays of its own.  Node numbering is DFS preorder, so a deferred subtree occupies
// one contiguous index range whose start is known once the sizes of everything before it are: the stitch below
// reproduces the arrays of the one-thread recursion exactly.
struct FlattenDefer {
    uint32_t threshold = 0;
    struct Job { uint32_t node, after_top; };   // after_top: top nodes allocated before the subtree was met
    std::vector<Job> jobs;
};

namespace {

template <class F> inline void each_ref(Bvh2Node& n, F fn){ fn(n.ref0); fn(n.ref1); }
template <class F> inline void each_ref(Bvh4Node& n, F fn){ for(int k = 0; k < 4; k++)fn(n.ref[k]); }
template <class F> inline void each_ref(Bvh8Node& n, F fn){ for(int k = 0; k < 8; k++)fn(n.ref[k]); }

// true (and *ref set) if ni was handed to a worker
inline bool deferred(FlatBvh& f, const BinaryBvh& b, uint32_t ni, size_t top_nodes, uint32_t* ref){
    if(!f.defer || subtree_items(b, ni) > f.defer->threshold)return false;
    *ref = REF_JOB | (uint32_t)f.defer->jobs.size();
    f.defer->jobs.push_back({ ni, (uint32_t)top_nodes });
    return true;
}

// BVH2: DFS preorder over inner nodes; node k holds the boxes + refs of its two
// children (what BVH2::Intersect reads from nodes[idx+1] and nodes[right]).
uint32_t flatten2(FlatBvh& f, const BinaryBvh& b, uint32_t ni){
    const BuildNode& n = b.nodes[ni];
    if(n.count != 0){ mark_leaf(f, n); return leaf_ref(n.right); }
    { uint32_t job; if(deferred(f, b, ni, f.n2.size(), &job))return job; }
    const uint32_t me = (uint32_t)f.n2.size();
    f.n2.emplace_back();
    Bvh2Node o;
    uint32_t lane_node[2];
    fill_cluster2(b.nodes.data(), ni, o, lane_node);        // bvh_flatten28.hpp, shared with the device flatten
    o.ref0 = flatten2(f, b, lane_node[0]);
    o.ref1 = flatten2(f, b, lane_node[1]);
    f.n2[me] = o;
    return me;
}

// BVH4 collapse: the per-cluster logic (which binary nodes become lanes, visit orders) lives in bvh_flatten4.hpp,
// shared with the device flatten.  Children are emitted lane by lane in increasing lane order, as the reference recursion does.
uint32_t flatten4(FlatBvh& f, const BinaryBvh& b, uint32_t ni){
    const BuildNode& n = b.nodes[ni];
    if(n.count != 0){ mark_leaf(f, n); return leaf_ref(n.right); }
    { uint32_t job; if(deferred(f, b, ni, f.n4.size(), &job))return job; }
    const uint32_t me = (uint32_t)f.n4.size();
    f.n4.emplace_back();
    ClusterPlan p;
    plan_cluster4(b.nodes.data(), ni, p);
    Bvh4Node o;
    uint32_t lane_node[4];
    bool lane_used[4];
    fill_cluster4(b.nodes.data(), p, o, lane_node, lane_used);
    for(int k = 0; k < 4; k++)if(lane_used[k])o.ref[k] = flatten4(f, b, lane_node[k]);
    f.n4[me] = o;
    return me;
}

// BVH8 collapse: the per-cluster logic (lane choice by area, visit orders) lives in bvh_flatten28.hpp, shared with the device flatten.
uint32_t flatten8(FlatBvh& f, const BinaryBvh& b, uint32_t ni){
    const BuildNode& n = b.nodes[ni];
    if(n.count != 0){ mark_leaf(f, n); return leaf_ref(n.right); }
    { uint32_t job; if(deferred(f, b, ni, f.n8.size(), &job))return job; }
    const uint32_t me = (uint32_t)f.n8.size();
    f.n8.emplace_back();
    Bvh8Node o;
    uint32_t lanes[8];
    const int end = fill_cluster8(b.nodes.data(), ni, o, lanes);
    for(int i = 0; i < end; i++)o.ref[i] = flatten8(f, b, lanes[i]);     // children in lane order, as the reference recursion
    f.n8[me] = o;
    return me;
}

} // namespace

namespace {

template <class Node, class Fn>
void flatten_threaded(FlatBvh& f, const BinaryBvh& b, std::vector<Node> FlatBvh::* arr, Fn flatten, unsigned threads){
    FlattenDefer D;
    D.threshold = std::max<uint32_t>(4096u, (uint32_t)(b.order.size() / (8 * (size_t)threads)));
    f.defer = &D;
    f.root = flatten(f, b, 0u);                 // the root holds more than `threshold` items: never a job itself
    f.defer = nullptr;
    std::vector<Node>& top = f.*arr;
    const size_t nj = D.jobs.size();
    std::vector<FlatBvh> part(nj);
    std::atomic<size_t> next { 0 };
    auto work = [&]{
        for(size_t j; (j = next.fetch_add(1)) < nj; ){
            part[j].last_out = f.last_out;
            flatten(part[j], b, D.jobs[j].node);  // an inner node: becomes node 0 of its part
        }
    };
    std::vector<std::thread> th;
    for(unsigned t = 1; t < threads; t++)th.emplace_back(work);
    work();
    for(std::thread& t : th)t.join();
    // final index of top node t = t + sizes of the jobs met before it; of job j = after_top + sizes of the jobs before j
    std::vector<uint32_t> before(nj + 1, 0), start(nj);
    for(size_t j = 0; j < nj; j++){ before[j + 1] = before[j] + (uint32_t)(part[j].*arr).size(); start[j] = D.jobs[j].after_top + before[j]; }
    auto top_final = [&](uint32_t t){
        size_t lo = 0, hi = nj;                   // jobs with after_top <= t
        while(lo < hi){ const size_t mid = (lo + hi) / 2; if(D.jobs[mid].after_top <= t)lo = mid + 1; else hi = mid; }
        return t + before[lo];
    };
    std::vector<Node> out(top.size() + before[nj]);
    for(size_t t = 0; t < top.size(); t++){
        Node nd = top[t];
        each_ref(nd, [&](uint32_t& r){
            if(r & REF_LEAF)return;               // leaves and REF_EMPTY
            r = (r & REF_JOB) ? start[r & ~REF_JOB] : top_final(r);
        });
        out[top_final((uint32_t)t)] = nd;
    }
    next = 0;
    auto copy = [&]{
        for(size_t j; (j = next.fetch_add(1)) < nj; ){
            const std::vector<Node>& src = part[j].*arr;
            const uint32_t off = start[j];
            for(size_t q = 0; q < src.size(); q++){
                Node nd = src[q];
                each_ref(nd, [&](uint32_t& r){ if(!(r & REF_LEAF))r += off; });
                out[off + q] = nd;
            }
        }
    };
    th.clear();
    for(unsigned t = 1; t < threads; t++)th.emplace_back(copy);
    copy();
    for(std::thread& t : th)t.join();
    if(f.root & REF_JOB)f.root = start[f.root & ~REF_JOB]; else if(!(f.root & REF_LEAF))f.root = top_final(f.root);
    top.swap(out);
}

} // namespace

FlatBvh flatten_bvh(const BinaryBvh& b, int width, unsigned threads){
    FlatBvh f;
    f.width = width;
    f.order = b.order;
    f.leaf_last.assign(b.order.size(), 0);
    f.last_out = f.leaf_last.data();
    if(b.nodes.empty())return f;
    f.bounds = b.nodes[0].box;
    if(threads == 0){
        threads = std::min(16u, std::max(1u, std::thread::hardware_concurrency()));
        if(b.order.size() < ((size_t)1 << 17))threads = 1;
    }
    if(b.nodes.size() >= REF_JOB)threads = 1;     // the placeholder bit must stay clear of node indices
    if(threads > 1 && b.nodes[0].count == 0){
        if(width == 2)flatten_threaded<Bvh2Node>(f, b, &FlatBvh::n2, flatten2, threads);
        else if(width == 4)flatten_threaded<Bvh4Node>(f, b, &FlatBvh::n4, flatten4, threads);
        else flatten_threaded<Bvh8Node>(f, b, &FlatBvh::n8, flatten8, threads);
        f.last_out = nullptr;
        return f;
    }
    if(width == 2){
        f.n2.reserve(b.nodes.size() / 2 + 1);
        f.root = flatten2(f, b, 0);
    } else if(width == 4){
        f.n4.reserve(b.nodes.size() / 3 + 1);
        f.root = flatten4(f, b, 0);
    } else{
        f.n8.reserve(b.nodes.size() / 5 + 1);
        f.root = flatten8(f, b, 0);
    }
    f.last_out = nullptr;
    return f;
}

BinaryBvh canonical_bvh(const BinaryBvh& b){
    BinaryBvh out;
    out.order = b.order;
    if(b.nodes.empty())return out;
    out.nodes.reserve(b.nodes.size());
    struct Item { uint32_t src, parent; bool is_right; };
    std::vector<Item> stack;
    stack.push_back({ 0u, 0xFFFFFFFFu, false });
    while(!stack.empty()){
        const Item it = stack.back(); stack.pop_back();
        const uint32_t me = (uint32_t)out.nodes.size();
        out.nodes.push_back(b.nodes[it.src]);
        if(it.parent != 0xFFFFFFFFu){ if(it.is_right)out.nodes[it.parent].right = me; else out.nodes[it.parent].left = me; }
        const BuildNode& n = b.nodes[it.src];
        if(n.count == 0){ stack.push_back({ n.right, me, true }); stack.push_back({ n.left, me, false }); }
    }
    return out;
}

// ---- small matrix helpers (column-major, glm operation order) -----------------------

void mat4_inverse(const float m[16], float out[16]){
#define M(c, r) m[(c) * 4 + (r)]
    const float c00 = M(2, 2) * M(3, 3) - M(3, 2) * M(2, 3);
    const float c02 = M(1, 2) * M(3, 3) - M(3, 2) * M(1, 3);
    const float c03 = M(1, 2) * M(2, 3) - M(2, 2) * M(1, 3);
    const float c04 = M(2, 1) * M(3, 3) - M(3, 1) * M(2, 3);
    const float c06 = M(1, 1) * M(3, 3) - M(3, 1) * M(1, 3);
    const float c07 = M(1, 1) * M(2, 3) - M(2, 1) * M(1, 3);
    const float c08 = M(2, 1) * M(3, 2) - M(3, 1) * M(2, 2);
    const float c10 = M(1, 1) * M(3, 2) - M(3, 1) * M(1, 2);
    const float c11 = M(1, 1) * M(2, 2) - M(2, 1) * M(1, 2);
    const float c12 = M(2, 0) * M(3, 3) - M(3, 0) * M(2, 3);
    const float c14 = M(1, 0) * M(3, 3) - M(3, 0) * M(1, 3);
    const float c15 = M(1, 0) * M(2, 3) - M(2, 0) * M(1, 3);
    const float c16 = M(2, 0) * M(3, 2) - M(3, 0) * M(2, 2);
    const float c18 = M(1, 0) * M(3, 2) - M(3, 0) * M(1, 2);
    const float c19 = M(1, 0) * M(2, 2) - M(2, 0) * M(1, 2);
    const float c20 = M(2, 0) * M(3, 1) - M(3, 0) * M(2, 1);
    const float c22 = M(1, 0) * M(3, 1) - M(3, 0) * M(1, 1);
    const float c23 = M(1, 0) * M(2, 1) - M(2, 0) * M(1, 1);
    const float f0[4] = { c00, c00, c02, c03 }, f1[4] = { c04, c04, c06, c07 }, f2[4] = { c08, c08, c10, c11 };
    const float f3[4] = { c12, c12, c14, c15 }, f4[4] = { c16, c16, c18, c19 }, f5[4] = { c20, c20, c22, c23 };
    const float v0[4] = { M(1, 0), M(0, 0), M(0, 0), M(0, 0) };
    const float v1[4] = { M(1, 1), M(0, 1), M(0, 1), M(0, 1) };
    const float v2[4] = { M(1, 2), M(0, 2), M(0, 2), M(0, 2) };
    const float v3[4] = { M(1, 3), M(0, 3), M(0, 3), M(0, 3) };
    const float sa[4] = { 1, -1, 1, -1 }, sb[4] = { -1, 1, -1, 1 };
    float inv[16];
    for(int k = 0; k < 4; k++){
        inv[0 * 4 + k] = ((v1[k] * f0[k] - v2[k] * f1[k]) + v3[k] * f2[k]) * sa[k];
        inv[1 * 4 + k] = ((v0[k] * f0[k] - v2[k] * f3[k]) + v3[k] * f4[k]) * sb[k];
        inv[2 * 4 + k] = ((v0[k] * f1[k] - v1[k] * f3[k]) + v3[k] * f5[k]) * sa[k];
        inv[3 * 4 + k] = ((v0[k] * f2[k] - v1[k] * f4[k]) + v2[k] * f5[k]) * sb[k];
    }
    const float d0 = M(0, 0) * inv[0 * 4 + 0], d1 = M(0, 1) * inv[1 * 4 + 0], d2 = M(0, 2) * inv[2 * 4 + 0], d3 = M(0, 3) * inv[3 * 4 + 0];
    const float det = (d0 + d1) + (d2 + d3);
    const float inv_det = 1.0f / det;
    for(int k = 0; k < 16; k++)out[k] = inv[k] * inv_det;
#undef M
}

void mat3_normal_matrix(const float m4[16], float out[9]){
#define M(c, r) m4[(c) * 4 + (r)]
    const float ood = 1.0f / (
        +M(0, 0) * (M(1, 1) * M(2, 2) - M(2, 1) * M(1, 2))
        - M(1, 0) * (M(0, 1) * M(2, 2) - M(2, 1) * M(0, 2))
        + M(2, 0) * (M(0, 1) * M(1, 2) - M(1, 1) * M(0, 2)));
    float inv[9]; // inv[c*3+r]
    inv[0 * 3 + 0] = +(M(1, 1) * M(2, 2) - M(2, 1) * M(1, 2)) * ood;
    inv[1 * 3 + 0] = -(M(1, 0) * M(2, 2) - M(2, 0) * M(1, 2)) * ood;
    inv[2 * 3 + 0] = +(M(1, 0) * M(2, 1) - M(2, 0) * M(1, 1)) * ood;
    inv[0 * 3 + 1] = -(M(0, 1) * M(2, 2) - M(2, 1) * M(0, 2)) * ood;
    inv[1 * 3 + 1] = +(M(0, 0) * M(2, 2) - M(2, 0) * M(0, 2)) * ood;
    inv[2 * 3 + 1] = -(M(0, 0) * M(2, 1) - M(2, 0) * M(0, 1)) * ood;
    inv[0 * 3 + 2] = +(M(0, 1) * M(1, 2) - M(1, 1) * M(0, 2)) * ood;
    inv[1 * 3 + 2] = -(M(0, 0) * M(1, 2) - M(1, 0) * M(0, 2)) * ood;
    inv[2 * 3 + 2] = +(M(0, 0) * M(1, 1) - M(1, 0) * M(0, 1)) * ood;
    for(int c = 0; c < 3; c++)for(int r = 0; r < 3; r++)out[c * 3 + r] = inv[r * 3 + c]; // transpose
#undef M
}

Box transform_box(const float m[16], const Box& b){
    Box out = box_empty();
    for(int n = 0; n < 8; n++){
        const float x = (n & 1) ? b.hi[0] : b.lo[0];
        const float y = (n & 2) ? b.hi[1] : b.lo[1];
        const float z = (n & 4) ? b.hi[2] : b.lo[2];
        float p[3];
        for(int r = 0; r < 3; r++)p[r] = (m[0 * 4 + r] * x + m[1 * 4 + r] * y) + (m[2 * 4 + r] * z + m[3 * 4 + r] * 1.0f);
        box_grow(out, p);
    }
    return out;
}


// ---- PT_LAYOUT_FAST: Bvh4Node -> Bvh4QNode -----------------------------------------------------------------------------
// Per cluster and axis: step = 2^e with (extent + margins) / step <= 255 and step >= 2^-20 of the largest coordinate
// (so that a quarter step stays far above the float rounding of the traversal's plane distances), origin a quarter
// step below the cluster's lower corner.  Child planes: lower ones rounded down, upper ones up, each by at least a
// quarter step; the result is checked in float arithmetic (origin + q * step) against the exact box and widened if
// a rounding ever went the wrong way.
void encode_bvh4q(const std::vector<Bvh4Node>& n4, std::vector<Bvh4QNode>& out){
    out.resize(n4.size());
    const size_t n = n4.size();
    unsigned nt = std::thread::hardware_concurrency();
    if(nt == 0)nt = 1;
    if(nt > 16)nt = 16;
    if(n < 65536)nt = 1;
    auto work = [&](size_t lo_i, size_t hi_i){ for(size_t i = lo_i; i < hi_i; i++)encode_node4q(n4[i], out[i]); };     // bvh4q_encode.hpp
    if(nt == 1){ work(0, n); return; }
    std::vector<std::thread> th;
    for(unsigned t = 0; t < nt; t++)th.emplace_back(work, n * t / nt, n * (t + 1) / nt);
    for(auto& t : th)t.join();
}

} // namespace ptb
