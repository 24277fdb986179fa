// bvh_flatten4.hpp -- the BVH4 collapse of one cluster (BVH4::buildBVH4, BVH.hpp:788-1017), shared by the host
// flatten (bvh_build.cpp flatten4) and the device one (scene_pack_gpu.cu): which binary nodes become the four lanes, in
// which slots, and the visit order per ray octant.
#pragma once
#include <cstdint>
#include "bvh_build.hpp"

#if defined(__CUDACC__)
#define PTB_HD __host__ __device__
#else
#define PTB_HD
#endif

namespace ptb {

// A cluster is described by a tiny expression tree over the
// original binary nodes: Leafs of that tree become the cluster's lanes.
struct Expr {
    int lane;               // >= 0: a lane holding binary node `node`
    uint32_t node;
    uint32_t axis;          // split axis (inner expr)
    int l, r;               // children (indices into the per-cluster pool)
};

struct ClusterPlan {
    Expr pool[7];
    int used;
    int root;
    PTB_HD int add_lane(uint32_t node, int slot){ pool[used].node = node; pool[used].lane = slot; pool[used].axis = 0; pool[used].l = pool[used].r = -1; return used++; }
    PTB_HD int add_split(uint32_t axis, int l, int r){ pool[used].node = 0; pool[used].axis = axis; pool[used].l = l; pool[used].r = r; pool[used].lane = -1; return used++; }
};

// Visit order for one octant: at every binary split the near side comes
// first; a negative direction on the split axis makes the right side near
// (this is what the 5x27x8 LUT of BVH.hpp:562-718 enumerates).
PTB_HD inline void cluster_visit(const ClusterPlan& p, int e, unsigned octant, int* out, int& n){
    const Expr& x = p.pool[e];
    if(x.lane >= 0){ out[n++] = x.lane; return; }
    const bool neg = (octant >> x.axis) & 1u;
    cluster_visit(p, neg ? x.r : x.l, octant, out, n);
    cluster_visit(p, neg ? x.l : x.r, octant, out, n);
}

// lane numbers follow BVH4::buildBVH4's slot choices (BVH.hpp:801-1015); ni is an inner node
PTB_HD inline void plan_cluster4(const BuildNode* nodes, uint32_t ni, ClusterPlan& p){
    p.used = 0; p.root = -1;
    const BuildNode& n = nodes[ni];
    const BuildNode& L = nodes[n.left];
    const BuildNode& R = nodes[n.right];
    const bool lleaf = L.count != 0, rleaf = R.count != 0;
    if(lleaf && rleaf){
        const int a = p.add_lane(n.left, 0), c = p.add_lane(n.right, 2);
        p.root = p.add_split(n.axis, a, c);
    } else if(lleaf){
        const BuildNode& RL = nodes[R.left];
        const BuildNode& RR = nodes[R.right];
        if(RL.count != 0 && RR.count != 0){
            const int a = p.add_lane(R.left, 2), c = p.add_lane(R.right, 3);
            const int sub = p.add_split(R.axis, a, c);
            const int l0 = p.add_lane(n.left, 0);
            p.root = p.add_split(n.axis, l0, sub);
        } else if(RL.count != 0){
            const int a = p.add_lane(RR.left, 2), c = p.add_lane(RR.right, 3);
            const int low = p.add_split(RR.axis, a, c);
            const int l1 = p.add_lane(R.left, 1);
            const int sub = p.add_split(R.axis, l1, low);
            const int l0 = p.add_lane(n.left, 0);
            p.root = p.add_split(n.axis, l0, sub);
        } else{
            const int a = p.add_lane(RL.left, 1), c = p.add_lane(RL.right, 2);
            const int low = p.add_split(RL.axis, a, c);
            const int l3 = p.add_lane(R.right, 3);
            const int sub = p.add_split(R.axis, low, l3);
            const int l0 = p.add_lane(n.left, 0);
            p.root = p.add_split(n.axis, l0, sub);
        }
    } else if(rleaf){
        const BuildNode& LL = nodes[L.left];
        const BuildNode& LR = nodes[L.right];
        if(LL.count != 0 && LR.count != 0){
            const int a = p.add_lane(L.left, 0), c = p.add_lane(L.right, 1);
            const int sub = p.add_split(L.axis, a, c);
            const int l2 = p.add_lane(n.right, 2);
            p.root = p.add_split(n.axis, sub, l2);
        } else if(LL.count != 0){
            const int a = p.add_lane(LR.left, 1), c = p.add_lane(LR.right, 2);
            const int low = p.add_split(LR.axis, a, c);
            const int l0 = p.add_lane(L.left, 0);
            const int sub = p.add_split(L.axis, l0, low);
            const int l3 = p.add_lane(n.right, 3);
            p.root = p.add_split(n.axis, sub, l3);
        } else{
            const int a = p.add_lane(LL.left, 0), c = p.add_lane(LL.right, 1);
            const int low = p.add_split(LL.axis, a, c);
            const int l2 = p.add_lane(L.right, 2);
            const int sub = p.add_split(L.axis, low, l2);
            const int l3 = p.add_lane(n.right, 3);
            p.root = p.add_split(n.axis, sub, l3);
        }
    } else{
        const int a = p.add_lane(L.left, 0), b2 = p.add_lane(L.right, 1);
        const int x = p.add_split(L.axis, a, b2);
        const int c = p.add_lane(R.left, 2), d = p.add_lane(R.right, 3);
        const int y = p.add_split(R.axis, c, d);
        p.root = p.add_split(n.axis, x, y);
    }
}

// The cluster without its refs: child boxes (inactive lanes: zero boxes, BVH.hpp:46-48), visit-order bytes, and which binary
// node sits in which lane (lane_node[k], valid where lane_used[k])
PTB_HD inline void fill_cluster4(const BuildNode* nodes, const ClusterPlan& p, Bvh4Node& o, uint32_t lane_node[4], bool lane_used[4]){
    for(int k = 0; k < 4; k++){ lane_used[k] = false; lane_node[k] = 0; }
    for(int e = 0; e < p.used; e++)if(p.pool[e].lane >= 0){ lane_used[p.pool[e].lane] = true; lane_node[p.pool[e].lane] = p.pool[e].node; }
    for(int k = 0; k < 4; k++){
        float lo[3] = { 0.0f, 0.0f, 0.0f }, hi[3] = { 0.0f, 0.0f, 0.0f };
        if(lane_used[k]){ const Box& bx = nodes[lane_node[k]].box; for(int a = 0; a < 3; a++){ lo[a] = bx.lo[a]; hi[a] = bx.hi[a]; } }
        o.xmin[k] = lo[0]; o.xmax[k] = hi[0];
        o.ymin[k] = lo[1]; o.ymax[k] = hi[1];
        o.zmin[k] = lo[2]; o.zmax[k] = hi[2];
        o.ref[k] = REF_EMPTY;
    }
    for(unsigned oct = 0; oct < 8; oct++){
        int seq[4], cnt = 0;
        cluster_visit(p, p.root, oct, seq, cnt);
        bool seen[4] = { false, false, false, false };
        for(int k = 0; k < cnt; k++)seen[seq[k]] = true;
        for(int k = 0; k < 4; k++)if(!seen[k])seq[cnt++] = k;   // absent lanes last; never pushed
        o.order[oct] = (uint8_t)((seq[0] << 6) | (seq[1] << 4) | (seq[2] << 2) | seq[3]);
    }
    o.pad[0] = o.pad[1] = 0;
}

} // namespace ptb
