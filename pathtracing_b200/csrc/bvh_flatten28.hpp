// bvh_flatten28.hpp -- one BVH2 child-pair node / one BVH8 cluster from the binary tree, shared by the host flatten
// (bvh_build.cpp flatten2 / flatten8) and the device one (scene_pack_gpu.cu), like bvh_flatten4.hpp for BVH4.
#pragma once
#include <cstdint>
#include <cstring>
#include "bvh_build.hpp"
#include "bvh_flatten4.hpp"     // PTB_HD

namespace ptb {

// AABB::Area (AABB.hpp:126-129): xy + yz + zx of the extent (no fused multiply-add on either side)
PTB_HD inline float box_half_area_hd(const Box& b){
    const float ex = b.hi[0] - b.lo[0], ey = b.hi[1] - b.lo[1], ez = b.hi[2] - b.lo[2];
    return ex * ey + ey * ez + ez * ex;
}

// BVH2: node k holds the boxes + refs of its two children (what BVH2::Intersect reads from nodes[idx+1] and nodes[right])
PTB_HD inline void fill_cluster2(const BuildNode* nodes, uint32_t ni, Bvh2Node& o, uint32_t lane_node[2]){
    const BuildNode& n = nodes[ni];
    const BuildNode& l = nodes[n.left];
    const BuildNode& r = nodes[n.right];
    for(int k = 0; k < 3; k++){
        o.lo0[k] = l.box.lo[k]; o.hi0[k] = l.box.hi[k];
        o.lo1[k] = r.box.lo[k]; o.hi1[k] = r.box.hi[k];
    }
    o.ref0 = REF_EMPTY; o.ref1 = REF_EMPTY; o.axis = n.axis; o.pad = 0;
    lane_node[0] = n.left; lane_node[1] = n.right;
}

// visit order of one octant over the binary nodes between the cluster's root and its lanes (BVH8::buildPerm, BVH.hpp:1356-1372)
PTB_HD inline void perm8_hd(const BuildNode* nodes, const uint32_t* lanes, int end, uint32_t root, unsigned octant, int* out, int& n){
    uint32_t st[16];
    int sp = 0;
    st[sp++] = root;
    while(sp > 0){
        const uint32_t cur = st[--sp];
        int li = -1;
        for(int i = 0; i < end; i++)if(lanes[i] == cur){ li = i; break; }
        if(li >= 0){ out[n++] = li; continue; }
        const BuildNode& x = nodes[cur];
        const bool neg = (octant >> x.axis) & 1u;
        st[sp++] = neg ? x.left : x.right;      // visited second
        st[sp++] = neg ? x.right : x.left;      // visited first
    }
}

// BVH8 collapse (BVH8::buildBVH8, BVH.hpp:1312-1354): starting from the node itself, the open inner node with the largest
// AABB::Area (ties: the later slot, `>=`) is replaced by its left child and its right child is appended, until eight lanes
// exist or every lane is a leaf.  Returns the number of lanes; refs are left REF_EMPTY for the caller.
PTB_HD inline int fill_cluster8(const BuildNode* nodes, uint32_t ni, Bvh8Node& o, uint32_t lanes[8]){
    for(int i = 0; i < 8; i++)lanes[i] = 0;
    lanes[0] = ni;
    int end = 1;
    while(end < 8){
        int k = -1;
        float best = -1.0f;
        for(int i = 0; i < end; i++){
            const BuildNode& c = nodes[lanes[i]];
            if(c.count == 0){ const float a = box_half_area_hd(c.box); if(a >= best){ k = i; best = a; } }
        }
        if(k < 0)break;
        const BuildNode& c = nodes[lanes[k]];
        lanes[end++] = c.right;
        lanes[k] = c.left;
    }
    memset(&o, 0, sizeof o);                                            // absent lanes: zero boxes (BVH.hpp:1242-1244)
    for(int i = 0; i < end; i++){
        const Box& bx = nodes[lanes[i]].box;
        const int h = i >> 2, l = i & 3;
        o.box[h][0][l] = bx.lo[0]; o.box[h][1][l] = bx.hi[0];
        o.box[h][2][l] = bx.lo[1]; o.box[h][3][l] = bx.hi[1];
        o.box[h][4][l] = bx.lo[2]; o.box[h][5][l] = bx.hi[2];
    }
    for(int i = 0; i < 8; i++)o.ref[i] = REF_EMPTY;
    for(unsigned oct = 0; oct < 8; oct++){
        int seq[8], cnt = 0;
        perm8_hd(nodes, lanes, end, ni, oct, seq, cnt);
        bool seen[8] = { false, false, false, false, false, false, false, false };
        for(int k = 0; k < cnt; k++)seen[seq[k]] = true;
        for(int k = 0; k < 8; k++)if(!seen[k])seq[cnt++] = k;
        uint32_t code = 0;
        for(int k = 0; k < 8; k++)code |= (uint32_t)seq[k] << (3 * k);
        o.order[oct] = code;
    }
    return end;
}

} // namespace ptb
