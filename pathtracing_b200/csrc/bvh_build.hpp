// bvh_build.hpp — host-side binned-SAH BVH2 build and its flattening into the
// GPU layouts of pt_types.h.  north_star keeps the build on the host; this is a
// from-scratch restatement of the reference builder's *decisions*
// (BVHBase::BuildBaseThreaded, BVH.hpp:290-390: 8/16/32 bins by span, leaf size
// 2, split rejected when bestCost >= area*span, partition on centroid <=
// bestPos) so that the resulting tree topology — and therefore the order in
// which candidate triangles are tested — is the reference's own.
#pragma once
#include <cstdint>
#include <vector>
#include "pt_types.h"

namespace ptb {

struct Box {
    float lo[3], hi[3];
};

Box box_empty();
void box_grow(Box& b, const Box& o);
void box_grow(Box& b, const float p[3]);
float box_half_area(const Box& b);   // AABB::Area (AABB.hpp:126-129): xy+yz+zx of the extent

// Binary build node (one per BVH_NODE of BVH.hpp:26-36).
struct BuildNode {
    Box box;
    uint32_t left = 0, right = 0;  // children (inner) | right = first item (leaf)
    uint32_t count = 0;            // 0 = inner, else number of items in the leaf
    uint32_t axis = 0;
};

struct BinaryBvh {
    std::vector<BuildNode> nodes;  // nodes[0] = root; empty if there were no items
    std::vector<uint32_t> order;   // item permutation: leaf order -> original index
};

// Build over `n` item boxes.  Threads are forked for spans > 256 Ki items as
// the reference does (BVH.hpp:374-380); topology does not depend on timing.
BinaryBvh build_binary_bvh(const Box* boxes, uint32_t n, uint32_t leaf_size = 2);

// Flattened, GPU-ready hierarchy.  `root` is a ref (pt_types.h encoding)
// relative to this hierarchy's own arrays (offset it when packing several
// BLASes into one device array).
struct FlatBvh {
    int width = 0;                 // 2, 4 or 8
    std::vector<Bvh2Node> n2;
    std::vector<Bvh4Node> n4;
    std::vector<Bvh8Node> n8;
    uint32_t root = REF_EMPTY;
    Box bounds = box_empty();
    std::vector<uint32_t> order;   // leaf order -> original item index
    std::vector<uint32_t> leaf_last; // per item in leaf order: 1 if last of its leaf
    // transient, used while flattening on several threads (bvh_build.cpp): where leaf marks go, and the top pass's
    // list of subtrees handed to worker threads
    uint32_t* last_out = nullptr;
    struct FlattenDefer* defer = nullptr;
};

// threads: 0 = pick (one thread under 128 Ki items, else up to 16); the arrays do not depend on it
FlatBvh flatten_bvh(const BinaryBvh& b, int width, unsigned threads = 0);

// glm::inverse(mat4) restated (glm/detail/func_matrix.inl:347-404), column-major.
void mat4_inverse(const float m[16], float out[16]);
// glm::transpose(glm::inverse(glm::mat3(m))) (func_matrix.inl:322-344), column-major 3x3.
void mat3_normal_matrix(const float m[16], float out[9]);
// TransformedPrimitive::BoundingBox (Primitive.cpp:32-40)
Box transform_box(const float m[16], const Box& b);

} // namespace ptb
