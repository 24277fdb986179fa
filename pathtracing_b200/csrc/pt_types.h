// pt_types.h — device-resident scene layout shared by the host builder and the
// sm_100a kernels.  All records are 16-byte aligned so that every fetch is a
// 128-bit load (LDG.E.128); node sizes are one (BVH4) or half (BVH2) of a
// 128-byte cache line.  See DESIGN.md "Data layout in HBM".
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define PT_HD __host__ __device__
#else
#define PT_HD
#endif

namespace ptb {

// Child / stack reference encoding (32 bit):
//   bit31 = 0 : inner node; low bits = node index (BVH2 pair node / BVH4 cluster)
//   bit31 = 1 : leaf; low 31 bits = index of the first record (triangle in a
//               BLAS, instance in the TLAS); the leaf ends at the record whose
//               LAST flag is set (no count needed on the stack)
//   0xFFFFFFFF: empty lane of a BVH4 cluster (BVH.hpp:38-43 `count=0, active=0`)
constexpr uint32_t REF_LEAF = 0x80000000u;
constexpr uint32_t REF_EMPTY = 0xFFFFFFFFu;

// BVH2 "child-pair" node, 64 B.  Replaces the two 32 B BVH2_NODE loads at
// idx+1 / node.right of BVH2::Intersect (BVH.hpp:450-459,515-519): both child
// boxes and both child refs arrive in one half cache line.
struct alignas(16) Bvh2Node {
    float lo0[3], hi0[3];   // child 0 (the reference's `index+1`, left) box
    float lo1[3], hi1[3];   // child 1 (the reference's `node.right`) box
    uint32_t ref0, ref1;
    uint32_t axis;          // split axis of this node (BVH_NODE::axis, BVH.hpp:31)
    uint32_t pad;
};
static_assert(sizeof(Bvh2Node) == 64, "Bvh2Node must be 64 B");

// BVH4 cluster, 128 B = one cache line.  Same content as BVH4_CLUSTER
// (BVH.hpp:45-60): SoA child boxes + 4 child descriptors; the 135x8 traversal
// order LUT (BVH.hpp:562-718) is folded at build time into one visit-order
// byte per ray octant, so the kernel needs no lookup tables.
struct alignas(16) Bvh4Node {
    float xmin[4], xmax[4], ymin[4], ymax[4], zmin[4], zmax[4];
    uint32_t ref[4];
    uint8_t order[8];       // per octant (bit0 = dir.x<0, bit1 = y, bit2 = z):
                            // P(a,b,c,d) = a<<6|b<<4|c<<2|d, lane a visited first
    uint32_t pad[2];
};
static_assert(sizeof(Bvh4Node) == 128, "Bvh4Node must be 128 B");

// "Fast" BVH4 cluster, 64 B = two 32 B sectors (pt_scene_set_layout(PT_LAYOUT_FAST)).  The same tree, child
// order and visit-order bytes as Bvh4Node -- only the child boxes are stored as 8-bit offsets from a per-cluster
// origin in power-of-two steps per axis (after Ylitie et al. 2017, "Efficient Incoherent Ray Traversal on GPUs Through
// Compressed Wide BVHs"), rounded outwards by at least a quarter step, so every decoded box CONTAINS the exact one:
// the traversal visits a superset of the nodes the reference visits, in the same order, and the closest hit (exact
// triangle arithmetic, unchanged) is the same.  Why: k_trace is bound by L1 data-pipe wavefronts -- one per 32 B
// sector per lane for incoherent rays (profiles/r01_trace4_v7.md: 92 % of peak) -- and this halves the sectors per
// cluster visit (2 x LDG.256 instead of 4).
//   word[p], p = 0..5 (lo x, lo y, lo z, hi x, hi y, hi z): byte c = child c's plane p in steps from `org`
struct alignas(32) Bvh4QNode {
    float org[3];
    uint8_t ex[3];          // biased exponent of the step per axis: step = 2^(ex - 127)
    uint8_t pad0;
    uint32_t ref[4];
    uint32_t q[6];
    uint8_t order[8];
};
static_assert(sizeof(Bvh4QNode) == 64, "Bvh4QNode must be 64 B");
constexpr int kWidthFast4 = 44;          // template "width" of the kernels that walk Bvh4QNode + Tri64

// BVH8 cluster, 256 B = eight 32 B sectors.  Same content as BVH8_CLUSTER (BVH.hpp:1241-1257) with the child
// boxes stored as two 4-lane halves (half h holds lanes 4h..4h+3: xmin[4] xmax[4] ymin[4] ymax[4] zmin[4]
// zmax[4]) so that a half is three 256-bit loads and eight slab tests never hold 48 box floats at once.
// order[octant]: the visit sequence of BVH8::buildPerm (BVH.hpp:1356-1372), nearest lane in bits 0-2,
// absent lanes last (never pushed).
struct alignas(32) Bvh8Node {
    float box[2][6][4];
    uint32_t ref[8];
    uint32_t order[8];
};
static_assert(sizeof(Bvh8Node) == 256, "Bvh8Node must be 256 B");

// Triangle in leaf order, 48 B: vertex 0 and the two edges exactly as
// glm::intersectRayTriangle forms them (edge = v - v0 is a single correctly
// rounded subtraction, so precomputing it is bit-identical).
struct alignas(16) Tri {
    float v0[3]; uint32_t prim;   // BLAS-wide triangle index in original order
    float e1[3]; uint32_t mesh;   // scene mesh id (material / medium / emission)
    float e2[3]; uint32_t flags;
};
static_assert(sizeof(Tri) == 48, "Tri must be 48 B");
// fast layout: the same record padded to 64 B, so that it is two aligned 256-bit loads (2 wavefronts) instead of three 128-bit ones
struct alignas(32) Tri64 { Tri t; uint32_t pad[4]; };
static_assert(sizeof(Tri64) == 64, "Tri64 must be 64 B");
constexpr uint32_t TRI_LAST = 1u;        // last triangle of its leaf
constexpr uint32_t TRI_ALPHA_PATH = 2u;  // material->HasAlpha(): any-hit goes through
                                         // shape->Intersect (Primitive.cpp:6-14)
constexpr uint32_t TRI_ALPHA_TEX = 4u;   // alpha can be < 1: evaluate the alpha texture

// Instance record for traversal, 96 B (6 x 16 B).
struct alignas(16) Instance {
    float inv[12];       // rows 0..2 of the inverse transform, column-major source:
                         // inv[c*3+r] = invTransform[c][r], c = 0..3 (translation c=3)
    uint32_t root;       // BLAS root ref (already offset into the global arrays)
    uint32_t flags;      // INST_*
    uint32_t id;         // instance id in pt_scene_add_instance order
    uint32_t blas;
    float blo[3];        // BLAS root box (BVH2::Intersect tests it first, BVH.hpp:504)
    float bhi0;
    float bhi12[2];
    uint32_t pad[2];
};
static_assert(sizeof(Instance) == 96, "Instance must be 96 B");
constexpr uint32_t INST_LAST = 1u;
constexpr uint32_t INST_XF = 2u;         // TransformedPrimitive (Primitive.cpp:41-66)
constexpr uint32_t INST_SPHERE = 4u;     // not a BLAS: GeometricPrimitive(SphereShape), constants packed in the record (scene_build.cpp pack_shape)
constexpr uint32_t INST_QUAD = 8u;       // GeometricPrimitive(QuadShape)
// Wrapper records.  A top-level primitive is a chain of 1-3 consecutive records
//   [INST_ANIM]? [INST_WRAP | INST_XF]? (BLAS instance | sphere | quad)
// whose head carries INST_LAST, the primitive's id and, in bits 8-9 (INST_EXTRA_*), how many records
// follow it; the next leaf item is head + 1 + extra.
constexpr uint32_t INST_ANIM = 16u;      // AnimatedPrimitive (Primitive.cpp:77-96): inv[0..2] direction, inv[3..4] timeBounds; the child follows
constexpr uint32_t INST_WRAP = 32u;      // TransformedPrimitive over an analytic shape: inv[] = inverse transform; the shape record follows
constexpr uint32_t INST_ALPHA = 64u;     // shape record: its material can reject a hit (Material::Alpha, Primitive.cpp:6-26); `root` = material id
constexpr uint32_t INST_WBOX = 128u;     // BVH4 / BVH8, BLAS instance: blo / bhi hold the instance's WORLD box (slightly enlarged), tested before the ray is transformed
constexpr uint32_t INST_EXTRA_SHIFT = 8u;
constexpr uint32_t INST_EXTRA_MASK = 3u;

// Everything a traversal kernel needs, passed by value as a kernel argument.
struct DevScene {
    const Bvh2Node* n2;      // BVH2 nodes: all BLASes, then the TLAS
    const Bvh4Node* n4;      // BVH4 clusters: all BLASes, then the TLAS
    const Bvh8Node* n8;      // BVH8 clusters: all BLASes, then the TLAS
    const Tri* tris;         // all BLAS triangles in leaf order
    const Instance* inst;    // instances in TLAS leaf order
    const Bvh4QNode* n4q;    // fast layout (width == 4 only): quantised copy of n4, same indices
    const Tri64* tris64;     // fast layout: tris padded to 64 B, same indices
    uint32_t tlas_root;      // ref into n2/n4 (or a leaf ref into inst)
    int width;               // 2 | 4 | 8
    int fast;                // 1: the kernels walk n4q / tris64 (kWidthFast4)
    float tlo[3], thi[3];    // TLAS root box (BVH2::Intersect tests it first)
    uint32_t single_root;    // scene = one untransformed instance: its BLAS root ref (traversal starts
                             // inside it, no TLAS leaf / instance record round trip); else REF_EMPTY
    uint32_t single_id;      // that instance's id
};

} // namespace ptb
