// csrc/bvh.h -- SAH BVH in the GPU traversal layout (host builder + device consumer).
//
// Replaces the reference's uniform grid (Myobj::cal_scene_boundingbox + meshing, Myobj.cpp:78-162,
// storage Myobj.h:76-81).  Semantics the traversal must deliver are those of the exhaustive search
// the grid approximates (Myobj.cpp:194-213): minimum t, strict <, ascending prim on ties, with the
// origin triangle excluded by id (Myobj.cpp:417).
//
// Layout (all 16-byte vectors, SoA between node array and triangle array):
//   node (64 B): n0 = (c0.lo.x, c0.hi.x, c0.lo.y, c0.hi.y)
//                n1 = (c1.lo.x, c1.hi.x, c1.lo.y, c1.hi.y)
//                n2 = (c0.lo.z, c0.hi.z, c1.lo.z, c1.hi.z)
//                n3 = (child0, child1, -, -) as int bits; child >= 0 inner node index,
//                     child < 0 leaf: ~child = (first_tri << 3) | (count - 1), count in 1..8
//   triangle (48 B): v0.xyz | prim id bits, v1.xyz | 0, v2.xyz | 0  in leaf order
// An empty child has an inverted box (lo = +inf, hi = -inf) and is never entered.
//
// nodes4: the same tree collapsed to 4-wide nodes for the renderer's stream traversal (traverse.cuh,
// trace_stream): the child with the largest surface area is replaced by its own two children until the node has
// four (or only leaves are left).  112 B per node, seven 16-byte vectors, child k in component k:
//   q0 = lo.x[4], q1 = hi.x[4], q2 = lo.y[4], q3 = hi.y[4], q4 = lo.z[4], q5 = hi.z[4], q6 = child[4] (int bits,
//   same inner / leaf encoding as above; unused slots hold kBvhNoChild and an inverted box).  Half the levels of the binary tree
//   means half as many dependent node fetches per ray, which is what bounds incoherent rays in a deep BVH.
//
// nodes8 / tris8: the same tree once more as a COMPRESSED 8-WIDE BVH for large scenes (layout after Ylitie, Karras, Laine,
// "Efficient Incoherent Ray Traversal on GPUs Through Compressed Wide BVHs", HPG 2017; restated here from the paper):
// 80 B per node, five 16-byte vectors, against 2 x 112 B for the two levels of 4-wide nodes it replaces --
//   v0 = origin p (3 floats) | ex, ey, ez (IEEE exponent bytes of the per-axis quantisation step), imask (slots that hold an inner node)
//   v1 = child_base (index of the first inner child in nodes8), tri_base (first triangle of this node in tris8),
//        meta[8]: per slot  0 = empty,  inner: 0b001_11sss (low 5 bits = 24 + slot),  leaf: unary triangle count in the top
//        3 bits (0b001, 0b011, 0b111 = 1, 2, 3 triangles) | offset of its first triangle from tri_base in the low 5 (0..23)
//   v2 = qlo.x[8] qlo.y[8],  v3 = qlo.z[8] qhi.x[8],  v4 = qhi.y[8] qhi.z[8]   child boxes as bytes: p + q * 2^e, rounded outward
// Inner children are stored consecutively from child_base in slot order.  A child sits in the slot whose sign pattern
// (bit a set = towards -axis a) best matches its offset from the node centre, so a ray with direction-sign octant `oct`
// meets the slots in roughly front-to-back order when it takes them by descending (slot ^ oct) -- no distance sort.
// tris8 holds every triangle once more, grouped by node8 (a node's <= 24 leaf triangles are contiguous), 48 B each.
#pragma once
#include <cstdint>
#include <vector>

namespace mcpt {

struct f4 { float x, y, z, w; };

struct BvhNode { f4 n0, n1, n2; int32_t c0, c1, pad0, pad1; };
static_assert(sizeof(BvhNode) == 64, "node must be 64 bytes");

struct Bvh {
    std::vector<BvhNode> nodes;  // nodes[0] is the root (always an inner node)
    std::vector<f4> nodes4;      // 7 per 4-wide node; nodes4[0..6] is the root
    std::vector<f4> tris;        // 3 per triangle, leaf order
    std::vector<f4> nodes8;      // 5 per compressed 8-wide node; nodes8[0..4] is the root
    std::vector<f4> tris8;       // 3 per triangle, grouped by 8-wide node
    uint32_t num_leaves = 0, max_depth = 0, max_depth4 = 0, max_depth8 = 0;
    double sah_cost = 0;
    float lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};  // bounds of everything inside
};

#ifndef MCPT_MAX_LEAF
#define MCPT_MAX_LEAF 4
#endif
constexpr int kBvhMaxLeaf = MCPT_MAX_LEAF;
constexpr int32_t kBvhNoChild = INT32_MIN;  // child word of an unused nodes4 slot (its box is inverted, too)
constexpr int kBvhStackDepth = 64;  // traversal stack entries; the builder falls back to median splits to stay below
// Refill thresholds of the ray-stream tracers (traverse.cuh): measured per layout, r2y -- the short traversals of the 4-wide
// kernels want a low one (veach-mis trace side 25.2 -> 24.0 ms at 14 instead of 20, interior flat), the compressed 8-wide kernel
// wants 20 on cornell-box and 29 on the 1 M-triangle soup (+6 %): the 8-wide value is settled per scene with the layout.
#ifndef MCPT_REFILL8_HI
#define MCPT_REFILL8_HI 28
#endif
constexpr int kRefill4 = 14, kRefill8 = 24, kRefill8Lo = 20, kRefill8Hi = MCPT_REFILL8_HI;
constexpr int kBvh8StackDepth = 32; // node-group entries of the 8-wide traversal: at most one pending group per level
static_assert(kBvhMaxLeaf <= 6, "emit_wide8 splits a leaf into two slots of at most 3 triangles");

// positions: [P][3][3] floats; prims: the subset to build over (ids kept as triangle payload).
Bvh build_bvh(const float* positions, const int32_t* prims, size_t count);

}  // namespace mcpt
