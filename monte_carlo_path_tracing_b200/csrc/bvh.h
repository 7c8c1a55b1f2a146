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
    uint32_t num_leaves = 0, max_depth = 0, max_depth4 = 0;
    double sah_cost = 0;
    float lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};  // bounds of everything inside
};

constexpr int kBvhMaxLeaf = 4;
constexpr int32_t kBvhNoChild = INT32_MIN;  // child word of an unused nodes4 slot (its box is inverted, too)
constexpr int kBvhStackDepth = 64;  // traversal stack entries; the builder falls back to median splits to stay below

// positions: [P][3][3] floats; prims: the subset to build over (ids kept as triangle payload).
Bvh build_bvh(const float* positions, const int32_t* prims, size_t count);

}  // namespace mcpt
