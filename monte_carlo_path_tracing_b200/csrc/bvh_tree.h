// csrc/bvh_tree.h -- the builders' intermediate binary tree (host SAH builder, GPU LBVH builder) and its emission
// into the traversal layouts of bvh.h.  Internal to the library.
#pragma once
#include <cfloat>
#include <cstdint>
#include <vector>
#include "bvh.h"

namespace mcpt {

struct BvhBox {
    float lo[3], hi[3];
    void reset() { for (int a = 0; a < 3; a++) { lo[a] = FLT_MAX; hi[a] = -FLT_MAX; } }
    void grow(const float* p) { for (int a = 0; a < 3; a++) { lo[a] = p[a] < lo[a] ? p[a] : lo[a]; hi[a] = p[a] > hi[a] ? p[a] : hi[a]; } }
    void grow(const BvhBox& b) { for (int a = 0; a < 3; a++) { lo[a] = b.lo[a] < lo[a] ? b.lo[a] : lo[a]; hi[a] = b.hi[a] > hi[a] ? b.hi[a] : hi[a]; } }
    float half_area() const
    {
        float d[3] = {hi[0] - lo[0], hi[1] - lo[1], hi[2] - lo[2]};
        if (d[0] < 0 || d[1] < 0 || d[2] < 0) return 0.f;
        return d[0] * d[1] + d[1] * d[2] + d[2] * d[0];
    }
};

// left < 0: leaf over order[first .. first + count); else inner node with children tmp[left], tmp[right]
struct BvhTmpNode {
    BvhBox box;
    int32_t left = -1, right = -1;
    uint32_t first = 0, count = 0;
};
static_assert(sizeof(BvhTmpNode) == 40, "BvhTmpNode is copied from the device as plain bytes");

struct BvhTree {
    std::vector<BvhTmpNode> tmp;
    std::vector<int32_t> order;  // prim ids in leaf order
    int32_t root = 0;
};

bool build_tree_host(const float* positions, const int32_t* prims, size_t count, BvhTree& T);
// Morton-code LBVH on the device (bvh_gpu.cu).  false = not usable (count < 2): use the host builder.
bool build_tree_gpu(const float* positions, size_t num_prims, const int32_t* prims, size_t count, int device, BvhTree& T, float* device_ms);
void emit_bvh(const float* positions, const BvhTree& T, Bvh& out);

}  // namespace mcpt
