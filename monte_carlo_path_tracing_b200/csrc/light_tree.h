// csrc/light_tree.h -- light hierarchy over the emissive triangles.
//
// The reference selects a light at every vertex by looping over ALL light triangles (Mylight.cpp:169-259 /
// :329-421) and drawing one with probability sA_k L_k / sum.  With MCPT_LIGHT_TREE the renderer instead walks a
// binary tree over the lights: at every inner node it estimates how much each child can contribute to the
// vertex (power, distance, the cone that holds the children's normals, the vertex's horizon -- the importance
// of Conty & Kulla, "Importance Sampling of Many Lights with Adaptive Tree Splitting", 2018) and descends into
// one child at random.  O(log L) per vertex instead of O(L); the chosen triangle is then sampled by the same
// Arvo spherical-triangle sampler (Mylight.cpp:285-317), so pdf(direction) = P(triangle | vertex) / sA.
// A different (still unbiased) estimator than the reference's: never the default.
//
// Device layout: three float4 per node, node 0 = root, leaves are nodes too (2L - 1 nodes):
//   t0 = centre of the node's bounding box | radius of the sphere around that box (holds every light vertex below)
//   t1 = cone axis | cos(theta_o)       every light normal below lies within theta_o of the axis
//   t2 = power (sum of area * (R+G+B)) | left | right | parent   (int bits; left < 0: leaf of light ~left; root's parent = -1)
// plus leaf_of_light[L] for the bottom-up probability query (MIS: pdf of a BRDF-sampled direction that hits a light).
#pragma once
#include <cstdint>
#include <vector>
#include "bvh.h"

namespace mcpt {

struct LightTree {
    std::vector<f4> nodes;              // 3 per node
    std::vector<int32_t> leaf_of_light; // [L]
    uint32_t num_nodes = 0, max_depth = 0;
};

// positions [P][3][3]; light_prim [L]; normal [P][3] geometric normals (emitting side); area [L]; lsum [L] = R+G+B.
LightTree build_light_tree(const float* positions, const int32_t* light_prim, const double* geom_normals, const double* area,
                           const double* lsum, uint32_t L);

// Structural check for tests: every light in exactly one leaf, children's boxes / normals inside the parent's
// sphere / cone, powers add up, parent links consistent.  Returns the number of violations.
uint32_t check_light_tree(const LightTree& T, const float* positions, const int32_t* light_prim, const double* geom_normals, uint32_t L);

}  // namespace mcpt
