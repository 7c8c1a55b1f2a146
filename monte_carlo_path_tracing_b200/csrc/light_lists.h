// csrc/light_lists.h -- per-primitive survivor lists for the O(L) spherical-triangle light selection.
//
// The reference evaluates every light triangle at every vertex (Mylight.cpp:169-259 / :329-421) and drops the ones that
// fail one of two culls BEFORE any trigonometry: the vertex is behind the light (`nl . (x1 - v0) < eps`, :178-185), or
// the whole light is below the vertex's tangent plane (`n . (v_k - x1) < eps` for k = 0..2, :187-197).  Both tests are
// (bi)linear in the position on the primitive the vertex lies on:
//   * nl . (x - v0) with x = sum b_j p_j is linear in the barycentrics, so its maximum over the primitive is at one of
//     its three corners;
//   * N(x) . (v_k - x) with N = normalize(sum b_i n_i) has the sign of sum_ij b_i b_j n_i . (v_k - p_j), a convex
//     combination of the nine corner terms, so it cannot exceed their maximum.
// A light whose corner maxima are negative (by more than an FP32 safety margin) is culled at EVERY vertex of that
// primitive: it contributes an exact zero to weights_sum and can never be drawn.  At scene creation the host lists, for
// every non-emissive primitive, the lights that are not so excluded (about half of the 1282 on veach-mis: a sphere of
// light triangles shows any point its front half only).  Primitives with identical survivor sets share a list.  The sweep
// kernels (lights.cuh: warp_light_partial_list) then evaluate the list instead of the table -- same weights, same
// probabilities, half the evaluations.
//
// Device layout: llist = uint32 offsets into lsoa (lsoa_index(l), ascending l), every list padded to a multiple of 32
// entries with the offset of a padding light (weight 0); lrange[list] = (first entry, padded count); prim_list[prim] =
// list id, -1 for emissive primitives (never sampled from).
#pragma once
#include <cstdint>
#include <vector>

namespace mcpt {

// Layout of the SoA light table `lsoa` the sweeps read (lights.cuh, filled in api.cu): blocks of 32 lights, kLsoaPlanes
// float4 planes per block: v0 | e1 = v1 - v0 | e2 = v2 - v0 | -2 area ng, 2 Lsum | e1.e1, e2.e2, e1.e2, cull threshold.
constexpr int kLsoaPlanes = 5;
constexpr int kLsoaBlock = 32 * kLsoaPlanes;  // float4 per block of 32 lights

struct LightLists {
    std::vector<uint32_t> entries;    // lsoa offsets, lists back to back
    std::vector<int32_t> range;       // [nlists][2] first entry | padded count
    std::vector<int32_t> prim_list;   // [P]
    uint32_t nlists = 0;
    uint32_t max_len = 0;             // longest padded list
    double mean_fraction = 0;         // mean unpadded list length / L over the listed primitives
    double build_ms = 0;
};

// positions, normals [P][3][3]; light_of_prim [P] (-1 = not a light); light_prim [L]; geom_normals [P][3] (FP64, emitting
// side); light_area [L]; Lpad > L (lsoa has at least one padding light).  Returns an empty structure (nlists = 0) when the
// lists would not pay: few lights (the per-lane loop of k_lightsel handles <= 32), too many primitive x light pairs to
// test on the host, or more distinct lists than the sort of k_lsel_* is sized for.
LightLists build_light_lists(const float* positions, const float* normals, const int32_t* light_of_prim, uint32_t P,
                             const int32_t* light_prim, const double* geom_normals, const double* light_area, uint32_t L, uint32_t Lpad);

}  // namespace mcpt
