// csrc/scene_dev.cuh -- the scene as the kernels see it: plain device pointers, passed by value.
//
// HBM layout (everything 16-byte vectors, read through the read-only path):
//   nodes / tris, lnodes / ltris : the two BVHs of bvh.h (whole scene; emissive triangles only)
//   nodes4 [N4][7] float4 : the scene BVH collapsed to 4-wide nodes (bvh.h) for k_extend / k_shadow
//   nodes8 [N8][5] float4, tris8 [P][3] float4 : the same tree as a compressed 8-wide BVH (80 B nodes, byte boxes) with its
//                          own node-grouped copy of the triangles; the trace kernels use it for large scenes (use8)
//   pverts [P][3] float4 : prim-order vertices; v0.w = material id bits, v1.w = light index bits (-1 = not a light)
//   pnorms [P][3] float4 : prim-order vertex normals (shading normal = normalised barycentric blend,
//                          main.cpp:273-274)
//   mats   [M][2] float4 : (kd.rgb, ns), (ks.rgb, 0)              (main.cpp:292,313)
//   lights [L][5] float4 : v0|prim bits, v1|class bits, v2|area, ng|Lsum (R+G+B), Le|0 -- in the
//                          reference's light order (Mylight.cpp:169-173); FP32 copy for the renderer
//   lsoa   [Lpad/32][4][32] float4 : the same lights as four planes per block of 32 (v0, v1, v2 | cull threshold, -2 area ng | 2 Lsum) for
//                          the O(L) spherical-triangle weight loop (lights.cuh); Lpad = L rounded up to 64
//   lng64 [L][3], larea64 [L], crad64 [C][3] : FP64 originals for the parity entry points
//   ltree [2L-1][3] float4, ltree_leaf [L] : light hierarchy of MCPT_LIGHT_TREE (light_tree.h)
//   llist, lrange [nlists], prim_list [P] : per-primitive survivor lists of the O(L) sweeps (light_lists.h)
//   class_begin [C+1], class_cdf [C], tri_cdf [L] : the two discrete distributions of
//                          select_a_point_from_lights (Mylight.cpp:112-135) as normalised cumulative tables
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace mcpt {

struct DScene {
    const float4* nodes;
    const float4* nodes4;  // the scene BVH collapsed to 4-wide nodes (bvh.h), what the trace kernels traverse
    const float4* tris;
    const float4* nodes8;  // compressed 8-wide BVH of the scene (bvh.h), traversed instead of nodes4 when use8 != 0
    const float4* tris8;   // the triangles once more, grouped by 8-wide node
    const float4* lnodes;
    const float4* ltris;
    const float4* pverts;
    const float4* pnorms;
    const float4* mats;
    const float4* lights;
    const float4* lsoa;
    const double* lng64;
    const double* larea64;
    const double* crad64;
    const int32_t* class_begin;
    const float* class_cdf;
    const float* tri_cdf;
    const float4* ltree;        // [2L-1][3] light hierarchy nodes (light_tree.h), node 0 = root
    const int32_t* ltree_leaf;  // [L] leaf node of each light
    const uint32_t* llist;      // per-primitive survivor lists of the light sweeps (light_lists.h): lsoa offsets
    const int2* lrange;         // [nlists] first entry | padded count (multiple of 32)
    const int32_t* prim_list;   // [P] list of each primitive, -1 for emissive ones
    int32_t nlists;             // 0: no lists, the sweeps evaluate the whole table
    int32_t list_max;           // longest padded list (entries)
    int32_t P, L, C, M, Lpad;
    int32_t use8;  // trace kernels walk nodes8 / tris8 (large scenes) instead of nodes4 / tris
    int32_t refill4, refill8;  // ray-stream tracers: idle lanes fetch new rays when fewer than this many lanes of the warp are traversing
    float tmin;  // FP32 leaf test: reject t < tmin (scene-scale aware stand-in for the reference's 1e-8)
};

struct DCamera {
    float eye[3];
    float U[3], V[3], N[3];  // T = [U V N], main.cpp:549-553
    float pixellen, wlen;    // main.cpp:547, |w|
    float ci, cj;            // (H-1)/2, (W-1)/2 (359.5 / 639.5 at 720x1280, main.cpp:563)
    int32_t W, H;
};

}  // namespace mcpt
