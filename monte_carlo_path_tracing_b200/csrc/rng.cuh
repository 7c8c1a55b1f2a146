// csrc/rng.cuh -- counter-based per-(pixel, sample, path-node) RNG (Philox4x32-10).
//
// Replaces the reference's clock-seeded std::default_random_engine built inside every sampling
// call (main.cpp:323-326, BRDF.cpp:38-39, Mylight.cpp:107-108,268-269,432-433).
// counter = (pixel, global sample index, node code lo, node code hi | call#), key = seed.
// The node code identifies the vertex inside the path (tree): 1 for the primary vertex, 2n for the
// light-sample child and 2n+1 for the BRDF-sample child of node n (shade_with_mis recurses twice,
// main.cpp:464,491); non-branching estimators only ever take the 2n+1 edge.  Results therefore do
// not depend on batch size, queue order or the number of GPUs the spp range is sharded over.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace mcpt {

struct Philox {
    uint32_t c[4];
    uint32_t k[2];
};

__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key)
{
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; r++) {
        uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
        uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += W0;
        key.y += W1;
    }
    return ctr;
}

// 24-bit uniforms in [0,1)
__device__ __forceinline__ float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }

struct Rng4 { float u[4]; };

// `call` selects one of up to 4 independent blocks of 4 uniforms for the same vertex
__device__ __forceinline__ Rng4 rng_draw4(uint64_t seed, uint32_t pixel, uint32_t sample, uint64_t node, uint32_t call)
{
    uint4 ctr = make_uint4(pixel, sample, (uint32_t)node, (uint32_t)(node >> 32) * 4u + call);
    // nodes deeper than 62 levels alias in the top word; with RR p=0.6 that is never reached
    uint4 r = philox4x32_10(ctr, make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
    Rng4 o;
    o.u[0] = u01(r.x); o.u[1] = u01(r.y); o.u[2] = u01(r.z); o.u[3] = u01(r.w);
    return o;
}

}  // namespace mcpt
