// csrc/traverse.cuh -- ray/triangle tests and BVH traversal (closest hit, any hit).
//
// Replaces Myobj::intersect_with_triangle (Myobj.cpp:165-192), closet_ray_intersect (:334-474) and
// closet_ray_intersect_light_triangle (:476-622).  Result semantics = the exhaustive search the grid
// approximates (Myobj.cpp:194-213): minimum t, strict <, ascending prim on ties, origin triangle
// excluded by id (:417), a hit needs |detA| >= 1e-8, beta,gamma >= 0, beta+gamma <= 1, t >= 1e-8.
//
// Two leaf tests:
//   EXACT : FP64, the reference's Cramer's-rule expression tree evaluated with round-to-nearest
//           intrinsics so nothing is contracted into FMAs -> beta, gamma, t and therefore the winning
//           prim are bit-identical to the g++ x86-64 build of the reference.
//   FAST  : FP32 Moeller-Trumbore, what the renderer runs by default.
// Traversal stack: per-thread, first kStackSm entries in shared memory ([entry][thread] so a warp's
// pushes hit 32 different banks), deeper entries spill to a local array.
#pragma once
#include "bvh.h"
#include "device_math.cuh"
#include "scene_dev.cuh"
#include <math_constants.h>

namespace mcpt {

#ifndef MCPT_STACK_SM
#define MCPT_STACK_SM 8
#endif
constexpr int kStackSm = MCPT_STACK_SM;
constexpr int kStackLocal = kBvhStackDepth - kStackSm;

struct Hit { int prim; float b, g, t; };       // a2: intersec_result as one float4 (prim = -1: miss)
struct HitD { int prim; double b, g, t; };

// ---- exact FP64 leaf test -----------------------------------------------------------------------
__device__ __forceinline__ double rn_dot(d3 a, d3 b)  // vec.cpp:73-81: ((0 + a0 b0) + a1 b1) + a2 b2
{
    return __dadd_rn(__dadd_rn(__dmul_rn(a.x, b.x), __dmul_rn(a.y, b.y)), __dmul_rn(a.z, b.z));
}
__device__ __forceinline__ d3 rn_cross(d3 a, d3 b)    // vec.cpp:67-70
{
    return mk<double>(__dsub_rn(__dmul_rn(a.y, b.z), __dmul_rn(a.z, b.y)), __dsub_rn(__dmul_rn(a.z, b.x), __dmul_rn(a.x, b.z)),
                      __dsub_rn(__dmul_rn(a.x, b.y), __dmul_rn(a.y, b.x)));
}
__device__ __forceinline__ d3 rn_sub(d3 a, d3 b) { return mk<double>(__dsub_rn(a.x, b.x), __dsub_rn(a.y, b.y), __dsub_rn(a.z, b.z)); }
__device__ __forceinline__ double rn_det(d3 a, d3 b, d3 c) { return rn_dot(rn_cross(a, b), c); }  // vec.cpp:84-87

__device__ __forceinline__ bool tri_test_exact(d3 ro, d3 rd, d3 a, d3 b, d3 c, double* beta, double* gamma, double* t)
{
    d3 ab = rn_sub(a, b), ac = rn_sub(a, c), ao = rn_sub(a, ro);
    double detA = rn_det(ab, ac, rd);                    // Myobj.cpp:172
    if (fabs(detA) < 1e-8) return false;                 // :175
    double be = __ddiv_rn(rn_det(ao, ac, rd), detA);     // :180
    double ga = __ddiv_rn(rn_det(ab, ao, rd), detA);     // :181
    double tt = __ddiv_rn(rn_det(ab, ac, ao), detA);     // :182
    if (be < 0 || ga < 0 || __dadd_rn(be, ga) > 1 || tt < 0 || fabs(tt) < 1e-8) return false;  // :185
    *beta = be; *gamma = ga; *t = tt;
    return true;
}

// ---- fast FP32 leaf test --------------------------------------------------------------------------
__device__ __forceinline__ bool tri_test_fast(f3 ro, f3 rd, f3 a, f3 b, f3 c, float tmin, float* beta, float* gamma, float* t)
{
    f3 e1 = b - a, e2 = c - a;
    f3 pv = cross(rd, e2);
    float det = dot(e1, pv);
    if (fabsf(det) < 1e-8f) return false;
    float inv = 1.0f / det;
    f3 tv = ro - a;
    float u = dot(tv, pv) * inv;
    f3 qv = cross(tv, e1);
    float v = dot(rd, qv) * inv;
    float tt = dot(e2, qv) * inv;
    if (u < 0.f || v < 0.f || u + v > 1.f || tt < tmin) return false;
    *beta = u; *gamma = v; *t = tt;
    return true;
}

// ---- traversal stack --------------------------------------------------------------------------------
struct TStack {
    int* sm;      // &smem[threadIdx] ; entry k at sm[k * stride]
    int stride;
    int sp;
    int local[kStackLocal];
    __device__ __forceinline__ void init(int* base, int stride_) { sm = base; stride = stride_; sp = 0; }
    // A tree deeper than the stack is refused at scene creation (api.cu); should an entry ever not fit, it is DROPPED --
    // the traversal then ends early instead of popping a bogus node and walking the tree again from the root.
    __device__ __forceinline__ void push(int v)
    {
        if (sp < kStackSm) sm[sp * stride] = v;
        else if (sp < kBvhStackDepth) local[sp - kStackSm] = v;
        else return;
        sp++;
    }
    __device__ __forceinline__ int pop()
    {
        sp--;
        return sp < kStackSm ? sm[sp * stride] : local[sp - kStackSm];
    }
};

__device__ __forceinline__ float safe_rcp(float d)
{
    return 1.0f / (fabsf(d) > 1e-30f ? d : copysignf(1e-30f, d));
}

// Closest hit (ANY = false) or first hit with t < tmax (ANY = true), FP32.
// Ties on t resolve to the lowest prim id (the exhaustive loop's strict <).
template <bool ANY>
__device__ __forceinline__ Hit trace_fast(const float4* __restrict__ nodes, const float4* __restrict__ tris, f3 o, f3 d,
                                          int exclude, float tmin, float tmax, int* stack_sm, int stack_stride, int init_prim = -1)
{
    TStack st;
    st.init(stack_sm, stack_stride);
    f3 id = mk<float>(safe_rcp(d.x), safe_rcp(d.y), safe_rcp(d.z));
    f3 ood = mk<float>(o.x * id.x, o.y * id.y, o.z * id.z);
    // init_prim >= 0: the caller already holds a hit (init_prim at t = tmax); only something that beats it
    // under the closest-hit order (smaller t, or equal t and lower prim) replaces it -- the visibility test
    // "closest hit IS the sampled light triangle" of main.cpp:309-311 with an any-hit early out.
    Hit best;
    best.prim = init_prim; best.b = 0.f; best.g = 0.f; best.t = tmax;
    int node = 0;
    for (;;) {
        if (node >= 0) {
            const float4 n0 = __ldg(nodes + 4 * node + 0);
            const float4 n1 = __ldg(nodes + 4 * node + 1);
            const float4 n2 = __ldg(nodes + 4 * node + 2);
            const float4 n3 = __ldg(nodes + 4 * node + 3);
            float ax = n0.x * id.x - ood.x, bx = n0.y * id.x - ood.x;
            float ay = n0.z * id.y - ood.y, by = n0.w * id.y - ood.y;
            float az = n2.x * id.z - ood.z, bz = n2.y * id.z - ood.z;
            float t0n = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.f));
            float t0f = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), best.t));
            ax = n1.x * id.x - ood.x; bx = n1.y * id.x - ood.x;
            ay = n1.z * id.y - ood.y; by = n1.w * id.y - ood.y;
            az = n2.z * id.z - ood.z; bz = n2.w * id.z - ood.z;
            float t1n = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.f));
            float t1f = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), best.t));
            bool h0 = t0n <= t0f, h1 = t1n <= t1f;
            int c0 = __float_as_int(n3.x), c1 = __float_as_int(n3.y);
            if (h0 && h1) {
                bool swap = t1n < t0n;
                st.push(swap ? c0 : c1);
                node = swap ? c1 : c0;
                continue;
            }
            if (h0) { node = c0; continue; }
            if (h1) { node = c1; continue; }
        } else {
            int code = ~node;
            int first = code >> 3, cnt = (code & 7) + 1;
            for (int i = 0; i < cnt; i++) {
                const float4 v0 = __ldg(tris + 3 * (first + i) + 0);
                int prim = __float_as_int(v0.w);
                if (prim == exclude) continue;
                const float4 v1 = __ldg(tris + 3 * (first + i) + 1);
                const float4 v2 = __ldg(tris + 3 * (first + i) + 2);
                float be, ga, tt;
                if (tri_test_fast(o, d, to_f3(v0), to_f3(v1), to_f3(v2), tmin, &be, &ga, &tt)) {
                    if (tt < best.t || (tt == best.t && (unsigned)prim < (unsigned)best.prim)) {
                        best.prim = prim; best.b = be; best.g = ga; best.t = tt;
                        if (ANY) return best;
                    }
                }
            }
        }
        if (st.sp == 0) break;
        node = st.pop();
    }
    return best;
}

// Closest hit, FP64 boxes + exact leaf test (parity mode).  `ANY`/tmax as above.
template <bool ANY>
__device__ __forceinline__ HitD trace_exact(const float4* __restrict__ nodes, const float4* __restrict__ tris, d3 o, d3 d,
                                            int exclude, double tmax, int* stack_sm, int stack_stride)
{
    TStack st;
    st.init(stack_sm, stack_stride);
    auto rcp = [](double x) { return 1.0 / (fabs(x) > 1e-300 ? x : copysign(1e-300, x)); };
    d3 id = mk<double>(rcp(d.x), rcp(d.y), rcp(d.z));
    HitD best;
    best.prim = -1; best.b = 0; best.g = 0; best.t = tmax;
    int node = 0;
    for (;;) {
        if (node >= 0) {
            const float4 n0 = __ldg(nodes + 4 * node + 0);
            const float4 n1 = __ldg(nodes + 4 * node + 1);
            const float4 n2 = __ldg(nodes + 4 * node + 2);
            const float4 n3 = __ldg(nodes + 4 * node + 3);
            double ax = ((double)n0.x - o.x) * id.x, bx = ((double)n0.y - o.x) * id.x;
            double ay = ((double)n0.z - o.y) * id.y, by = ((double)n0.w - o.y) * id.y;
            double az = ((double)n2.x - o.z) * id.z, bz = ((double)n2.y - o.z) * id.z;
            double t0n = fmax(fmax(fmin(ax, bx), fmin(ay, by)), fmax(fmin(az, bz), 0.0));
            double t0f = fmin(fmin(fmax(ax, bx), fmax(ay, by)), fmin(fmax(az, bz), best.t));
            ax = ((double)n1.x - o.x) * id.x; bx = ((double)n1.y - o.x) * id.x;
            ay = ((double)n1.z - o.y) * id.y; by = ((double)n1.w - o.y) * id.y;
            az = ((double)n2.z - o.z) * id.z; bz = ((double)n2.w - o.z) * id.z;
            double t1n = fmax(fmax(fmin(ax, bx), fmin(ay, by)), fmax(fmin(az, bz), 0.0));
            double t1f = fmin(fmin(fmax(ax, bx), fmax(ay, by)), fmin(fmax(az, bz), best.t));
            bool h0 = t0n <= t0f, h1 = t1n <= t1f;
            int c0 = __float_as_int(n3.x), c1 = __float_as_int(n3.y);
            if (h0 && h1) {
                bool swap = t1n < t0n;
                st.push(swap ? c0 : c1);
                node = swap ? c1 : c0;
                continue;
            }
            if (h0) { node = c0; continue; }
            if (h1) { node = c1; continue; }
        } else {
            int code = ~node;
            int first = code >> 3, cnt = (code & 7) + 1;
            for (int i = 0; i < cnt; i++) {
                const float4 v0 = __ldg(tris + 3 * (first + i) + 0);
                int prim = __float_as_int(v0.w);
                if (prim == exclude) continue;
                const float4 v1 = __ldg(tris + 3 * (first + i) + 1);
                const float4 v2 = __ldg(tris + 3 * (first + i) + 2);
                double be, ga, tt;
                if (tri_test_exact(o, d, to_d3(v0), to_d3(v1), to_d3(v2), &be, &ga, &tt)) {
                    if (tt < best.t || (tt == best.t && (unsigned)prim < (unsigned)best.prim)) {
                        best.prim = prim; best.b = be; best.g = ga; best.t = tt;
                        if (ANY) return best;
                    }
                }
            }
        }
        if (st.sp == 0) break;
        node = st.pop();
    }
    return best;
}

// ---- ray-stream tracing with dynamic refill ---------------------------------------------------------------
// The wavefront trace kernels (k_extend: closest hit, k_shadow: any hit) run persistent warps over a compact
// array of rays.  A lane keeps its ray until the traversal ends; when fewer than `refill` lanes of the warp (bvh.h: kRefill4 / kRefill8) are
// still traversing, the idle lanes pull the next rays of the stream (one atomicAdd per refill, ballot/popc
// ranks) instead of waiting for the slowest ray of their original group of 32 -- incoherent rays through a
// deep BVH differ by an order of magnitude in node visits.  Results are written by ray index, so the order in
// which rays finish is irrelevant.
//   load(idx, o, d, exclude, tmax, init_prim)   fetches ray idx
//   store(idx, hit)                             consumes its result

#ifndef MCPT_VOTE_A
#define MCPT_VOTE_A 1
#define MCPT_VOTE_B 1
#endif
#ifndef MCPT_VOTE8_A
#define MCPT_VOTE8_A 1
#define MCPT_VOTE8_B 4
#endif
template <bool ANY, class Load, class Store>
__device__ __forceinline__ void trace_stream(const float4* __restrict__ nodes, const float4* __restrict__ tris, float tmin, int n,
                                             int* counter, int* stack_sm, int stack_stride, int refill, Load load, Store store)
{
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    TStack st;
    st.init(stack_sm, stack_stride);
    bool active = false, more = true;
    int idx = -1, exclude = -1, node = 0;
    f3 o = mk<float>(0.f, 0.f, 0.f), d = mk<float>(0.f, 0.f, 1.f), id = d, ood = o;
    int sx = 0, sy = 0, sz = 0;  // 1 where the ray runs towards -axis: the near plane of a box is then its `hi` plane
    Hit best;
    best.prim = -1; best.b = best.g = 0.f; best.t = 0.f;
    for (;;) {
        const unsigned act = __ballot_sync(full, active);
        if (more && __popc(act) < refill) {
            const unsigned idle = ~act;
            int base = 0;
            if (lane == 0) base = atomicAdd(counter, __popc(idle));
            base = __shfl_sync(full, base, 0);
            if (!active) {
                const int my = base + __popc(idle & ((1u << lane) - 1u));
                if (my < n) {
                    float tmax;
                    int init_prim;
                    idx = my;
                    load(idx, o, d, exclude, tmax, init_prim);
                    id = mk<float>(safe_rcp(d.x), safe_rcp(d.y), safe_rcp(d.z));
                    ood = mk<float>(o.x * id.x, o.y * id.y, o.z * id.z);
                    sx = id.x < 0.f; sy = id.y < 0.f; sz = id.z < 0.f;
                    best.prim = init_prim; best.b = 0.f; best.g = 0.f; best.t = tmax;
                    st.sp = 0;
                    node = 0;
                    active = true;
                }
            }
            if (base + __popc(idle) >= n) more = false;
        }
        if (!__any_sync(full, active)) {
            if (!more) break;
            continue;
        }
        for (;;) {
            // One warp-synchronous step.  The lanes vote: if at least as many sit at an inner node as at a leaf, the
            // inner lanes test their two child boxes, otherwise the leaf lanes test their triangles; the minority waits
            // one step.  The vote keeps the warp converged and every step at >= 50 % of the active lanes.  Measured on
            // the 1 M-triangle soup: a per-lane `while (node >= 0)` lets the lanes drift apart for good (4.9 of 32
            // threads per instruction), a two-phase while-while idles at the 1-triangle leaves (6.5).
            const bool at_inner = active && node >= 0, at_leaf = active && node < 0;
            const unsigned mi = __ballot_sync(full, at_inner), ml = __ballot_sync(full, at_leaf);
            bool done = false;
            if (MCPT_VOTE_A * __popc(mi) >= MCPT_VOTE_B * __popc(ml)) {
                if (at_inner) {
                    // 4-wide node (bvh.h nodes4): seven 16-byte vectors, child k in component k.  The ray's direction
                    // signs pick the near and the far plane of every slab by ADDRESS (lo and hi planes are neighbouring
                    // vectors), so a child costs 6 FMAs + 4 min/max instead of 6 + 10: the step is ALU-pipe-bound.
                    const float4* q = nodes + 7 * node;
                    const float4 nx = __ldg(q + sx), fx = __ldg(q + (sx ^ 1)), ny = __ldg(q + 2 + sy), fy = __ldg(q + 2 + (sy ^ 1));
                    const float4 nz = __ldg(q + 4 + sz), fz = __ldg(q + 4 + (sz ^ 1)), cq = __ldg(q + 6);
                    float dn[4];
                    int cc[4];
#define MCPT_SLAB(k, C)                                                                                      \
                    {                                                                                        \
                        const float tn = fmaxf(fmaxf(nx.C * id.x - ood.x, ny.C * id.y - ood.y), fmaxf(nz.C * id.z - ood.z, 0.f));    \
                        const float tf = fminf(fminf(fx.C * id.x - ood.x, fy.C * id.y - ood.y), fminf(fz.C * id.z - ood.z, best.t)); \
                        cc[k] = __float_as_int(cq.C);                                                        \
                        dn[k] = (tn <= tf && cc[k] != kBvhNoChild) ? tn : CUDART_INF_F;                      \
                    }
                    MCPT_SLAB(0, x) MCPT_SLAB(1, y) MCPT_SLAB(2, z) MCPT_SLAB(3, w)
#undef MCPT_SLAB
                    // sort the four (entry distance, child) pairs, nearest first (5-comparator network)
#define MCPT_CSWAP(i, j)                                                                                     \
                    {                                                                                        \
                        const bool sw = dn[j] < dn[i];                                                       \
                        const float td = sw ? dn[j] : dn[i], te = sw ? dn[i] : dn[j];                        \
                        const int tc = sw ? cc[j] : cc[i], tg = sw ? cc[i] : cc[j];                          \
                        dn[i] = td; dn[j] = te; cc[i] = tc; cc[j] = tg;                                      \
                    }
                    MCPT_CSWAP(0, 1) MCPT_CSWAP(2, 3) MCPT_CSWAP(0, 2) MCPT_CSWAP(1, 3) MCPT_CSWAP(1, 2)
#undef MCPT_CSWAP
                    if (dn[0] < CUDART_INF_F) {
                        if (dn[3] < CUDART_INF_F) st.push(cc[3]);
                        if (dn[2] < CUDART_INF_F) st.push(cc[2]);
                        if (dn[1] < CUDART_INF_F) st.push(cc[1]);
                        node = cc[0];
                    } else if (st.sp == 0) done = true;
                    else node = st.pop();
                }
            } else if (at_leaf) {
                const int code = ~node;
                const int first = code >> 3, cnt = (code & 7) + 1;
                for (int i = 0; i < cnt; i++) {
                    const float4 v0 = __ldg(tris + 3 * (first + i) + 0);
                    const int prim = __float_as_int(v0.w);
                    if (prim == exclude) continue;
                    const float4 v1 = __ldg(tris + 3 * (first + i) + 1);
                    const float4 v2 = __ldg(tris + 3 * (first + i) + 2);
                    float be, ga, tt;
                    if (tri_test_fast(o, d, to_f3(v0), to_f3(v1), to_f3(v2), tmin, &be, &ga, &tt)) {
                        if (tt < best.t || (tt == best.t && (unsigned)prim < (unsigned)best.prim)) {
                            best.prim = prim; best.b = be; best.g = ga; best.t = tt;
                            if (ANY) { done = true; break; }
                        }
                    }
                }
                if (!done) {
                    if (st.sp == 0) done = true;
                    else node = st.pop();
                }
            }
            if (done) {
                store(idx, best);
                active = false;
            }
            const unsigned m = __ballot_sync(full, active);
            if (m == 0u || (more && __popc(m) < refill)) break;
        }
    }
}

// ---- compressed 8-wide BVH (bvh.h: nodes8 / tris8): the same ray-stream tracer for large scenes ---------------------------
// One node visit = five 16-byte loads (80 B) for eight children, against 2 x seven for the two levels of 4-wide nodes it
// replaces: the trace kernels on the 1 M-triangle soup were bound by the L1 pipeline (98.8 % busy: every divergent LDG.128
// of a warp costs 32 wavefronts), so bytes and loads per visit are what counts.  Child boxes are bytes on the node's own
// grid (p + q * 2^e); the slab test runs on them directly: t = q * (2^e / d) + (p - o) / d, one FMA per plane after the
// byte-to-float conversion.  Children are visited in octant order (slot ^ ray octant, highest first) taken straight from
// the bit positions of the hit mask -- no distance sort.  State per lane: a node group G = (first child node, hit bits
// 24..31 | inner-child mask) and a triangle group T = (first triangle, hit bits 0..23); the stack holds node groups.
#ifndef MCPT_STACK8_SM
#define MCPT_STACK8_SM 6
#endif
constexpr int kStack8Sm = MCPT_STACK8_SM;
struct TStack8 {
    uint2* sm;
    int stride, sp;
    uint2 local[kBvh8StackDepth - kStack8Sm];
    __device__ __forceinline__ void init(uint2* base, int stride_) { sm = base; stride = stride_; sp = 0; }
    __device__ __forceinline__ void push(uint2 v)
    {
        if (sp < kStack8Sm) sm[sp * stride] = v;
        else if (sp < kBvh8StackDepth) local[sp - kStack8Sm] = v;
        else return;
        sp++;
    }
    __device__ __forceinline__ uint2 pop()
    {
        sp--;
        return sp < kStack8Sm ? sm[sp * stride] : local[sp - kStack8Sm];
    }
};

// byte i of w as a float.  MCPT_BYTE_MAGIC: one PRMT drops the byte into the mantissa of 2^23 and one FADD takes 2^23
// off again (both full rate) instead of the I2F conversion.
#ifndef MCPT_BYTE_MAGIC
#define MCPT_BYTE_MAGIC 2
#endif
__device__ __forceinline__ float byte_f_magic(unsigned w, int i) { return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7650u + (unsigned)i)) - 8388608.0f; }
__device__ __forceinline__ float byte_f_cvt(unsigned w, int i) { return (float)((w >> (8 * i)) & 0xffu); }
// MCPT_BYTE_MAGIC = how many of the six plane conversions of a child use PRMT + FADD instead of I2F, in the order far z, far y,
// far x, near z, near y, near x (0 = all I2F, 6 = none): the 48 conversions of a visit keep the XU pipe 65 % busy, the magic form
// moves some of them to the ALU / FMA pipes.
template <int PLANE> __device__ __forceinline__ float byte_f(unsigned w, int i)
{
    if (PLANE < MCPT_BYTE_MAGIC) return byte_f_magic(w, i);
    return byte_f_cvt(w, i);
}

template <bool ANY, class Load, class Store>
__device__ __forceinline__ void trace_stream8(const float4* __restrict__ nodes8, const float4* __restrict__ tris8, float tmin, int n,
                                              int* counter, uint2* stack_sm, int stack_stride, int refill, Load load, Store store)
{
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    TStack8 st;
    st.init(stack_sm, stack_stride);
    bool active = false, more = true;
    int idx = -1, exclude = -1;
    f3 o = mk<float>(0.f, 0.f, 0.f), d = mk<float>(0.f, 0.f, 1.f), id = d, ood = o;
    unsigned oct = 0;
    uint2 G = make_uint2(0u, 0u), T = make_uint2(0u, 0u);
    Hit best;
    best.prim = -1; best.b = best.g = 0.f; best.t = 0.f;
    for (;;) {
        const unsigned act = __ballot_sync(full, active);
        if (more && __popc(act) < refill) {
            const unsigned idle = ~act;
            int base = 0;
            if (lane == 0) base = atomicAdd(counter, __popc(idle));
            base = __shfl_sync(full, base, 0);
            if (!active) {
                const int my = base + __popc(idle & ((1u << lane) - 1u));
                if (my < n) {
                    float tmax;
                    int init_prim;
                    idx = my;
                    load(idx, o, d, exclude, tmax, init_prim);
                    id = mk<float>(safe_rcp(d.x), safe_rcp(d.y), safe_rcp(d.z));
                    ood = mk<float>(o.x * id.x, o.y * id.y, o.z * id.z);
                    oct = (id.x < 0.f ? 1u : 0u) | (id.y < 0.f ? 2u : 0u) | (id.z < 0.f ? 4u : 0u);
                    best.prim = init_prim; best.b = 0.f; best.g = 0.f; best.t = tmax;
                    st.sp = 0;
                    G = make_uint2(0u, 0x80000000u);  // the root as a group of one: node 0, no siblings
                    T = make_uint2(0u, 0u);
                    active = true;
                }
            }
            if (base + __popc(idle) >= n) more = false;
        }
        if (!__any_sync(full, active)) {
            if (!more) break;
            continue;
        }
        for (;;) {
            // invariant: an active lane has triangles waiting in T or inner hits waiting in G.  Majority vote as in trace_stream.
            const bool at_leaf = active && T.y != 0u, at_inner = active && T.y == 0u;
            const unsigned mi = __ballot_sync(full, at_inner), ml = __ballot_sync(full, at_leaf);
            bool done = false;
            if (MCPT_VOTE8_A * __popc(mi) >= MCPT_VOTE8_B * __popc(ml)) {
                if (at_inner) {
                    // nearest pending child of G in octant order = highest hit bit
                    const unsigned hits = G.y & 0xff000000u;
                    const int b = 31 - __clz(hits);
                    const unsigned imask = G.y & 0xffu;
                    G.y &= ~(1u << b);
                    if (G.y & 0xff000000u) st.push(G);
                    const unsigned slot = (unsigned)(b - 24) ^ oct;
                    const unsigned nidx = G.x + __popc(imask & ((1u << slot) - 1u));
                    const float4* q = nodes8 + 5 * (size_t)nidx;
                    const float4 n0 = __ldg(q), n1 = __ldg(q + 1), n2 = __ldg(q + 2), n3 = __ldg(q + 3), n4 = __ldg(q + 4);
                    const unsigned w3 = __float_as_uint(n0.w);
                    const float sx = __uint_as_float((w3 & 0xffu) << 23) * id.x, sy = __uint_as_float(((w3 >> 8) & 0xffu) << 23) * id.y,
                                sz = __uint_as_float(((w3 >> 16) & 0xffu) << 23) * id.z;
                    const float bx = __fmaf_rn(n0.x, id.x, -ood.x), by = __fmaf_rn(n0.y, id.y, -ood.y), bz = __fmaf_rn(n0.z, id.z, -ood.z);
                    unsigned hitmask = 0u;
#pragma unroll
                    for (int h = 0; h < 2; h++) {
                        const unsigned meta4 = __float_as_uint(h ? n1.w : n1.z);
                        const unsigned inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
                        const unsigned inner_mask4 = (inner4 >> 4) * 0xffu;
                        const unsigned bit4 = (meta4 ^ ((oct * 0x01010101u) & inner_mask4)) & 0x1f1f1f1fu;
                        const unsigned cnt4 = (meta4 >> 5) & 0x07070707u;
                        const unsigned qlx = __float_as_uint(h ? n2.y : n2.x), qly = __float_as_uint(h ? n2.w : n2.z), qlz = __float_as_uint(h ? n3.y : n3.x);
                        const unsigned qhx = __float_as_uint(h ? n3.w : n3.z), qhy = __float_as_uint(h ? n4.y : n4.x), qhz = __float_as_uint(h ? n4.w : n4.z);
                        const unsigned nx = (oct & 1u) ? qhx : qlx, fx = (oct & 1u) ? qlx : qhx;
                        const unsigned ny = (oct & 2u) ? qhy : qly, fy = (oct & 2u) ? qly : qhy;
                        const unsigned nz = (oct & 4u) ? qhz : qlz, fz = (oct & 4u) ? qlz : qhz;
#pragma unroll
                        for (int i = 0; i < 4; i++) {
                            const float tn = fmaxf(fmaxf(__fmaf_rn(byte_f<5>(nx, i), sx, bx), __fmaf_rn(byte_f<4>(ny, i), sy, by)),
                                                   fmaxf(__fmaf_rn(byte_f<3>(nz, i), sz, bz), 0.f));
                            const float tf = fminf(fminf(__fmaf_rn(byte_f<2>(fx, i), sx, bx), __fmaf_rn(byte_f<1>(fy, i), sy, by)),
                                                   fminf(__fmaf_rn(byte_f<0>(fz, i), sz, bz), best.t));
                            if (tn <= tf) hitmask |= ((cnt4 >> (8 * i)) & 0xffu) << ((bit4 >> (8 * i)) & 0xffu);
                        }
                    }
                    G = make_uint2(__float_as_uint(n1.x), (hitmask & 0xff000000u) | (w3 >> 24));
                    T = make_uint2(__float_as_uint(n1.y), hitmask & 0x00ffffffu);
                }
            } else if (at_leaf) {
                while (T.y) {
                    const int bit = __ffs(T.y) - 1;
                    T.y &= T.y - 1u;
                    const float4* tp = tris8 + 3 * (size_t)(T.x + (unsigned)bit);
                    const float4 v0 = __ldg(tp);
                    const int prim = __float_as_int(v0.w);
                    if (prim == exclude) continue;
                    const float4 v1 = __ldg(tp + 1), v2 = __ldg(tp + 2);
                    float be, ga, tt;
                    if (tri_test_fast(o, d, to_f3(v0), to_f3(v1), to_f3(v2), tmin, &be, &ga, &tt)) {
                        if (tt < best.t || (tt == best.t && (unsigned)prim < (unsigned)best.prim)) {
                            best.prim = prim; best.b = be; best.g = ga; best.t = tt;
                            if (ANY) { done = true; break; }
                        }
                    }
                }
            }
            // restore the invariant: nothing waiting here -> next group from the stack, or the ray is finished
            if (active && !done && T.y == 0u && !(G.y & 0xff000000u)) {
                if (st.sp == 0) done = true;
                else G = st.pop();
            }
            if (done) {
                store(idx, best);
                active = false;
            }
            const unsigned m = __ballot_sync(full, active);
            if (m == 0u || (more && __popc(m) < refill)) break;
        }
    }
}

}  // namespace mcpt
