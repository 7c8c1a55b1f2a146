// csrc/render.cu -- the wavefront path tracer: replacement for the pixel/sample loop of main()
// (main.cpp:557-588) and the three recursive estimators shade / shade_with_brdf / shade_with_mis
// (main.cpp:269-494), as bounce-synchronous kernels over a queue of rays.
//
// One pass renders `batch_spp` samples of every pixel:  slot = sample_in_batch * npix + pixel.
//   k_primary   camera rays (main.cpp:563-564) + closest hit, once per pixel (no jitter: every sample of a
//               pixel shoots the same ray, F4) -> primary_hit[pixel]; with MCPT_JITTER k_raygen + k_extend instead
//   k_lightprep spherical strategies: the O(L) light-weight sweep of the first vertex, once per pixel PAIR, and the
//               per-sample light draws from its CDF
//   then per bounce, over the dense ray queue Q[0] and the hit records k_extend left for it:
//   k_lsel_classify / _offsets / _scatter + k_lightsel_sorted
//               spherical strategies, later vertices, scenes with per-primitive survivor lists (light_lists.h: > 32
//               lights): the sampling vertices grouped by list, packed sweeps of four vertices per list -> sel4[i]
//   k_lightsel  the same selection without lists: warp-cooperative sweep of the whole table, or (<= 32 lights) one vertex
//               per lane over the table in shared memory       (k_lighttree instead with MCPT_LIGHT_TREE)
//   k_shade     shade prologue (main.cpp:273-288), Russian roulette, light and BRDF sampling, the MIS light-only ray;
//               children are packed with __ballot_sync / popc prefix sums into the chunk's window of the staging
//               queue Q[1], visibility rays with their pending contributions go to the shadow queue
//   k_gather    exact offsets from chunk counts and group sums; staged children move to the dense queue in input
//               order (wait-free; keeps the rays of one path tree adjacent)
//   k_extend    closest hit of every ray of the dense queue -> hit[i]   (k_extend_exact with MCPT_LEAF_EXACT)
//   k_shadow    NEE visibility (main.cpp:309-311); visible rays add their contribution to the slot's accumulator
//               (k_shadow_exact: closest-hit-and-compare-ids, the reference's own formulation)
//   k_runsum    MIS path tree only: children of one path are adjacent in the queue, so the head of each
//               run of equal slots sums the run's emission contributions into that slot
//   the host reads the new queue length (4 bytes) and launches the next bounce
//   k_accumulate / k_finalize   per-pixel owner sums its slots (atomic-free), scales by 1/spp
// Depth is implied by the bounce index: the vertex shaded at bounce b has depth b+1, because every
// earlier vertex of its path was a non-emissive, front-facing vertex that passed RR.
// MCPT_PARITY_FP64 leaves this file altogether: render_parity.cu.
#include <chrono>
#include <cstdlib>
#include <cstdio>
#include <algorithm>
#include <cmath>
#include <cstring>
#include <mutex>
#include <string>
#include "brdf.cuh"
#include "lights.cuh"
#include "mcpt_internal.h"
#include "rng.cuh"
#include "traverse.cuh"

namespace mcpt {
namespace {

constexpr int kBlk = 128;             // threads per block (4 warps; a warp owns one chunk of queue entries at a time)
constexpr unsigned kFull = 0xffffffffu;
constexpr size_t kTargetSlots = 128u << 20;  // auto batch: up to 128 Mi path slots in flight per pass (memory permitting)

enum { CNT_VERTICES = 0, CNT_RAYS, CNT_SHADOW, CNT_LIGHTRAYS, CNT_PRIMARY_HITS, CNT_OVERFLOW, CNT_SWEEPS, CNT_EVALS, CNT_SQ_HOLES, CNT_N };

struct Queue {
    float4* org;   // o.xyz | exclude prim (int bits)          (Myobj.cpp:417 rotri)
    float4* dir;   // d.xyz | slot (int bits)
    float4* thr;   // throughput rgb | unused
    uint2* node;   // path-tree node code (MIS only)
};

struct ShadowQueue {
    float4* org;   // o.xyz | t of the sampled light point (the ray's tmax)
    float4* dir;   // d.xyz | slot (int bits)
    float4* ctr;   // contribution if visible, rgb | unused
    int2* ids;     // origin prim (excluded, Myobj.cpp:417) | sampled light prim
    int* count;
};

struct BounceArgs {
    DScene s;
    DCamera cam;
    Queue in, out;
    float4* Lacc;              // [slots] per-slot radiance accumulator (owner-exclusive)
    float4* contrib;           // [capacity] MIS: per-element emission contribution | slot
    const float4* primary_hit; // [npix] prim | beta | gamma | t
    const float4* hit;         // [capacity] closest hit of in-queue ray i, written by k_extend: prim | beta | gamma | t
    ShadowQueue sq;            // NEE visibility rays of this bounce with their pending contributions
    const int2* sel;           // [slots] first-vertex light drawn by k_lightprep: index | weight bits
    const float* wsum_pix;     // [npix] first-vertex weights_sum
    float4* sel4;              // [capacity] later vertices, drawn by k_lightsel for in-queue ray i: index bits | weight | weights_sum
    int* counts;               // [chunks] children staged by each chunk
    int* groupsum;             // [chunks / kGroup] sums of counts, filled with one integer atomic per chunk
    int* ticket;               // chunk ticket counter
    int* skey;                 // [capacity] survivor list of the vertex of in-queue ray i that samples a light, -1 otherwise (k_lsel_classify)
    int* sorted;               // [capacity + 4 nlists] those queue indices grouped by list, runs padded to whole groups with -1
    int* lhist;                // [nlists] vertices per list, then (k_lsel_offsets) the scatter cursors
    int* ngroups;              // [1] groups of kSelV sorted entries
    unsigned long long* counters;
    uint64_t seed;
    int in_count, capacity, npix, batch_spp, sample_base, bounce, max_depth;
    uint32_t flags;
    float p_rr;
    int by_lists;              // this bounce's light selection runs over the survivor lists (k_lsel_*): skey[] is valid
};

// ---- small helpers --------------------------------------------------------------------------------------
__device__ __forceinline__ PhongMat<float> load_mat(const DScene& s, int m)
{
    const float4 a = __ldg(s.mats + 2 * m), b = __ldg(s.mats + 2 * m + 1);
    PhongMat<float> r;
    r.kd = to_f3(a); r.ns = a.w; r.ks = to_f3(b);
    return r;
}

// camera ray of pixel (i,j), main.cpp:563-564; (di,dj) = optional sub-pixel jitter in [-0.5,0.5)
__device__ __forceinline__ f3 camera_dir(const DCamera& c, int pixel, float di, float dj)
{
    int i = pixel / c.W, j = pixel - i * c.W;
    float dx = -c.pixellen * ((float)i + di - c.ci);
    float dy = c.pixellen * ((float)j + dj - c.cj);
    f3 d = mk<float>(c.U[0] * dx + c.V[0] * dy + c.N[0] * c.wlen, c.U[1] * dx + c.V[1] * dy + c.N[1] * c.wlen,
                     c.U[2] * dx + c.V[2] * dy + c.N[2] * c.wlen);
    return normalize_fast(d);
}

// ---- warp-cooperative spherical-triangle light selection --------------------------------------------------
// The reference loops over ALL light triangles at every vertex (Mylight.cpp:169-259 / :329-421) and then
// draws one with probability weight/weights_sum (:271-276).  Here a warp serves one vertex at a time: lanes
// stride over the SoA light table (warp_light_partial, lights.cuh), a shuffle scan gives weights_sum and the
// lane whose partial interval contains u*weights_sum, and a second short pass over that lane's lights resolves
// the triangle.  Same distribution, O(L/32) per lane per vertex, and no lane idles while others loop.
struct LightPick { int light; float w; };

// `part` = this lane's partial sum, `incl` its inclusive scan over lanes, `total` the sum over all lanes.
// LIST: the sweep ran over a survivor list of n lsoa offsets (light_lists.h); otherwise over the whole table (n = s.L).
template <bool LIST>
__device__ __forceinline__ LightPick warp_light_pick(const DScene& s, const uint32_t* __restrict__ list, int n, f3 X, f3 Nn, float part,
                                                     float incl, float total, float uu)
{
    const int lane = threadIdx.x & 31;
    LightPick r;
    r.light = -1; r.w = 0.f;
    if (!(total > 0.f)) return r;
    const float target = uu * total;
    const unsigned nz = __ballot_sync(kFull, part > 0.f);
    const unsigned ge = __ballot_sync(kFull, incl >= target && part > 0.f);
    const int j = ge ? (__ffs(ge) - 1) : (31 - __clz(nz));
    const float rem = target - (__shfl_sync(kFull, incl, j) - __shfl_sync(kFull, part, j));
    float running = 0.f;
    int last_l = -1;
    float last_w = 0.f;
    for (int k0 = 0; j + 32 * k0 < n && r.light < 0; k0 += 32) {
        const int pos = j + 32 * (k0 + lane);
        const int off = pos < n ? (LIST ? (int)__ldg(list + pos) : lsoa_index(pos)) : -1;
        const float w = (off >= 0) ? light_weight_at(s, off, X, Nn) : 0.f;
        float wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            float t = __shfl_up_sync(kFull, wi, o);
            if (lane >= o) wi += t;
        }
        const unsigned pos_m = __ballot_sync(kFull, w > 0.f);
        const unsigned hit = __ballot_sync(kFull, w > 0.f && running + wi >= rem);
        if (hit) {
            const int f = __ffs(hit) - 1;
            r.light = lsoa_light(__shfl_sync(kFull, off, f));
            r.w = __shfl_sync(kFull, w, f);
        } else if (pos_m) {
            const int f = 31 - __clz(pos_m);
            last_l = lsoa_light(__shfl_sync(kFull, off, f));
            last_w = __shfl_sync(kFull, w, f);
        }
        running += __shfl_sync(kFull, wi, 31);
    }
    if (r.light < 0) { r.light = last_l; r.w = last_w; }  // rounding: fall back to the lane's last survivor
    return r;
}

__device__ __forceinline__ float warp_scan_incl(float v)
{
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        float t = __shfl_up_sync(kFull, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

// ---- order-preserving queue compaction, wait-free ------------------------------------------------------------
// A chunk is the kChunk = 256 queue entries one warp takes per ticket.  The warp packs the 0..2 children of each entry
// with __ballot_sync / popc prefix sums, in input order, into the chunk's own window of a STAGING queue
// (stage[chunk * stride + k], stride = 256 or, for the MIS path tree, 512), records counts[chunk] and adds the count to the
// sum of its group of kGroup = 64 chunks (one integer
// atomic per warp: commutative, so the result is deterministic).  k_gather then turns group sums + counts into
// exact offsets and moves the chunk windows into the dense queue, preserving input order -- which keeps all
// rays of one path-tree adjacent (k_runsum) without any warp ever waiting for another.
#ifndef MCPT_SQ_GRAB
#define MCPT_SQ_GRAB 256
#endif
constexpr int kSqGrab = MCPT_SQ_GRAB;  // visibility-queue slots a warp of k_shade reserves per atomic
constexpr int kSerialLights = 32;  // at or below: per-lane light loop in k_lightsel instead of the warp-cooperative sweep
constexpr int kGroup = 64;    // chunks per k_gather block (16 Ki queue entries; threads 0..63 scan one chunk count each)

__global__ void __launch_bounds__(256) k_gather(const Queue stage, const Queue out, const int* __restrict__ counts,
                                                const int* __restrict__ groupsum, int nch, int stride, int capacity, int tree,
                                                int* out_count, unsigned long long* counters)
{
    __shared__ int s_red[8];
    __shared__ int s_base[kGroup];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = blockIdx.x;
    // 1. offset of this group = sum of the sums of all earlier groups
    int acc = 0;
    for (int j = tid; j < g; j += 256) acc += groupsum[j];
    acc = __reduce_add_sync(kFull, acc);
    if (lane == 0) s_red[warp] = acc;
    __syncthreads();
    int gbase = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) gbase += s_red[w];
    __syncthreads();
    // 2. exclusive scan of this group's chunk counts (one chunk per thread)
    const int c0 = g * kGroup + tid;
    const int tsum = (tid < kGroup && c0 < nch) ? counts[c0] : 0;
    int incl = tsum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(kFull, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_red[warp] = incl;
    __syncthreads();
    int wbase = 0, gtotal = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) { if (w < warp) wbase += s_red[w]; gtotal += s_red[w]; }
    if (tid < kGroup) s_base[tid] = gbase + wbase + incl - tsum;
    __syncthreads();
    if (g == gridDim.x - 1 && tid == 0) {
        const int total = gbase + gtotal;
        *out_count = total <= capacity ? total : -1;  // -1: the host abandons the pass and retries it with more room per slot
        if (total > capacity) atomicAdd(counters + CNT_OVERFLOW, (unsigned long long)(total - capacity));
    }
    // 3. move the staged entries: one thread per dense destination, its chunk found by bisection of the bases
    //    (largest chunk whose base <= destination; empty chunks share a base with their successor and lose)
    for (int e = tid; e < gtotal; e += 256) {
        const int dst = gbase + e;
        int lo = 0, hi = kGroup - 1;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (s_base[mid] <= dst) lo = mid; else hi = mid - 1;
        }
        const int src = (g * kGroup + lo) * stride + (dst - s_base[lo]);
        if (dst < capacity) {
            out.org[dst] = stage.org[src];
            out.dir[dst] = stage.dir[src];
            out.thr[dst] = stage.thr[src];
            if (tree) out.node[dst] = stage.node[src];
        }
    }
}

// ---- shade prologue shared by k_lightprep, k_lightsel and k_shade: p, N, back-face / emissive classification ----------
struct Vertex { f3 p, N; int mat, light; bool front; };

__device__ __forceinline__ Vertex make_vertex(const DScene& s, int prim, float hb, float hg, f3 wo)
{
    const float4 v0 = __ldg(s.pverts + 3 * prim), v1 = __ldg(s.pverts + 3 * prim + 1), v2 = __ldg(s.pverts + 3 * prim + 2);
    const float4 n0 = __ldg(s.pnorms + 3 * prim), n1 = __ldg(s.pnorms + 3 * prim + 1), n2 = __ldg(s.pnorms + 3 * prim + 2);
    const float al = 1.0f - hb - hg;
    Vertex v;
    v.p = to_f3(v0) * al + to_f3(v1) * hb + to_f3(v2) * hg;                  // main.cpp:26-29,273
    v.N = normalize_fast(to_f3(n0) * al + to_f3(n1) * hb + to_f3(n2) * hg);  // main.cpp:274
    v.front = !(dot(v.N, wo) < 0.f);                                         // main.cpp:277-280
    v.mat = __float_as_int(v0.w);
    v.light = __float_as_int(v1.w);
    return v;
}

// ---- primary rays ----------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlk) k_primary(DScene s, DCamera cam, int npix, float4* primary_hit)
{
    __shared__ int stack[kStackSm * kBlk];
    int pixel = blockIdx.x * kBlk + threadIdx.x;
    if (pixel >= npix) return;
    f3 o = mk<float>(cam.eye[0], cam.eye[1], cam.eye[2]);
    f3 d = camera_dir(cam, pixel, 0.f, 0.f);
    Hit h = trace_fast<false>(s.nodes, s.tris, o, d, -1, s.tmin, 3.0e38f, stack + threadIdx.x, kBlk);
    primary_hit[pixel] = make_float4(__int_as_float(h.prim), h.b, h.g, h.t);
}

// MCPT_LEAF_EXACT: the same ray resolved by the reference's FP64 triangle test (traverse.cuh: trace_exact)
__global__ void __launch_bounds__(kBlk) k_primary_exact(DScene s, DCamera cam, int npix, float4* primary_hit)
{
    __shared__ int stack[kStackSm * kBlk];
    int pixel = blockIdx.x * kBlk + threadIdx.x;
    if (pixel >= npix) return;
    f3 o = mk<float>(cam.eye[0], cam.eye[1], cam.eye[2]);
    f3 d = camera_dir(cam, pixel, 0.f, 0.f);
    HitD h = trace_exact<false>(s.nodes, s.tris, to_d3(o), to_d3(d), -1, 1.7976931348623157e308, stack + threadIdx.x, kBlk);
    primary_hit[pixel] = make_float4(__int_as_float(h.prim), (float)h.b, (float)h.g, (float)h.t);
}

// ---- first-vertex light selection, once per PIXEL ----------------------------------------------------------
// Without jitter every sample of a pixel shoots the same camera ray (F4, main.cpp:563-577), so the first path
// vertex -- and with it all L spherical-triangle weights and weights_sum -- is identical for the whole batch.
// One warp sweeps the light table once for a PAIR of neighbouring pixels (O(L), loads shared), leaving the
// per-lane running sums in shared memory; then lane k draws the light of sample k by inverting that two-level
// CDF (lane totals, then the lane's column) with two short binary searches -- O(log L) per sample instead of
// O(L).  sel[slot] = (light index | weight bits), -1 when the sample draws no light (unshadable vertex, or
// shade_with_mis's roulette, main.cpp:429-437, already ended it); wsum_pix[pixel] = weights_sum.
__global__ void __launch_bounds__(kBlk, 5) k_lightprep(DScene s, DCamera cam, const float4* __restrict__ primary_hit, int npix,
                                                    int batch_spp, int sample_base, uint64_t seed, float p_rr, int rr_first,
                                                    int max_depth, int use_cdf, int2* sel, float* wsum_pix, unsigned long long* counters)
{
    extern __shared__ float sm_prep[];
    const int lane = threadIdx.x & 31;
    float* cdf = sm_prep + (threadIdx.x >> 5) * (2 * s.Lpad + 64);  // [2][Lpad] running sums of the lane columns
    float* sinc = cdf + 2 * s.Lpad;                                  // [2][32] inclusive scan of the lane totals
    const int warp = (blockIdx.x * kBlk + threadIdx.x) >> 5, nwarps = (gridDim.x * kBlk) >> 5;
    const bool rr_forced = max_depth > 0 && 1 >= max_depth;
    unsigned c_sweeps = 0;
    unsigned c_evals = 0;  // in units of 32 (every sweep length is a multiple of 32)
    for (int pair = warp; 2 * pair < npix; pair += nwarps) {
        f3 X[2], Nn[2];
        bool shad[2];
        int lid[2] = {-1, -1};
#pragma unroll
        for (int v = 0; v < 2; v++) {
            const int pixel = 2 * pair + v;
            shad[v] = false;
            X[v] = Nn[v] = mk<float>(0.f, 0.f, 1.f);
            if (pixel < npix) {
                const float4 ph = __ldg(primary_hit + pixel);
                const int prim = __float_as_int(ph.x);
                if (prim >= 0) {
                    const Vertex vx = make_vertex(s, prim, ph.y, ph.z, -camera_dir(cam, pixel, 0.f, 0.f));
                    shad[v] = vx.front && vx.light < 0;
                    X[v] = vx.p; Nn[v] = vx.N;
                    if (shad[v] && s.nlists) lid[v] = __ldg(s.prim_list + prim);
                }
            }
        }
        // the sweep runs over the survivor list of the pixels' primitive (light_lists.h) when both pixels share one --
        // neighbouring pixels mostly do -- and over the whole table otherwise; n = entries swept (a multiple of 32)
        if (!shad[0]) lid[0] = lid[1];
        if (!shad[1]) lid[1] = lid[0];
        const bool by_list = s.nlists && lid[0] >= 0 && lid[0] == lid[1];
        const uint32_t* list = nullptr;
        int n = s.Lpad;
        if (by_list) { const int2 rg = __ldg(s.lrange + lid[0]); list = s.llist + rg.x; n = rg.y; }
        const int K = n >> 5;
        float part[2] = {0.f, 0.f}, incl[2] = {0.f, 0.f}, total[2] = {0.f, 0.f};
        if (shad[0] || shad[1]) {
            c_sweeps += (unsigned)shad[0] + (unsigned)shad[1];
            c_evals += (unsigned)(n >> 5) * ((unsigned)shad[0] + (unsigned)shad[1]);
            if (!shad[0]) { X[0] = X[1]; Nn[0] = Nn[1]; }
            if (!shad[1]) { X[1] = X[0]; Nn[1] = Nn[0]; }
            if (by_list) warp_light_partial_list<2>(s, list, n, X, Nn, part, use_cdf ? cdf : nullptr, s.Lpad);
            else warp_light_partial<2>(s, X, Nn, part, use_cdf ? cdf : nullptr);
#pragma unroll
            for (int v = 0; v < 2; v++) {
                incl[v] = warp_scan_incl(part[v]);
                total[v] = __shfl_sync(kFull, incl[v], 31);
                if (use_cdf) sinc[32 * v + lane] = incl[v];
            }
        }
        __syncwarp();
#pragma unroll
        for (int v = 0; v < 2; v++) {
            const int pixel = 2 * pair + v;
            if (pixel >= npix) break;
            if (lane == 0) wsum_pix[pixel] = shad[v] ? total[v] : 0.f;
            for (int s0 = 0; s0 < batch_spp; s0 += 32) {
                const int k = s0 + lane;
                const uint32_t sample = (uint32_t)(sample_base + k);
                bool need = false;
                float u = 0.f;
                if (k < batch_spp && shad[v] && total[v] > 0.f) {
                    need = true;
                    if (rr_first) need = !rr_forced && !(rng_draw4(seed, (uint32_t)pixel, sample, 1, 0).u[0] > p_rr);
                    if (need) u = rng_draw4(seed, (uint32_t)pixel, sample, 1, 1).u[0];
                }
                int mine = -1;
                float wmine = 0.f;
                if (use_cdf) {
                    if (need && K > 0) {
                        const float* si = sinc + 32 * v;
                        const float target = fmaxf(u, 1e-7f) * total[v];
                        int j = 0;  // first lane whose inclusive total reaches the target
#pragma unroll
                        for (int step = 16; step; step >>= 1)
                            if (si[j + step - 1] < target) j += step;
                        const float* col = cdf + v * s.Lpad + j;  // col[32 k]: running sum of lane j's lights
                        const float rem = fminf(target - (j ? si[j - 1] : 0.f), col[32 * (K - 1)]);
                        int lo = 0, hi = K - 1;
                        while (lo < hi) {
                            const int mid = (lo + hi) >> 1;
                            if (col[32 * mid] < rem) lo = mid + 1; else hi = mid;
                        }
                        const int pos = j + 32 * lo;  // position in the sweep: list entry or table index
                        const int off = by_list ? (int)__ldg(list + pos) : lsoa_index(pos);
                        mine = lsoa_light(off);
                        // the CDF difference only SELECTS; the weight handed on (sA = w / Lsum feeds arvo_direct) is
                        // re-evaluated, bit-identical to the sweep's and to what warp_light_pick returns at later vertices
                        wmine = (mine < s.L) ? light_weight_at(s, off, X[v], Nn[v]) : 0.f;
                        if (!(wmine > 0.f) || mine >= s.L) mine = -1;
                    }
                } else {
                    unsigned todo = __ballot_sync(kFull, need);
                    while (todo) {
                        const int src = __ffs(todo) - 1;
                        todo &= todo - 1;
                        const float us = __shfl_sync(kFull, u, src);
                        const LightPick pk = by_list ? warp_light_pick<true>(s, list, n, X[v], Nn[v], part[v], incl[v], total[v], us)
                                                     : warp_light_pick<false>(s, nullptr, s.L, X[v], Nn[v], part[v], incl[v], total[v], us);
                        if (lane == src) { mine = pk.light; wmine = pk.w; }
                    }
                }
                if (k < batch_spp) sel[(size_t)k * npix + pixel] = make_int2(mine, __float_as_int(wmine));
            }
        }
        __syncwarp();  // the next pair overwrites cdf / sinc
    }
    if (lane == 0 && c_sweeps) {
        atomicAdd(counters + CNT_SWEEPS, (unsigned long long)c_sweeps);  // vertices whose light table was swept
        atomicAdd(counters + CNT_EVALS, (unsigned long long)c_evals * 32);  // (light, vertex) weights evaluated for them
    }
}

// ---- light selection at the later vertices ---------------------------------------------------------------------
// The O(L) sweeps are the most expensive thing the spherical strategies do.  This kernel does nothing else -- rebuild
// the vertex, redo the (counter-based, hence repeatable) roulette and uniform draws that decide whether the vertex
// samples a light, sweep, pick -- and leaves sel4[i] = (light index bits, its weight, weights_sum) for k_shade, which
// then needs 94 instead of 126 registers.  Measured (veach-mis, B200): sweeps inside k_shade 34.2 ms per 64-spp step,
// here 33.0 ms with two lights in flight per lane (unroll 2); 4, 5, 6, 7 or 8 resident blocks per SM make no
// difference -- the loop is bound by the FMA-heavy pipe (FFMA2 occupies it for two cycles: 68 % busy at 61 % issue-
// active) and the MUFU bursts, not by latency.  Four vertices per sweep lost 20 % to padded groups.
#ifndef MCPT_SEL_MINBLK
#define MCPT_SEL_MINBLK 5
#endif
template <int STRAT>
__global__ void __launch_bounds__(kBlk, MCPT_SEL_MINBLK) k_lightsel(const BounceArgs a, int* cursor)
{
    constexpr bool kTree = (STRAT == MCPT_MIS);
    const DScene& s = a.s;
    const int lane = threadIdx.x & 31;
    const int nchunks = (a.in_count + 31) >> 5;
    const bool rr_forced = a.max_depth > 0 && a.bounce + 1 >= a.max_depth;
    // few lights: the table as pairs of lights in shared memory for the packed per-lane loop below (light_weight_2lights)
    __shared__ float2 sm_lp[(kSerialLights / 2) * kLightPairFields];
    if (s.L <= kSerialLights) {
        const int npairs = (s.L + 1) >> 1;
        for (int t = threadIdx.x; t < npairs * kLightPairFields; t += kBlk) {
            const int j = t / kLightPairFields, f = t - j * kLightPairFields;
            // field f of a light record: planes 0..2 xyz, plane 3 xyzw, plane 4 xyzw (lights.cuh); light L is a padding light (weight 0)
            const int plane = f < 9 ? f / 3 : (f < 13 ? 3 : 4), comp = f < 9 ? f % 3 : (f < 13 ? f - 9 : f - 13);
            float v[2];
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const float4 r = __ldg(s.lsoa + lsoa_index(2 * j + h) + 32 * plane);
                v[h] = comp == 0 ? r.x : comp == 1 ? r.y : comp == 2 ? r.z : r.w;
            }
            sm_lp[t] = make_float2(v[0], v[1]);
        }
        __syncthreads();
    }
    // Warp path: a sweep serves two vertices, and only ~1 queue entry in 9 needs one at the second bounce of veach-mis
    // (3.6 per chunk of 32), so pairing inside a chunk would leave every other chunk with a half-empty sweep.  The odd
    // vertex is carried (warp-uniform registers) and paired with the first one of the next chunk; results are written
    // by index from lane 0.
    bool pend = false;
    f3 PX = mk<float>(0.f, 0.f, 0.f), PN = mk<float>(0.f, 0.f, 1.f);
    float Pu = 0.f;
    int Pi = -1;
    unsigned c_evals = 0;  // in units of 32
    auto sweep_pair = [&](f3 X0, f3 N0, float u0, int i0, bool two, f3 X1, f3 N1, float u1, int i1) {
        f3 X[2] = {X0, two ? X1 : X0}, Nn[2] = {N0, two ? N1 : N0};
        const float uu[2] = {u0, u1};
        const int ii[2] = {i0, i1};
        float part[2];
        warp_light_partial<2>(s, X, Nn, part);
        c_evals += (unsigned)(s.Lpad >> 5) * (two ? 2u : 1u);
#pragma unroll
        for (int v = 0; v < 2; v++) {
            if (v == 1 && !two) break;
            const float incl = warp_scan_incl(part[v]);
            const float total = __shfl_sync(kFull, incl, 31);
            const LightPick pk = warp_light_pick<false>(s, nullptr, s.L, X[v], Nn[v], part[v], incl, total, uu[v]);
            if (lane == 0) a.sel4[ii[v]] = make_float4(__int_as_float(pk.light), pk.w, total, 0.f);
        }
    };
    // few lights: up to 63 waiting vertices per warp (position | uniform, normal | queue index)
    __shared__ float4 sm_wait[(kBlk / 32) * 64 * 2];
    float4* wbuf = sm_wait + (threadIdx.x >> 5) * 128;
    int q_n = 0;
    // one vertex per lane: the whole (short) table, two lights per packed evaluation (bit-identical to light_weight_fast),
    // selection in the same pass by weighted reservoir sampling with ONE recycled uniform -- P(l) = w_l / weights_sum exactly
    auto serial_pick = [&](const float4* e) {
        const float4 e0 = e[0], e1 = e[1];
        const f3 p = mk<float>(e0.x, e0.y, e0.z), N = mk<float>(e1.x, e1.y, e1.z);
        int chosen = -1;
        float wsum = 0.f, wch = 0.f;
        float uu = fminf(e0.w, 0.99999994f);
        for (int l = 0; l < s.L; l += 2) {
            const float2 w2 = light_weight_2lights(sm_lp + (l >> 1) * kLightPairFields, p, N);
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const float w = h ? w2.y : w2.x;
                if (w > 0.f && l + h < s.L) {
                    wsum += w;
                    const float pr = __fdividef(w, wsum);
                    if (uu < pr) { chosen = l + h; wch = w; uu = __fdividef(uu, pr); }
                    else uu = __fdividef(uu - pr, 1.f - pr);
                    uu = fminf(fmaxf(uu, 0.f), 0.99999994f);
                }
            }
        }
        a.sel4[__float_as_int(e1.w)] = make_float4(__int_as_float(chosen), wch, wsum, 0.f);
    };
    int next = 0;
    if (lane == 0) next = atomicAdd(cursor, 1);
    for (;;) {
        const int chunk = __shfl_sync(kFull, next, 0);
        if (chunk >= nchunks) break;
        if (lane == 0) next = atomicAdd(cursor, 1);
        const int i = chunk * 32 + lane;
        bool need = false;
        f3 p = mk<float>(0.f, 0.f, 0.f), N = mk<float>(0.f, 0.f, 1.f);
        float u = 0.f;
        if (i < a.in_count) {
            const float4 qh = a.hit[i];
            const int prim = __float_as_int(qh.x);
            if (prim >= 0) {
                const float4 qd = a.in.dir[i];
                const Vertex v = make_vertex(s, prim, qh.y, qh.z, -to_f3(qd));
                if (v.front && v.light < 0) {
                    const int slot = __float_as_int(qd.w);
                    const uint32_t pixel = (uint32_t)(slot % a.npix), sample = (uint32_t)(a.sample_base + slot / a.npix);
                    uint64_t node = (uint64_t)a.bounce + 1;
                    if (kTree) { const uint2 nd = a.in.node[i]; node = ((uint64_t)nd.y << 32) | nd.x; }
                    need = true;
                    // MIS (both forms): roulette before any sampling (main.cpp:429-437); NEE samples the light first (:300-316)
                    if (kTree || STRAT == MCPT_MIS_SINGLE) need = !rr_forced && !(rng_draw4(a.seed, pixel, sample, node, 0).u[0] > a.p_rr);
                    if (need) { u = rng_draw4(a.seed, pixel, sample, node, 1).u[0]; p = v.p; N = v.N; }
                }
            }
        }
        if (s.L <= kSerialLights) {
            // Few lights: the warp-cooperative sweep would spend most of its instructions on broadcasts, scans and the
            // second pass (measured on the 16-light interior: 13 ms per 32-spp pass).  Each lane walks the whole table
            // for ONE vertex instead (serial_pick below).  Only ~60 % of the entries of a chunk sample a light, so the
            // sampling vertices are first collected in the warp's buffer and served 32 at a time, all lanes busy.
            if (i < a.in_count && !need) a.sel4[i] = make_float4(__int_as_float(-1), 0.f, 0.f, 0.f);
            const unsigned mneed = __ballot_sync(kFull, need);
            if (need) {
                float4* e = wbuf + 2 * (q_n + __popc(mneed & ((1u << lane) - 1u)));
                e[0] = make_float4(p.x, p.y, p.z, u);
                e[1] = make_float4(N.x, N.y, N.z, __int_as_float(i));
            }
            q_n += __popc(mneed);
            __syncwarp();
            if (q_n >= 32) {
                q_n -= 32;
                serial_pick(wbuf + 2 * (q_n + lane));
                __syncwarp();
            }
            continue;
        }
        if (i < a.in_count && !need) a.sel4[i] = make_float4(__int_as_float(-1), 0.f, 0.f, 0.f);
        unsigned todo = __ballot_sync(kFull, need);
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const f3 X1 = mk<float>(__shfl_sync(kFull, p.x, src), __shfl_sync(kFull, p.y, src), __shfl_sync(kFull, p.z, src));
            const f3 N1 = mk<float>(__shfl_sync(kFull, N.x, src), __shfl_sync(kFull, N.y, src), __shfl_sync(kFull, N.z, src));
            const float u1 = __shfl_sync(kFull, u, src);
            const int i1 = chunk * 32 + src;
            if (!pend) { PX = X1; PN = N1; Pu = u1; Pi = i1; pend = true; }
            else { sweep_pair(PX, PN, Pu, Pi, true, X1, N1, u1, i1); pend = false; }
        }
    }
    if (pend) sweep_pair(PX, PN, Pu, Pi, false, PX, PN, Pu, Pi);
    if (lane < q_n) serial_pick(wbuf + 2 * lane);  // few lights: what is left in the warp's buffer
    if (lane == 0 && c_evals) atomicAdd(a.counters + CNT_EVALS, (unsigned long long)c_evals * 32);
}

// ---- the same selection over per-primitive survivor lists (light_lists.h) ---------------------------------------------
// About half of the lights are culled at EVERY vertex of a given primitive (they face away from all of it, or lie below its
// horizon everywhere): the scene carries, per primitive, the list of the others.  A sweep over the list instead of the
// table halves the evaluations -- but a packed sweep serves a group of vertices with ONE list, so the vertices that sample
// a light are first grouped by list:
//   k_lsel_classify   per queue entry: does its vertex sample a light (same decisions as k_lightsel)?  key[i] = its list,
//                     histogram of the lists (one atomic per distinct list per warp)
//   k_lsel_offsets    one block: exclusive scan of the counts, each rounded up to whole groups of kSelV; the slack is
//                     filled with -1
//   k_lsel_scatter    queue indices into their list's run (order inside a run is whatever the atomics give: the result
//                     for a vertex depends on its own list only, never on its neighbours in the group)
//   k_lightsel_sorted persistent warps, one group of kSelV vertices of one list per ticket: rebuild the vertices, sweep
//                     the list for all of them at once (the loads of a light serve 2 x kSelV/2 packed evaluations),
//                     scan, pick, sel4[i]
// Weights, weights_sum and probabilities are those of the full sweep up to the order of the FP32 additions.
#ifndef MCPT_SEL_V
#define MCPT_SEL_V 4
#endif
constexpr int kSelV = MCPT_SEL_V;
// Every list's run is kept as kSelSub sub-runs with a histogram bin and a scatter cursor each, chosen by the 32-entry group
// an entry sits in: the cursor atomics of k_lsel_classify / k_lsel_scatter all go to one word per list, and a handful of
// lists (the floor under the lights) take a large share of the vertices -- same-address atomics serialise in L2.
#ifndef MCPT_SEL_SUB
#define MCPT_SEL_SUB 8
#endif
constexpr int kSelSub = MCPT_SEL_SUB;
// running-sum checkpoints kept per lane column in k_lightsel_sorted: every kCdfSub-th list position (power of two; 1 = all)
#ifndef MCPT_CDF_SUB
#define MCPT_CDF_SUB 8
#endif
constexpr int kCdfSub = MCPT_CDF_SUB;
__device__ __forceinline__ int sel_bin(int list, int i) { return list * kSelSub + ((i >> 5) & (kSelSub - 1)); }

// the vertex of in-queue ray i if it samples a light: the decisions k_shade takes before light sampling, repeated
template <int STRAT>
__device__ __forceinline__ bool sel_vertex(const BounceArgs& a, int i, bool rr_forced, bool want_u, f3* p, f3* N, float* u, int* prim_out)
{
    constexpr bool kTree = (STRAT == MCPT_MIS);
    const float4 qh = a.hit[i];
    const int prim = __float_as_int(qh.x);
    if (prim < 0) return false;
    const float4 qd = a.in.dir[i];
    const Vertex v = make_vertex(a.s, prim, qh.y, qh.z, -to_f3(qd));
    if (!(v.front && v.light < 0)) return false;
    const int slot = __float_as_int(qd.w);
    const uint32_t pixel = (uint32_t)(slot % a.npix), sample = (uint32_t)(a.sample_base + slot / a.npix);
    uint64_t node = (uint64_t)a.bounce + 1;
    if (kTree) { const uint2 nd = a.in.node[i]; node = ((uint64_t)nd.y << 32) | nd.x; }
    // MIS (both forms): roulette before any sampling (main.cpp:429-437); NEE samples the light first (:300-316)
    if (kTree || STRAT == MCPT_MIS_SINGLE)
        if (rr_forced || rng_draw4(a.seed, pixel, sample, node, 0).u[0] > a.p_rr) return false;
    if (want_u) *u = rng_draw4(a.seed, pixel, sample, node, 1).u[0];
    *p = v.p; *N = v.N; *prim_out = prim;
    return true;
}

template <int STRAT>
__global__ void __launch_bounds__(256) k_lsel_classify(const BounceArgs a)
{
    const bool rr_forced = a.max_depth > 0 && a.bounce + 1 >= a.max_depth;
    const int lane = threadIdx.x & 31;
    for (int base = (blockIdx.x * 256 + threadIdx.x) & ~31; base < a.in_count; base += gridDim.x * 256) {
        const int i = base + lane;
        int k = -1;
        if (i < a.in_count) {
            f3 p, N;
            float u;
            int prim;
            if (sel_vertex<STRAT>(a, i, rr_forced, false, &p, &N, &u, &prim)) k = __ldg(a.s.prim_list + prim);
            a.skey[i] = k;  // -1: no light drawn here; k_shade looks at the key before it trusts sel4[i] (16 bytes less written per entry)
        }
        const unsigned act = __ballot_sync(kFull, k >= 0);
        if (k >= 0) {
            const unsigned peers = __match_any_sync(act, k);
            if (lane == __ffs(peers) - 1) atomicAdd(a.lhist + sel_bin(k, base), __popc(peers));
        }
    }
}

__global__ void __launch_bounds__(1024) k_lsel_offsets(int* __restrict__ hist, int nlists, int* __restrict__ sorted, int* __restrict__ ngroups)
{
    __shared__ int s_warp[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int per = (nlists + 1023) / 1024, l0 = tid * per, l1 = min(nlists, l0 + per);
    int mine = 0;
    for (int l = l0; l < l1; l++) mine += (hist[l] + kSelV - 1) / kSelV * kSelV;
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(kFull, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int w = s_warp[lane], wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(kFull, wi, o);
            if (lane >= o) wi += t;
        }
        s_warp[lane] = wi - w;
        if (lane == 31) *ngroups = wi / kSelV;
    }
    __syncthreads();
    int at = s_warp[warp] + incl - mine;
    for (int l = l0; l < l1; l++) {
        const int c = hist[l], padded = (c + kSelV - 1) / kSelV * kSelV;
        hist[l] = at;  // from here on: the scatter cursor of list l
        for (int k = c; k < padded; k++) sorted[at + k] = -1;
        at += padded;
    }
}

__global__ void __launch_bounds__(256) k_lsel_scatter(const BounceArgs a)
{
    // Two dependent round trips per 32 entries (the key from DRAM, then the cursor atomic from L2) at 13 % issue-active: four
    // groups of 32 are in flight per warp, all keys loaded before the first vote, all atomics issued before the first store.
    constexpr int kU = 4;
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * 256 + threadIdx.x) >> 5, nwarps = (gridDim.x * 256) >> 5;
    for (int base = warp * 32 * kU; base < a.in_count; base += nwarps * 32 * kU) {
        int k[kU], at[kU], rank[kU];
#pragma unroll
        for (int u = 0; u < kU; u++) {
            const int i = base + 32 * u + lane;
            k[u] = i < a.in_count ? a.skey[i] : -1;
        }
#pragma unroll
        for (int u = 0; u < kU; u++) {
            const unsigned act = __ballot_sync(kFull, k[u] >= 0);
            at[u] = 0; rank[u] = 0;
            if (k[u] >= 0) {
                const unsigned peers = __match_any_sync(act, k[u]);
                const int leader = __ffs(peers) - 1;
                if (lane == leader) at[u] = atomicAdd(a.lhist + sel_bin(k[u], base + 32 * u), __popc(peers));
                rank[u] = (leader << 8) | __popc(peers & ((1u << lane) - 1u));
            }
        }
#pragma unroll
        for (int u = 0; u < kU; u++) {
            // every lane takes part in the shuffle; only the keyed ones use what they get
            const int base_at = __shfl_sync(kFull, at[u], rank[u] >> 8);
            if (k[u] >= 0) a.sorted[base_at + (rank[u] & 255)] = base + 32 * u + lane;
        }
    }
}

// One ticket = 32 sorted entries = 32 / kSelV groups: all lanes rebuild one vertex each, then the groups are swept one
// after the other.  With `cdf_stride` > 0 (dynamic shared memory, kSelV x cdf_stride floats per warp) the sweep leaves the
// running sums of every lane's column behind, and lane v picks vertex v's light by bisection of its column -- one weight
// re-evaluated per vertex instead of a second warp-wide pass (warp_light_pick) per vertex.
#ifndef MCPT_SORTED_MINBLK
#define MCPT_SORTED_MINBLK (kSelV == 4 ? 6 : MCPT_SEL_MINBLK)
#endif
template <int STRAT>
__global__ void __launch_bounds__(kBlk, MCPT_SORTED_MINBLK) k_lightsel_sorted(const BounceArgs a, int* cursor, int cdf_stride)
{
    extern __shared__ float sm_sel[];
    const DScene& s = a.s;
    const int lane = threadIdx.x & 31;
    float* cdf = sm_sel + (threadIdx.x >> 5) * (kSelV * cdf_stride);
    const int ngroups = *a.ngroups;
    const int ntickets = (ngroups * kSelV + 31) >> 5;
    const bool rr_forced = a.max_depth > 0 && a.bounce + 1 >= a.max_depth;
    unsigned c_evals = 0;  // in units of 32
    int next = 0;
    if (lane == 0) next = atomicAdd(cursor, 1);
    for (;;) {
        const int t = __shfl_sync(kFull, next, 0);
        if (t >= ntickets) break;
        if (lane == 0) next = atomicAdd(cursor, 1);
        const int e = 32 * t + lane;
        int i = e < ngroups * kSelV ? a.sorted[e] : -1;
        f3 p = mk<float>(0.f, 0.f, 0.f), N = mk<float>(0.f, 0.f, 1.f);
        float u = 0.f;
        int prim = 0, lid = 0;
        if (i >= 0) {
            if (sel_vertex<STRAT>(a, i, rr_forced, true, &p, &N, &u, &prim)) lid = __ldg(s.prim_list + prim);
            else i = -1;  // cannot happen: k_lsel_classify took the same decisions
        }
#pragma unroll 1
        for (int g0 = 0; g0 < 32; g0 += kSelV) {
            const int i0 = __shfl_sync(kFull, i, g0);  // the first entry of a group always exists (k_lsel_offsets)
            if (i0 < 0) continue;
            const int2 rg = __ldg(s.lrange + __shfl_sync(kFull, lid, g0));
            const uint32_t* list = s.llist + rg.x;
            f3 X[kSelV], Nn[kSelV];
            float uu[kSelV];
            int ii[kSelV];
            int nvalid = 0;
#pragma unroll
            for (int v = 0; v < kSelV; v++) {
                ii[v] = __shfl_sync(kFull, i, g0 + v);
                const int src = ii[v] >= 0 ? g0 + v : g0;
                X[v] = mk<float>(__shfl_sync(kFull, p.x, src), __shfl_sync(kFull, p.y, src), __shfl_sync(kFull, p.z, src));
                Nn[v] = mk<float>(__shfl_sync(kFull, N.x, src), __shfl_sync(kFull, N.y, src), __shfl_sync(kFull, N.z, src));
                uu[v] = __shfl_sync(kFull, u, src);
                nvalid += ii[v] >= 0;
            }
            float part[kSelV];
            warp_light_partial_list<kSelV, kCdfSub>(s, list, rg.y, X, Nn, part, cdf_stride ? cdf : nullptr, cdf_stride);
            c_evals += (unsigned)(rg.y >> 5) * (unsigned)nvalid;
            if (cdf_stride) {
                // lanes kCdfSub v .. kCdfSub v + kCdfSub - 1 resolve vertex v: lane totals -> lane column j, bisection of the
                // column's checkpoints -> one group of kCdfSub list positions, whose weights the lanes re-evaluate (one each,
                // bit-identical to the sweep's) and the first lane of the group walks
                __syncwarp();
                const int K = rg.y >> 5, M = K / kCdfSub;
                int my_i = -1, my_j = 0;
                float my_rem = 0.f, my_total = 0.f;
                f3 mX = X[0], mN = Nn[0];
#pragma unroll
                for (int v = 0; v < kSelV; v++) {
                    const float incl = warp_scan_incl(part[v]);
                    const float total = __shfl_sync(kFull, incl, 31);
                    const float target = fmaxf(uu[v], 1e-7f) * total;
                    const unsigned nz = __ballot_sync(kFull, part[v] > 0.f);
                    const unsigned ge = __ballot_sync(kFull, incl >= target && part[v] > 0.f);
                    const int j = ge ? (__ffs(ge) - 1) : (nz ? 31 - __clz(nz) : 0);
                    const float rem = target - (__shfl_sync(kFull, incl, j) - __shfl_sync(kFull, part[v], j));
                    if (lane / kCdfSub == v) { my_i = ii[v]; my_j = j; my_rem = rem; my_total = total; mX = X[v]; mN = Nn[v]; }
                }
                const int t = lane & (kCdfSub - 1);
                const bool mine = lane < kSelV * kCdfSub && my_i >= 0 && my_total > 0.f && K > 0;
                float w = 0.f, start = 0.f;
                int off = 0;
                if (mine) {
                    const float* col = cdf + (lane / kCdfSub) * cdf_stride + my_j;  // col[32 m]: running sum after the lane's position kCdfSub m + kCdfSub - 1
                    int lo = 0, hi = M;                                             // first checkpoint >= rem; M: the tail group
                    while (lo < hi) {
                        const int mid = (lo + hi) >> 1;
                        if (col[32 * mid] < my_rem) lo = mid + 1; else hi = mid;
                    }
                    start = lo ? col[32 * (lo - 1)] : 0.f;
                    const int it = kCdfSub * lo + t;
                    if (it < K) {
                        off = (int)__ldg(list + my_j + 32 * it);
                        w = light_weight_at(s, off, mX, mN);
                        if (!(w > 0.f) || lsoa_light(off) >= s.L) w = 0.f;
                    }
                }
                // the first lane of each group collects the group's weights (all lanes shuffle) and walks them in list order
                float wg[kCdfSub];
                int og[kCdfSub];
#pragma unroll
                for (int q = 0; q < kCdfSub; q++) { wg[q] = __shfl_down_sync(kFull, w, q); og[q] = __shfl_down_sync(kFull, off, q); }
                if (lane < kSelV * kCdfSub && t == 0 && my_i >= 0) {
                    int light = -1;
                    float wsel = 0.f, acc = start;
                    bool found = false;
#pragma unroll
                    for (int q = 0; q < kCdfSub; q++) {
                        if (wg[q] > 0.f && !found) {
                            acc += wg[q];
                            light = lsoa_light(og[q]); wsel = wg[q];  // the last survivor so far: the fallback against rounding
                            found = acc >= my_rem;
                        }
                    }
                    if (light < 0 && mine) {
                        // rounding put the target past the last positive weight of the column's tail: take the column's last
                        // survivor (rare: a backwards scan of the lane's positions)
                        for (int it = K - 1; it >= 0 && light < 0; it--) {
                            const int o2 = (int)__ldg(list + my_j + 32 * it);
                            const float w2 = light_weight_at(s, o2, mX, mN);
                            if (w2 > 0.f && lsoa_light(o2) < s.L) { light = lsoa_light(o2); wsel = w2; }
                        }
                    }
                    a.sel4[my_i] = make_float4(__int_as_float(light), wsel, my_total, 0.f);
                }
                __syncwarp();  // the next group overwrites the columns
            } else {
#pragma unroll
                for (int v = 0; v < kSelV; v++) {
                    if (ii[v] < 0) continue;
                    const float incl = warp_scan_incl(part[v]);
                    const float total = __shfl_sync(kFull, incl, 31);
                    const LightPick pk = warp_light_pick<true>(s, list, rg.y, X[v], Nn[v], part[v], incl, total, uu[v]);
                    if (lane == 0) a.sel4[ii[v]] = make_float4(__int_as_float(pk.light), pk.w, total, 0.f);
                }
            }
        }
    }
    if (lane == 0 && c_evals) atomicAdd(a.counters + CNT_EVALS, (unsigned long long)c_evals * 32);
}

// ---- light selection through the light hierarchy (MCPT_LIGHT_TREE) ------------------------------------------------
// Stands in for k_lightsel at every vertex after the first (the first vertex of a pixel keeps the exact selection: its
// O(L) sweep is shared by all samples of the pass, k_lightprep, and costs no more than a tree walk per sample).  Each
// thread rebuilds the vertex of one queue entry and repeats the roulette / uniform draws exactly like k_lightsel; the
// few entries that do sample a light (about 1 in 9 at the second bounce of veach-mis) are collected in a shared-memory
// queue, and whenever 256 of them are waiting the whole block walks the tree of light_tree.h for them (lt_pick,
// lights.cuh: O(log L) dependent node pairs per vertex instead of the O(L) sweep), one vertex per thread, so the walks
// run at full warps.  Leaves sel4[i] = (light, its spherical-triangle weight w, w / P(light)): k_shade's
// pl = Lsum / sel4.z is then the solid-angle pdf P(light) / sA of the direction it draws.
template <int STRAT>
__global__ void __launch_bounds__(256) k_lighttree(const BounceArgs a)
{
    constexpr bool kTree = (STRAT == MCPT_MIS);
    __shared__ float4 q_xi[512], q_nu[512];  // vertex | queue index bits,  normal | uniform
    __shared__ int q_n;
    const DScene& s = a.s;
    const bool rr_forced = a.max_depth > 0 && a.bounce + 1 >= a.max_depth;
    const int tid = threadIdx.x, lane = tid & 31;
    const int ntiles = (a.in_count + 255) >> 8;
    if (tid == 0) q_n = 0;
    __syncthreads();
    auto flush = [&](int n) {  // walks for the top min(n, 256) waiting vertices; returns how many are left
        const int m = min(n, 256);
        if (tid < m) {
            const float4 xi = q_xi[n - m + tid], nu = q_nu[n - m + tid];
            const f3 X = to_f3(xi), Nn = to_f3(nu);
            float prob = 0.f;
            const int l = lt_pick(s, X, Nn, nu.w, &prob);
            const float w = light_weight_fast(s, l, X, Nn);
            float4 out = make_float4(__int_as_float(-1), 0.f, 0.f, 0.f);
            if (w > 0.f && prob > 0.f) out = make_float4(__int_as_float(l), w, w / prob, 0.f);
            a.sel4[__float_as_int(xi.w)] = out;
        }
        __syncthreads();
        if (tid == 0) q_n = n - m;
        __syncthreads();
        return n - m;
    };
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int i = (tile << 8) + tid;
        bool need = false;
        f3 p = mk<float>(0.f, 0.f, 0.f), N = mk<float>(0.f, 0.f, 1.f);
        float u = 0.f;
        if (i < a.in_count) {
            const float4 qh = a.hit[i];
            const int prim = __float_as_int(qh.x);
            if (prim >= 0) {
                const float4 qd = a.in.dir[i];
                const Vertex v = make_vertex(s, prim, qh.y, qh.z, -to_f3(qd));
                if (v.front && v.light < 0) {
                    const int slot = __float_as_int(qd.w);
                    const uint32_t pixel = (uint32_t)(slot % a.npix), sample = (uint32_t)(a.sample_base + slot / a.npix);
                    uint64_t node = (uint64_t)a.bounce + 1;
                    if (kTree) { const uint2 nd = a.in.node[i]; node = ((uint64_t)nd.y << 32) | nd.x; }
                    need = true;
                    if (kTree || STRAT == MCPT_MIS_SINGLE) need = !rr_forced && !(rng_draw4(a.seed, pixel, sample, node, 0).u[0] > a.p_rr);
                    if (need) { u = rng_draw4(a.seed, pixel, sample, node, 1).u[0]; p = v.p; N = v.N; }
                }
            }
            if (!need) a.sel4[i] = make_float4(__int_as_float(-1), 0.f, 0.f, 0.f);
        }
        const unsigned m = __ballot_sync(kFull, need);
        if (m) {
            int base = 0;
            if (lane == 0) base = atomicAdd(&q_n, __popc(m));
            base = __shfl_sync(kFull, base, 0);
            if (need) {
                const int at = base + __popc(m & ((1u << lane) - 1u));
                q_xi[at] = make_float4(p.x, p.y, p.z, __int_as_float(i));
                q_nu[at] = make_float4(N.x, N.y, N.z, u);
            }
        }
        __syncthreads();
        const int n = q_n;  // < 512: at most 255 were waiting before this tile
        __syncthreads();    // everyone has read q_n before the next tile's pushes move it
        if (n >= 256) flush(n);
    }
    __syncthreads();
    for (int n = q_n; n > 0;) n = flush(n);
}

// ---- one bounce --------------------------------------------------------------------------------------------
// ray from a vertex to its sampled light point: distance along wl (Mylight.cpp:311 re-intersects the same way)
template <bool EXACT>
__device__ __forceinline__ bool light_point_test(f3 p, f3 wl, f3 v0, f3 v1, f3 v2, float tmin, float* be, float* ga, float* tl)
{
    if (!EXACT) return tri_test_fast(p, wl, v0, v1, v2, tmin, be, ga, tl);
    double b, g, t;
    if (!tri_test_exact(to_d3(p), to_d3(wl), to_d3(v0), to_d3(v1), to_d3(v2), &b, &g, &t)) return false;
    *be = (float)b; *ga = (float)g; *tl = (float)t;
    return true;
}

struct Child { f3 o, d, thr; int exclude; uint64_t node; float misw; };

// k_shade: everything shade* does at one vertex except the scene-BVH traversals, which the wavefront hands to
// k_extend (closest hit of the children) and k_shadow (NEE visibility).  FIRST = primary vertices without jitter:
// the ray comes from the camera formula and the hit from k_primary's per-pixel table.
// A warp owns a chunk of kChunk = 256 queue entries and works on it in two phases.  Phase A (every entry, cheap): hit
// record, vertex, back-face / emission classification, the accumulator write, the roulette taken before sampling
// (MIS / BRDF-only); the entries that go on to sample -- 60 % at the first vertex of veach-mis, about 1 in 9 at the second --
// are listed, in order, in shared memory.  Phase B (32 listed entries at a time, all lanes busy): light sampling, BRDF
// sampling, the MIS light-only ray, emission of the children into the chunk's staging window.  The list keeps the input
// order, so the children land exactly where the one-phase kernel put them (it ran 6.9 threads per instruction at the
// later bounces, 17.8 at the first).
constexpr int kChunk = 256;  // queue entries per warp ticket / staging window (128: veach-mis 1378 instead of 1399 Msamples/s)
#ifndef MCPT_SHADE_MINBLK
#define MCPT_SHADE_MINBLK 6
#endif
// EXACT (MCPT_LEAF_EXACT): the two ray/triangle tests k_shade makes itself -- the sampled light point's distance and the
// MIS light-only ray -- go through the reference's FP64 test like every other ray of the render.
#ifndef MCPT_SHADE_PREFETCH
#define MCPT_SHADE_PREFETCH 1
#endif
#ifndef MCPT_PREFETCH_AHEAD
#define MCPT_PREFETCH_AHEAD 512
#endif
constexpr int kPrefetchAhead = MCPT_PREFETCH_AHEAD;
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

template <int STRAT, bool FIRST, bool EXACT>
__global__ void __launch_bounds__(kBlk, (STRAT == MCPT_MIS && !EXACT) ? 5 : MCPT_SHADE_MINBLK) k_shade(const BounceArgs a)
{
    constexpr bool kTree = (STRAT == MCPT_MIS);
    constexpr bool kNee = (STRAT == MCPT_NEE_SPHERICAL || STRAT == MCPT_NEE_UNIFORM);
    constexpr bool kSingle = (STRAT == MCPT_MIS_SINGLE);  // single-path MIS: NEE's visibility ray + MIS weights
    constexpr bool kShadow = kNee || kSingle;              // strategies that queue a visibility ray per vertex
    constexpr bool kSph = (STRAT == MCPT_NEE_SPHERICAL || STRAT == MCPT_MIS || kSingle);
    __shared__ int stack[kStackSm * kBlk];
    __shared__ int s_list[kBlk / 32][kChunk];
    int* list = s_list[threadIdx.x >> 5];
    const DScene& s = a.s;
    const int lane = threadIdx.x & 31;
    const unsigned below = (1u << lane) - 1u;
    unsigned c_vert = 0, c_lrays = 0, c_phits = 0, c_sweeps = 0;
    const int nchunks = (a.in_count + kChunk - 1) / kChunk;
    const int depth = a.bounce + 1;
    const bool rr_forced = a.max_depth > 0 && depth >= a.max_depth;

    // visibility-ray queue: the warp reserves kSqGrab slots at a time (one atomic on the queue's one counter per kSqGrab
    // records instead of one per 32 entries) and hands them out batch by batch; what is left at the end is marked as holes
    int sq_rb = 0, sq_re = 0;
    int next = 0;
    if (lane == 0) next = atomicAdd(a.ticket, 1);
    for (;;) {
        const int chunk = __shfl_sync(kFull, next, 0);
        if (chunk >= nchunks) break;
        if (lane == 0) next = atomicAdd(a.ticket, 1);  // prefetch the next ticket behind this chunk's work
#if MCPT_SHADE_PREFETCH
        if (!FIRST) {
            // phase A is a dependent chain per 32 entries (queue record -> vertex gather -> accumulator) at 20 warps per SM:
            // bound by DRAM latency (ncu: 10 of 15 warp-cycles per issue on the long scoreboard).  Whoever takes the chunk
            // kPrefetchAhead tickets from here (a few microseconds away) finds its queue records in L2: one line per lane and array.
            const int pc = chunk + kPrefetchAhead;
            if (pc < nchunks) {
                const size_t e = (size_t)pc * kChunk + 8 * (size_t)lane;
                prefetch_l2(a.in.dir + e); prefetch_l2(a.hit + e); prefetch_l2(a.in.thr + e);
                if (kTree && lane < 16) prefetch_l2(a.in.node + (size_t)pc * kChunk + 16 * (size_t)lane);
            }
        }
#endif

        // ==== phase A: classify every entry, write its accumulator, list the ones that sample ===========================
        int n_list = 0;
#pragma unroll 1
        for (int sub = 0; sub < kChunk / 32; sub++) {
            const int i = chunk * kChunk + sub * 32 + lane;
            bool keep = false;
            if (i < a.in_count) {
                // shade prologue: p, N, back-face, emission (main.cpp:273-288 / 351-366 / 406-421)
                int slot = i, prim;
                float hb, hg;
                f3 wo;
                uint32_t pixel;
                if (FIRST) {
                    pixel = (uint32_t)(i % a.npix);
                    const float4 ph = __ldg(a.primary_hit + pixel);
                    prim = __float_as_int(ph.x); hb = ph.y; hg = ph.z;
                    wo = -camera_dir(a.cam, (int)pixel, 0.f, 0.f);
                    if (prim >= 0) c_phits++;
                } else {
                    const float4 qd = a.in.dir[i], qh = a.hit[i];
                    wo = -to_f3(qd);
                    slot = __float_as_int(qd.w);
                    pixel = (uint32_t)(slot % a.npix);
                    prim = __float_as_int(qh.x); hb = qh.y; hg = qh.z;
                    if (a.bounce == 0 && prim >= 0) c_phits++;  // jittered camera rays
                }
                f3 contrib = mk<float>(0.f, 0.f, 0.f);
                if (prim >= 0) {
                    const Vertex v = make_vertex(s, prim, hb, hg, wo);
                    if (v.front) {
                        if (v.light >= 0) {  // emissive: self-emission only
                            // shade(): an indirect ray that lands on a light contributes 0 (main.cpp:338); the emission
                            // return of main.cpp:284-288 is only reachable from the camera ray.
                            if (!kNee || a.bounce == 0) {
                                f3 thr = mk<float>(1.f, 1.f, 1.f);
                                float misw = 1.f;  // single-path MIS: weight of emission found by this ray (1 for camera rays)
                                if (!FIRST) {
                                    const float4 qt = a.in.thr[i];
                                    thr = to_f3(qt);
                                    if (kSingle && a.bounce > 0) misw = qt.w;
                                }
                                contrib = thr * to_f3(__ldg(s.lights + 5 * v.light + 4)) * misw;
                            }
                        } else {
                            keep = true;
                            c_vert++;
                            if (!kNee) {  // MIS / BRDF-only: roulette before any sampling (main.cpp:375-383 / 429-437)
                                uint64_t node = FIRST ? 1 : (uint64_t)a.bounce + 1;
                                if (kTree && !FIRST) { const uint2 nd = a.in.node[i]; node = ((uint64_t)nd.y << 32) | nd.x; }
                                const uint32_t sample = (uint32_t)(a.sample_base + slot / a.npix);
                                keep = !rr_forced && !(rng_draw4(a.seed, pixel, sample, node, 0).u[0] > a.p_rr);
                            }
                        }
                    }
                }
                // accumulate
                if (FIRST) a.Lacc[slot] = make_float4(contrib.x, contrib.y, contrib.z, 0.f);
                else if (kTree) a.contrib[i] = make_float4(contrib.x, contrib.y, contrib.z, __int_as_float(slot));
                else if (contrib.x != 0.f || contrib.y != 0.f || contrib.z != 0.f) {
                    float4 L = a.Lacc[slot];
                    L.x += contrib.x; L.y += contrib.y; L.z += contrib.z;
                    a.Lacc[slot] = L;
                }
            }
            const unsigned mk_ = __ballot_sync(kFull, keep);
            if (keep) list[n_list + __popc(mk_ & below)] = i;
            n_list += __popc(mk_);
        }
        __syncwarp();
        // the listed entries of a later bounce are exactly the vertices k_lightsel swept the light table for (warp-uniform count)
        if (kSph && !FIRST && !(a.flags & MCPT_LIGHT_TREE)) c_sweeps += (unsigned)n_list;

        // ==== phase B: the listed entries, 32 at a time ===================================================================
        int emitted = 0;
#pragma unroll 1
        for (int b0 = 0; b0 < n_list; b0 += 32) {
        const bool valid = b0 + lane < n_list;
        const int i = valid ? list[b0 + lane] : 0;

        // -- A. ray + path state, closest hit (extend) ------------------------------------------------
        f3 o = mk<float>(0.f, 0.f, 0.f), d = mk<float>(0.f, 0.f, 1.f), thr = mk<float>(1.f, 1.f, 1.f);
        int slot = 0;
        uint64_t node = 1;
        Hit h;
        h.prim = -1; h.b = h.g = h.t = 0.f;
        uint32_t pixel = 0, sample = 0;
        if (valid) {
            if (FIRST) {
                slot = i;
                pixel = (uint32_t)(i % a.npix);
                sample = (uint32_t)(a.sample_base + i / a.npix);
                o = mk<float>(a.cam.eye[0], a.cam.eye[1], a.cam.eye[2]);
                d = camera_dir(a.cam, (int)pixel, 0.f, 0.f);
                const float4 ph = __ldg(a.primary_hit + pixel);
                h.prim = __float_as_int(ph.x); h.b = ph.y; h.g = ph.z; h.t = ph.w;
            } else {
                const float4 qo = a.in.org[i], qd = a.in.dir[i], qt = a.in.thr[i];
                o = to_f3(qo); d = to_f3(qd); thr = to_f3(qt);
                slot = __float_as_int(qd.w);
                pixel = (uint32_t)(slot % a.npix);
                sample = (uint32_t)(a.sample_base + slot / a.npix);
                if (kTree) { const uint2 nd = a.in.node[i]; node = ((uint64_t)nd.y << 32) | nd.x; }
                else node = (uint64_t)a.bounce + 1;
                const float4 qh = a.hit[i];
                h.prim = __float_as_int(qh.x); h.b = qh.y; h.g = qh.z; h.t = qh.w;
            }
        }

        // -- B. the vertex again (phase A established: front-facing, not emissive, roulette survived) -------------------
        bool alive = valid;
        f3 p = o, N = d, wo = -d;
        PhongMat<float> mat;
        mat.kd = mat.ks = mk<float>(0.f, 0.f, 0.f); mat.ns = 1.f;
        if (alive) {
            const Vertex v = make_vertex(s, h.prim, h.b, h.g, wo);
            p = v.p; N = v.N;
            mat = load_mat(s, v.mat);
        }
        Rng4 r0, r1;
        r0.u[0] = r0.u[1] = r0.u[2] = r0.u[3] = 0.5f;
        r1 = r0;
        if (alive) { r0 = rng_draw4(a.seed, pixel, sample, node, 0); if (STRAT != MCPT_BRDF_ONLY) r1 = rng_draw4(a.seed, pixel, sample, node, 1); }

        // NEE: the roulette comes after the direct term (main.cpp:321-329); the others took it in phase A
        const bool rr_ok = alive && !rr_forced && !(r0.u[0] > a.p_rr);

        // -- C. light sampling ---------------------------------------------------------------------------
        Child ch[2];
        int n_out = 0;
        float wsum = 0.f;
        // NEE visibility ray (main.cpp:309-311): visible iff the closest hit along wl IS the sampled triangle
        bool sh_need = false;
        f3 sh_d = d, sh_c = mk<float>(0.f, 0.f, 0.f);
        float sh_t = 0.f;
        int sh_prim = -1;
        if (kSph) {
            int chosen = -1; float wch = 0.f;
            if (alive) {
                if (FIRST) {  // drawn once per pixel by k_lightprep
                    const int2 v = a.sel[slot];
                    chosen = v.x; wch = __int_as_float(v.y);
                    wsum = a.wsum_pix[pixel];
                } else if (!a.by_lists || a.skey[i] >= 0) {  // drawn by k_lightsel / k_lightsel_sorted / k_lighttree
                    const float4 v = a.sel4[i];
                    chosen = __float_as_int(v.x); wch = v.y; wsum = v.z;
                }
            }
            if (alive && chosen >= 0) {
                // Arvo sample of the chosen triangle (Mylight.cpp:285-317) through arvo_direct (lights.cuh), in FP64:
                // the q of :292 is a ratio of two O(sA^2) differences, and FP32 misses the 1e-5 direction bar even on
                // large lights (measured 5e-5 on the Cornell ceiling light)
                const float4 l0 = __ldg(s.lights + 5 * chosen), l1 = __ldg(s.lights + 5 * chosen + 1), l2 = __ldg(s.lights + 5 * chosen + 2);
                const float4 l3 = __ldg(s.lights + 5 * chosen + 3);
                const float lsum_f = l3.w;
                const int lprim = __float_as_int(l0.w);
                const f3 Le = to_f3(__ldg(s.lights + 5 * chosen + 4)), ngl = to_f3(l3), lv0 = to_f3(l0);
                d3 wd = to_d3(N);
                const bool ok = arvo_direct<double>(to_d3(l0), to_d3(l1), to_d3(l2), to_d3(p), to_d3(N), (double)(wch / lsum_f),
                                                    (double)r1.u[1], (double)r1.u[2], &wd);
                const f3 wl = to_f3(wd);
                const float pl = lsum_f / wsum;  // solid-angle pdf, Mylight.cpp:317
                const float cosN = dot(wl, N);
                if (ok && cosN > 0.f) {
                    if (STRAT == MCPT_MIS) {
                        // light-sample branch (main.cpp:453-466): the ray continues into whatever it hits
                        const float pp = phong_pdf<float>(N, wl, wo, mat);
                        const f3 f = phong_eval<float>(N, wl, wo, mat);
                        Child& c = ch[n_out++];
                        c.o = p; c.d = wl; c.exclude = h.prim; c.node = node * 2;
                        c.thr = thr * f * (cosN / (pl + pp) / a.p_rr);
                        c.misw = 0.f;
                    } else {
                        // NEE (main.cpp:300-316): visible iff the closest hit along wl IS the sampled triangle
                        const float cosL = -dot(wl, ngl);
                        if (cosL > 0.f) {
                            const float4 q1 = __ldg(s.lights + 5 * chosen + 1), q2 = __ldg(s.lights + 5 * chosen + 2);
                            float be, ga, tl;
                            if (light_point_test<EXACT>(p, wl, lv0, to_f3(q1), to_f3(q2), s.tmin, &be, &ga, &tl)) {
                                const f3 f = phong_eval<float>(N, wl, wo, mat);
                                float g = (a.flags & MCPT_FIX_NEE_MEASURE) ? cosN / pl : cosN * cosL / (tl * tl) / pl;  // F9
                                // single-path MIS: solid-angle estimator with the balance weight pl / (pl + pp), after the roulette
                                if (kSingle) g = cosN / (pl + phong_pdf<float>(N, wl, wo, mat)) / a.p_rr;
                                sh_need = true; sh_d = wl; sh_t = tl; sh_prim = lprim;
                                sh_c = thr * Le * f * g;
                            }
                        }
                    }
                }
            }
        } else if (STRAT == MCPT_NEE_UNIFORM) {
            if (alive) {
                LightPoint<float> lp = light_sample_uniform_fast(s, r1.u[0], r1.u[1], r1.u[2], r1.u[3]);
                const f3 dl = lp.coord - p;
                const float d2 = dot(dl, dl);
                const f3 wl = dl * rsqrtf(d2);
                const f3 ngl = to_f3(__ldg(s.lights + 5 * lp.light + 3));
                const float cosN = dot(wl, N), cosL = -dot(wl, ngl);
                if (cosN > 0.f && cosL > 0.f) {
                    const float4 q0 = __ldg(s.lights + 5 * lp.light), q1 = __ldg(s.lights + 5 * lp.light + 1), q2 = __ldg(s.lights + 5 * lp.light + 2);
                    float be, ga, tl;
                    if (light_point_test<EXACT>(p, wl, to_f3(q0), to_f3(q1), to_f3(q2), s.tmin, &be, &ga, &tl)) {
                        const f3 f = phong_eval<float>(N, wl, wo, mat);
                        sh_need = true; sh_d = wl; sh_t = tl; sh_prim = lp.prim;
                        sh_c = thr * lp.Le * f * (cosN * cosL / d2 / lp.prob);  // main.cpp:314
                    }
                }
            }
        }

        // NEE: roulette after the direct term (main.cpp:321-329)
        if (kNee && !rr_ok) alive = false;

        // -- D. BRDF sample (main.cpp:331-341 / 385-395 / 471-491) ---------------------------------------
        if (alive) {
            float pdf;
            const f3 wi = phong_sample_fast(N, wo, mat, r0.u[1], r0.u[2], r0.u[3], &pdf);
            const float cosN = dot(wi, N);
            if (!(cosN < 0.f) && pdf > 0.f) {
                // F8; single-path MIS is not a reference estimator and is only unbiased with the mixture pdf: always use it
                if (kSingle || (a.flags & MCPT_FIX_MIXTURE_PDF)) pdf = phong_pdf<float>(N, wi, wo, mat);
                float lpdf = 0.f;
                const bool ltree = !FIRST && (a.flags & MCPT_LIGHT_TREE) != 0;  // the first vertex keeps the exact selection
                if ((STRAT == MCPT_MIS || kSingle) && (wsum > 0.f || ltree)) {
                    // pdf of wi under light sampling (main.cpp:484-489): nearest EMISSIVE triangle along wi, occluders
                    // ignored; it counts iff it survived this vertex's culls (Mylight.cpp:484-493, evaluated for the
                    // current vertex -- F10 not reproduced)
                    Hit lh;
                    if (EXACT) {
                        const HitD hx = trace_exact<false>(s.lnodes, s.ltris, to_d3(p), to_d3(wi), h.prim, 1.7976931348623157e308, stack + threadIdx.x, kBlk);
                        lh.prim = hx.prim; lh.b = (float)hx.b; lh.g = (float)hx.g; lh.t = (float)hx.t;
                    } else lh = trace_fast<false>(s.lnodes, s.ltris, p, wi, h.prim, s.tmin, 3.0e38f, stack + threadIdx.x, kBlk);
                    c_lrays++;
                    if (lh.prim >= 0) {
                        const int li = __float_as_int(__ldg(s.pverts + 3 * lh.prim + 1).w);
                        const float wli = light_weight_fast(s, li, p, N);
                        // exact selection: Lsum / weights_sum (Mylight.cpp:484-493); light tree: P(li | vertex) / sA(li)
                        if (wli > 0.f) lpdf = __ldg(s.lights + 5 * li + 3).w * (ltree ? lt_prob(s, li, p, N) / wli : 1.f / wsum);
                    }
                }
                const f3 f = phong_eval<float>(N, wi, wo, mat);
                Child& c = ch[n_out++];
                c.o = p; c.d = wi; c.exclude = h.prim; c.node = kTree ? node * 2 + 1 : 0;
                c.thr = thr * f * (cosN / (pdf + lpdf) / a.p_rr);
                c.misw = 0.f;
                if (kSingle) {  // reflected light keeps weight 1; only emission found by this ray takes the balance weight
                    c.thr = thr * f * (cosN / pdf / a.p_rr);
                    c.misw = pdf / (pdf + lpdf);
                }
            }
        }

        // -- E. emit children into the chunk's staging window, in list (= input) order ---------------------------------------
        {
            const unsigned m1 = __ballot_sync(kFull, n_out >= 1), m2 = __ballot_sync(kFull, n_out >= 2);
            const int off = chunk * (kTree ? 2 * kChunk : kChunk) + emitted + __popc(m1 & below) + __popc(m2 & below);
            for (int k = 0; k < n_out; k++) {
                const int dst = off + k;
                const Child& c = ch[k];
                a.out.org[dst] = make_float4(c.o.x, c.o.y, c.o.z, __int_as_float(c.exclude));
                a.out.dir[dst] = make_float4(c.d.x, c.d.y, c.d.z, __int_as_float(slot));
                a.out.thr[dst] = make_float4(c.thr.x, c.thr.y, c.thr.z, kSingle ? c.misw : 0.f);
                if (kTree) a.out.node[dst] = make_uint2((uint32_t)c.node, (uint32_t)(c.node >> 32));
            }
            emitted += __popc(m1) + __popc(m2);
        }
        if (kShadow) {
            // visibility rays: unordered append (every slot owns at most one per bounce, so k_shadow's
            // Lacc[slot] += contribution needs no ordering), one atomic per warp
            const unsigned ms = __ballot_sync(kFull, sh_need);
            if (ms) {
                const int m = __popc(ms), have = sq_re - sq_rb;
                int nb = 0;
                if (have < m) {  // the batch runs over the end of the reservation: the overflow goes to the head of a new one
                    if (lane == 0) nb = atomicAdd(a.sq.count, kSqGrab);
                    nb = __shfl_sync(kFull, nb, 0);
                }
                const int r = __popc(ms & ((1u << lane) - 1u));
                const int dst_ = r < have ? sq_rb + r : nb + (r - have);
                if (have < m) { sq_rb = nb + (m - have); sq_re = nb + kSqGrab; }
                else sq_rb += m;
                if (sh_need) {
                    const int dst = dst_;
                    a.sq.org[dst] = make_float4(p.x, p.y, p.z, sh_t);
                    a.sq.dir[dst] = make_float4(sh_d.x, sh_d.y, sh_d.z, __int_as_float(slot));
                    a.sq.ctr[dst] = make_float4(sh_c.x, sh_c.y, sh_c.z, 0.f);
                    a.sq.ids[dst] = make_int2(h.prim, sh_prim);
                }
            }
        }
        }  // batches of the list
        if (lane == 0) {
            a.counts[chunk] = emitted;
            if (emitted) atomicAdd(a.groupsum + chunk / kGroup, emitted);
        }
        __syncwarp();  // the next chunk overwrites the list
    }
    if (kShadow && sq_re > sq_rb) {  // the unused tail of the last reservation: holes k_shadow skips (light prim -2, tmax < 0)
        for (int j = sq_rb + lane; j < sq_re; j += 32) {
            a.sq.org[j] = make_float4(0.f, 0.f, 0.f, -1.f);
            a.sq.dir[j] = make_float4(0.f, 0.f, 1.f, __int_as_float(0));
            a.sq.ids[j] = make_int2(-2, -2);
        }
        if (lane == 0) atomicAdd(a.counters + CNT_SQ_HOLES, (unsigned long long)(sq_re - sq_rb));
    }
    // warp-level counter reduction: one atomic per counter per warp
    c_vert = __reduce_add_sync(kFull, c_vert);
    c_lrays = __reduce_add_sync(kFull, c_lrays); c_phits = __reduce_add_sync(kFull, c_phits);
    if (lane == 0) {
        if (c_vert) atomicAdd(a.counters + CNT_VERTICES, (unsigned long long)c_vert);
        if (c_lrays) atomicAdd(a.counters + CNT_LIGHTRAYS, (unsigned long long)c_lrays);
        if (c_phits) atomicAdd(a.counters + CNT_PRIMARY_HITS, (unsigned long long)c_phits);
        if (c_sweeps) atomicAdd(a.counters + CNT_SWEEPS, (unsigned long long)c_sweeps);
    }
}


// ---- wavefront trace kernels: persistent warps, dynamic ray refill (traverse.cuh: trace_stream) -------------------
// k_extend: closest hit of every ray of the dense queue (closet_ray_intersect, Myobj.cpp:334-474) -> hit[i]
__global__ void __launch_bounds__(kBlk) k_extend(DScene s, const float4* __restrict__ org, const float4* __restrict__ dir,
                                                 const int* __restrict__ n_ptr, int* counter, float4* hit)
{
    __shared__ int stack[kStackSm * kBlk];
    const int n = *n_ptr;
    trace_stream<false>(
        s.nodes4, s.tris, s.tmin, n, counter, stack + threadIdx.x, kBlk, s.refill4,
        [&](int i, f3& o, f3& d, int& exclude, float& tmax, int& init_prim) {
            const float4 qo = org[i], qd = dir[i];
            o = to_f3(qo); d = to_f3(qd);
            exclude = __float_as_int(qo.w); tmax = 3.0e38f; init_prim = -1;
        },
        [&](int i, const Hit& h) { hit[i] = make_float4(__int_as_float(h.prim), h.b, h.g, h.t); });
}

// the same over the compressed 8-wide BVH (traverse.cuh: trace_stream8), chosen per scene (DScene.use8)
// nine resident blocks (56 registers): soup-1m +0.7 %, cornell-box +1.5 % over the compiler's own 64; 6 / 7 / 10 blocks lose 2-9 %
__global__ void __launch_bounds__(kBlk, 9) k_extend8(DScene s, const float4* __restrict__ org, const float4* __restrict__ dir,
                                                  const int* __restrict__ n_ptr, int* counter, float4* hit)
{
    __shared__ uint2 stack[kStack8Sm * kBlk];
    const int n = *n_ptr;
    trace_stream8<false>(
        s.nodes8, s.tris8, s.tmin, n, counter, stack + threadIdx.x, kBlk, s.refill8,
        [&](int i, f3& o, f3& d, int& exclude, float& tmax, int& init_prim) {
            const float4 qo = org[i], qd = dir[i];
            o = to_f3(qo); d = to_f3(qd);
            exclude = __float_as_int(qo.w); tmax = 3.0e38f; init_prim = -1;
        },
        [&](int i, const Hit& h) { hit[i] = make_float4(__int_as_float(h.prim), h.b, h.g, h.t); });
}

__global__ void __launch_bounds__(kBlk, 9) k_shadow8(DScene s, const ShadowQueue sq, int* counter, float4* Lacc, unsigned long long* counters)
{
    __shared__ uint2 stack[kStack8Sm * kBlk];
    const int n = *sq.count;
    if (blockIdx.x == 0 && threadIdx.x == 0 && n) atomicAdd(counters + CNT_SHADOW, (unsigned long long)n);
    trace_stream8<true>(
        s.nodes8, s.tris8, s.tmin, n, counter, stack + threadIdx.x, kBlk, s.refill8,
        [&](int i, f3& o, f3& d, int& exclude, float& tmax, int& init_prim) {
            const float4 qo = sq.org[i], qd = sq.dir[i];
            const int2 id = sq.ids[i];
            o = to_f3(qo); d = to_f3(qd);
            exclude = id.x; tmax = qo.w; init_prim = id.y;
        },
        [&](int i, const Hit& h) {
            if (h.prim >= 0 && h.prim == sq.ids[i].y) {  // (holes of the queue carry light prim -2)
                const float4 c = sq.ctr[i];
                const int slot = __float_as_int(sq.dir[i].w);
                float4 L = Lacc[slot];
                L.x += c.x; L.y += c.y; L.z += c.z;
                Lacc[slot] = L;
            }
        });
}

// k_shadow: the visibility test of main.cpp:309-311 -- "the closest hit along wl IS the sampled light triangle" -- as
// an any-hit search for something that beats the light point under the closest-hit order; survivors add their
// pending contribution to the slot's accumulator (each slot has at most one visibility ray per bounce).
__global__ void __launch_bounds__(kBlk) k_shadow(DScene s, const ShadowQueue sq, int* counter, float4* Lacc, unsigned long long* counters)
{
    __shared__ int stack[kStackSm * kBlk];
    const int n = *sq.count;
    if (blockIdx.x == 0 && threadIdx.x == 0 && n) atomicAdd(counters + CNT_SHADOW, (unsigned long long)n);
    trace_stream<true>(
        s.nodes4, s.tris, s.tmin, n, counter, stack + threadIdx.x, kBlk, s.refill4,
        [&](int i, f3& o, f3& d, int& exclude, float& tmax, int& init_prim) {
            const float4 qo = sq.org[i], qd = sq.dir[i];
            const int2 id = sq.ids[i];
            o = to_f3(qo); d = to_f3(qd);
            exclude = id.x; tmax = qo.w; init_prim = id.y;
        },
        [&](int i, const Hit& h) {
            if (h.prim >= 0 && h.prim == sq.ids[i].y) {  // (holes of the queue carry light prim -2)
                const float4 c = sq.ctr[i];
                const int slot = __float_as_int(sq.dir[i].w);
                float4 L = Lacc[slot];
                L.x += c.x; L.y += c.y; L.z += c.z;
                Lacc[slot] = L;
            }
        });
}

// MCPT_LEAF_EXACT versions: one ray per thread through the binary BVH with FP64 slab tests and the reference's FP64
// Cramer triangle test (trace_exact) -- the hit ids are those of the reference's exhaustive search for the very ray the
// wavefront holds.  Visibility is the reference's own formulation: closest hit along wl, then compare ids.
__global__ void __launch_bounds__(kBlk) k_extend_exact(DScene s, const float4* __restrict__ org, const float4* __restrict__ dir,
                                                       const int* __restrict__ n_ptr, float4* hit)
{
    __shared__ int stack[kStackSm * kBlk];
    const int n = *n_ptr;
    for (int i = blockIdx.x * kBlk + threadIdx.x; i < n; i += gridDim.x * kBlk) {
        const float4 qo = org[i], qd = dir[i];
        const HitD h = trace_exact<false>(s.nodes, s.tris, to_d3(qo), to_d3(qd), __float_as_int(qo.w), 1.7976931348623157e308,
                                          stack + threadIdx.x, kBlk);
        hit[i] = make_float4(__int_as_float(h.prim), (float)h.b, (float)h.g, (float)h.t);
    }
}

__global__ void __launch_bounds__(kBlk) k_shadow_exact(DScene s, const ShadowQueue sq, float4* Lacc, unsigned long long* counters)
{
    __shared__ int stack[kStackSm * kBlk];
    const int n = *sq.count;
    if (blockIdx.x == 0 && threadIdx.x == 0 && n) atomicAdd(counters + CNT_SHADOW, (unsigned long long)n);
    for (int i = blockIdx.x * kBlk + threadIdx.x; i < n; i += gridDim.x * kBlk) {
        const float4 qo = sq.org[i], qd = sq.dir[i];
        const int2 id = sq.ids[i];
        if (id.y < 0) continue;  // a hole of the queue (k_shade's unused reservation)
        const HitD h = trace_exact<false>(s.nodes, s.tris, to_d3(qo), to_d3(qd), id.x, 1.7976931348623157e308, stack + threadIdx.x, kBlk);
        if (h.prim == id.y) {  // main.cpp:309-311
            const float4 c = sq.ctr[i];
            const int slot = __float_as_int(qd.w);
            float4 L = Lacc[slot];
            L.x += c.x; L.y += c.y; L.z += c.z;
            Lacc[slot] = L;
        }
    }
}

// jittered camera rays (extension F4): one ray per slot into the dense queue; the first k_extend traces them
__global__ void k_raygen(DCamera cam, int npix, int slots, int sample_base, uint64_t seed, Queue q, int tree, float4* Lacc, int* out_count)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) *out_count = slots;
    if (i >= slots) return;
    const uint32_t pixel = (uint32_t)(i % npix), sample = (uint32_t)(sample_base + i / npix);
    const Rng4 r = rng_draw4(seed, pixel, sample, 1, 2);
    const f3 d = camera_dir(cam, (int)pixel, r.u[0] - 0.5f, r.u[1] - 0.5f);
    q.org[i] = make_float4(cam.eye[0], cam.eye[1], cam.eye[2], __int_as_float(-1));
    q.dir[i] = make_float4(d.x, d.y, d.z, __int_as_float(i));
    q.thr[i] = make_float4(1.f, 1.f, 1.f, 0.f);
    if (tree) q.node[i] = make_uint2(1u, 0u);
    Lacc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}

// ---- MIS path tree: per-slot sums of per-element contributions ----------------------------------------------
// The compaction is order preserving and the initial queue is sorted by slot, so all elements of one slot
// are adjacent.  The head of each run owns the run: no two threads ever write the same slot.
__global__ void k_runsum(const float4* __restrict__ contrib, int n, float4* Lacc)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 c = contrib[i];
    const int slot = __float_as_int(c.w);
    if (i > 0 && __float_as_int(contrib[i - 1].w) == slot) return;  // not a run head
    float sx = c.x, sy = c.y, sz = c.z;
    for (int k = i + 1; k < n; k++) {
        const float4 e = contrib[k];
        if (__float_as_int(e.w) != slot) break;
        sx += e.x; sy += e.y; sz += e.z;
    }
    if (sx != 0.f || sy != 0.f || sz != 0.f) {
        float4 L = Lacc[slot];
        L.x += sx; L.y += sy; L.z += sz;
        Lacc[slot] = L;
    }
}

// film accumulation (main.cpp:566,575-576,582): the pixel's owner thread adds its samples of this pass to the running
// sum ONE BY ONE, in sample order -- the reference's `sum = sum + I` -- with Kahan compensation carried in `comp` from
// pass to pass.  The sequence of operations does not depend on where the pass boundaries fall, so a film is bit-identical
// for every batch_spp (and across an overflow retry), and 1024+ spp lose nothing to FP32 accumulation.
__global__ void k_accumulate(const float4* __restrict__ Lacc, int npix, int batch_spp, float4* image, float4* comp, int first_pass)
{
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npix) return;
    float4 s = first_pass ? make_float4(0.f, 0.f, 0.f, 0.f) : image[p];
    float4 c = first_pass ? make_float4(0.f, 0.f, 0.f, 0.f) : comp[p];
    for (int k = 0; k < batch_spp; k++) {
        const float4 v = Lacc[(size_t)k * npix + p];
        float y, t;
        y = __fsub_rn(v.x, c.x); t = __fadd_rn(s.x, y); c.x = __fsub_rn(__fsub_rn(t, s.x), y); s.x = t;
        y = __fsub_rn(v.y, c.y); t = __fadd_rn(s.y, y); c.y = __fsub_rn(__fsub_rn(t, s.y), y); s.y = t;
        y = __fsub_rn(v.z, c.z); t = __fadd_rn(s.z, y); c.z = __fsub_rn(__fsub_rn(t, s.z), y); s.z = t;
    }
    image[p] = s;
    comp[p] = c;
}

__global__ void k_finalize(const float4* __restrict__ image, int npix, float inv_spp, int accumulate, float* out_rgb)
{
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npix) return;
    const float4 v = image[p];
    float r = v.x * inv_spp, g = v.y * inv_spp, b = v.z * inv_spp;
    if (accumulate) { r += out_rgb[3 * p]; g += out_rgb[3 * p + 1]; b += out_rgb[3 * p + 2]; }
    out_rgb[3 * p] = r; out_rgb[3 * p + 1] = g; out_rgb[3 * p + 2] = b;
}

// RadianceRGB::tone_mapping, RadianceRGB.cpp:51-67
__global__ void k_tonemap(const float* __restrict__ rgb, size_t n3, double A, double gamma, uint8_t* out)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n3) return;
    double r = A * pow((double)rgb[i], gamma);
    int v = (int)floor(r * 255 + 0.5);
    out[i] = (uint8_t)(v > 255 ? 255 : (v < 0 ? 0 : v));
}

// experiment knob (tools/diag/overlap_probe.py): resident blocks per SM of the persistent kernels divided by this, so that
// two renders on two streams can share the SMs
static int g_grid_div = 1;
template <int STRAT>
int launch_bounce(bool first, bool exact, const BounceArgs& a, int grid, int sel_grid, int* sel_cursor, cudaStream_t st, cudaEvent_t ev_sweep0, cudaEvent_t ev_sweep1)
{
    constexpr bool kSph = (STRAT == MCPT_NEE_SPHERICAL || STRAT == MCPT_MIS || STRAT == MCPT_MIS_SINGLE);
    if (first) {
        if (exact) k_shade<STRAT, true, true><<<grid, kBlk, 0, st>>>(a);
        else k_shade<STRAT, true, false><<<grid, kBlk, 0, st>>>(a);
        return 1;
    }
    if constexpr (kSph) {
        if (a.flags & MCPT_LIGHT_TREE) {
            cudaEventRecord(ev_sweep0, st);
            k_lighttree<STRAT><<<std::max(1, std::min((a.in_count + 255) / 256, 148 * 8)), 256, 0, st>>>(a);
            cudaEventRecord(ev_sweep1, st);
        }
        else if (a.by_lists) {
            // survivor lists: group the sampling vertices by list, then sweep list by list (k_lsel_* above)
            const int cgrid = std::max(1, std::min((a.in_count + 255) / 256, 148 * 16));
            cudaMemsetAsync(a.lhist, 0, (size_t)a.s.nlists * kSelSub * 4, st);
            k_lsel_classify<STRAT><<<cgrid, 256, 0, st>>>(a);
            k_lsel_offsets<<<1, 1024, 0, st>>>(a.lhist, a.s.nlists * kSelSub, a.sorted, a.ngroups);
            k_lsel_scatter<<<cgrid, 256, 0, st>>>(a);
            // running-sum columns in shared memory (kSelV x list_max floats per warp) while >= 3 blocks per SM still fit
            const int cdf_stride = ((a.s.list_max / 32 + kCdfSub - 1) / kCdfSub) * 32;  // checkpoints per column x 32 lanes
            const size_t dyn = (size_t)(kBlk / 32) * kSelV * (size_t)cdf_stride * sizeof(float);
            const bool use_cdf = dyn > 0 && dyn <= 64 * 1024;
            if (use_cdf && dyn > 48 * 1024) cudaFuncSetAttribute(k_lightsel_sorted<STRAT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
            int sb = 1;
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&sb, k_lightsel_sorted<STRAT>, kBlk, use_cdf ? dyn : 0);
            const int sgrid = std::min(std::max(1, (a.in_count + 127) / 128), 148 * std::max(1, sb / g_grid_div));
            cudaEventRecord(ev_sweep0, st);
            k_lightsel_sorted<STRAT><<<sgrid, kBlk, use_cdf ? dyn : 0, st>>>(a, sel_cursor, use_cdf ? cdf_stride : 0);
            cudaEventRecord(ev_sweep1, st);
            if (exact) k_shade<STRAT, false, true><<<grid, kBlk, 0, st>>>(a);
            else k_shade<STRAT, false, false><<<grid, kBlk, 0, st>>>(a);
            return 5;
        }
        else {
            cudaEventRecord(ev_sweep0, st);
            k_lightsel<STRAT><<<std::min(grid, sel_grid), kBlk, 0, st>>>(a, sel_cursor);
            cudaEventRecord(ev_sweep1, st);
        }
    }
    if (exact) k_shade<STRAT, false, true><<<grid, kBlk, 0, st>>>(a);
    else k_shade<STRAT, false, false><<<grid, kBlk, 0, st>>>(a);
    return kSph ? 2 : 1;
}

template <int STRAT> int bounce_blocks_per_sm()
{
    int a = 0, b = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&a, k_shade<STRAT, true, false>, kBlk, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, k_shade<STRAT, false, false>, kBlk, 0);
    return std::max(1, std::min(a, b));
}

}  // namespace

// ---- test hooks (batch.cu): the production trace kernels on explicit rays ------------------------------------------------
namespace {
__global__ void k_vis_prepare(DScene s, int n, const float4* __restrict__ org, const float4* __restrict__ dir, const int2* __restrict__ ids,
                              int exact, ShadowQueue sq, float4* Lacc, int* visible)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Lacc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    const int2 id = ids[i];
    const int li = __float_as_int(__ldg(s.pverts + 3 * id.y + 1).w);
    bool ok = false;
    float be, ga, tl = 0.f;
    const f3 p = to_f3(org[i]), wl = to_f3(dir[i]);
    if (li >= 0) {
        // what k_shade does before it queues a visibility ray: distance to the sampled point on the light triangle
        const f3 v0 = to_f3(__ldg(s.lights + 5 * li)), v1 = to_f3(__ldg(s.lights + 5 * li + 1)), v2 = to_f3(__ldg(s.lights + 5 * li + 2));
        ok = exact ? light_point_test<true>(p, wl, v0, v1, v2, s.tmin, &be, &ga, &tl) : light_point_test<false>(p, wl, v0, v1, v2, s.tmin, &be, &ga, &tl);
    }
    visible[i] = ok ? 0 : -1;
    if (ok) {
        const int dst = atomicAdd(sq.count, 1);
        sq.org[dst] = make_float4(p.x, p.y, p.z, tl);
        sq.dir[dst] = make_float4(wl.x, wl.y, wl.z, __int_as_float(i));
        sq.ctr[dst] = make_float4(1.f, 0.f, 0.f, 0.f);
        sq.ids[dst] = id;
    }
}
__global__ void k_vis_finish(int n, const float4* __restrict__ Lacc, int* visible)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && visible[i] >= 0) visible[i] = Lacc[i].x > 0.5f ? 1 : 0;
}
}  // namespace

// closest hits of n explicit FP32 rays through k_extend (leaf_mode FAST) or k_extend_exact (EXACT)
void trace_rays_device(McptScene* scene, int n, const float4* d_org, const float4* d_dir, int leaf_mode, float4* d_hit)
{
    if (n <= 0) return;
    int* d_small = nullptr;  // [0] n, [1] cursor
    MCPT_CUDA(cudaMalloc((void**)&d_small, 8));
    const int init[2] = {n, 0};
    cudaError_t e = cudaMemcpy(d_small, init, 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        const int grid = std::min(148 * 8, (n + kBlk - 1) / kBlk);
        if (leaf_mode == MCPT_LEAF_EXACT) k_extend_exact<<<grid, kBlk>>>(scene->d, d_org, d_dir, d_small, d_hit);
        else if (scene->d.use8) k_extend8<<<grid, kBlk>>>(scene->d, d_org, d_dir, d_small, d_small + 1, d_hit);
        else k_extend<<<grid, kBlk>>>(scene->d, d_org, d_dir, d_small, d_small + 1, d_hit);
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaDeviceSynchronize();
    }
    cudaFree(d_small);
    MCPT_CUDA(e);
}

// the NEE visibility test of n explicit rays aimed at light triangles, through k_shadow / k_shadow_exact:
// visible[i] = 1 / 0, or -1 when the ray misses its light triangle (k_shade queues no visibility ray then)
void visibility_device(McptScene* scene, int n, const float4* d_org, const float4* d_dir, const int2* d_ids, int leaf_mode, int* d_visible)
{
    if (n <= 0) return;
    char* buf = nullptr;
    const size_t N = (size_t)n;
    MCPT_CUDA(cudaMalloc((void**)&buf, N * (16 * 4 + 8) + 256));
    ShadowQueue sq;
    sq.org = (float4*)buf; sq.dir = sq.org + N; sq.ctr = sq.dir + N;
    float4* Lacc = sq.ctr + N;
    sq.ids = (int2*)(Lacc + N);
    int* small = (int*)(sq.ids + N);  // [0] count, [1] cursor, [8..] counters
    sq.count = small;
    cudaError_t e = cudaMemset(small, 0, 256);
    if (e == cudaSuccess) {
        const int exact = leaf_mode == MCPT_LEAF_EXACT;
        k_vis_prepare<<<(n + 255) / 256, 256>>>(scene->d, n, d_org, d_dir, d_ids, exact, sq, Lacc, d_visible);
        const int grid = std::min(148 * 8, (n + kBlk - 1) / kBlk);
        if (exact) k_shadow_exact<<<grid, kBlk>>>(scene->d, sq, Lacc, (unsigned long long*)(small + 8));
        else if (scene->d.use8) k_shadow8<<<grid, kBlk>>>(scene->d, sq, small + 1, Lacc, (unsigned long long*)(small + 8));
        else k_shadow<<<grid, kBlk>>>(scene->d, sq, small + 1, Lacc, (unsigned long long*)(small + 8));
        k_vis_finish<<<(n + 255) / 256, 256>>>(n, Lacc, d_visible);
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaDeviceSynchronize();
    }
    cudaFree(buf);
    MCPT_CUDA(e);
}

// ---- the same visibility test without any allocation (mcpt_visibility_noalloc, include/mcpt.h) ---------------------------
// For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it): the scratch
// visibility_device allocates per call -- the visibility-ray queue k_shadow consumes, the per-ray accumulators, the queue
// length / ticket / counter words -- lives in ONE file-scope device array of a fixed size, 72 bytes per ray for at most
// MCPT_NOALLOC_MAX_RAYS rays (4.5 MiB, held by every GPU that loads the library until exit).  In host code the array's name
// is only a stand-in for the symbol; its device address is looked up once per device with cudaGetSymbolAddress, with that
// device current, and that pointer is what the kernels get.  Every call on a device shares the array and finds the last
// call's contents in it: the words that must start at zero are cleared on the stream first, everything else is written
// before it is read.  Two calls that overlap on one GPU (different streams) race on it.
namespace {
constexpr size_t kNoAllocBytes = (size_t)MCPT_NOALLOC_MAX_RAYS * (16 * 4 + 8) + 256;
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];
constexpr int kNoAllocMaxDevices = 64;
void* g_noalloc_ptr[kNoAllocMaxDevices] = {};
std::mutex g_noalloc_mutex;

char* noalloc_scratch(int device)
{
    if (device < 0 || device >= kNoAllocMaxDevices) throw Error(MCPT_ERR_INVALID, "device index beyond the no-allocation scratch table");
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device]) {
        void* p = nullptr;
        MCPT_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));  // the caller made `device` current
        if (!p) throw Error(MCPT_ERR_CUDA, "cudaGetSymbolAddress returned a null scratch address");
        g_noalloc_ptr[device] = p;
    }
    return (char*)g_noalloc_ptr[device];
}
}  // namespace

void visibility_noalloc_device(McptScene* scene, size_t n, const float4* d_org, const float4* d_dir, const int2* d_ids, int leaf_mode,
                               int* d_visible, cudaStream_t st)
{
    if (n > (size_t)MCPT_NOALLOC_MAX_RAYS)
        throw Error(MCPT_ERR_INVALID, "mcpt_visibility_noalloc: " + std::to_string(n) + " rays exceed the fixed scratch of MCPT_NOALLOC_MAX_RAYS = " +
                                          std::to_string(MCPT_NOALLOC_MAX_RAYS) + " rays; split the batch or use mcpt_visibility_batch");
    if (n == 0) return;
    MCPT_CUDA(cudaSetDevice(scene->device));
    char* buf = noalloc_scratch(scene->device);
    const size_t N = (size_t)MCPT_NOALLOC_MAX_RAYS;  // fixed carve-up, whatever n is
    ShadowQueue sq;
    sq.org = (float4*)buf; sq.dir = sq.org + N; sq.ctr = sq.dir + N;
    float4* Lacc = sq.ctr + N;
    sq.ids = (int2*)(Lacc + N);
    int* small = (int*)(sq.ids + N);  // [0] count, [1] cursor, [8..] counters
    sq.count = small;
    MCPT_CUDA(cudaMemsetAsync(small, 0, 256, st));  // what must start at zero; Lacc[i] and visible[i] are written by k_vis_prepare
    const int ni = (int)n, exact = leaf_mode == MCPT_LEAF_EXACT;
    k_vis_prepare<<<(ni + 255) / 256, 256, 0, st>>>(scene->d, ni, d_org, d_dir, d_ids, exact, sq, Lacc, d_visible);
    const int grid = std::min(148 * 8, (ni + kBlk - 1) / kBlk);
    if (exact) k_shadow_exact<<<grid, kBlk, 0, st>>>(scene->d, sq, Lacc, (unsigned long long*)(small + 8));
    else if (scene->d.use8) k_shadow8<<<grid, kBlk, 0, st>>>(scene->d, sq, small + 1, Lacc, (unsigned long long*)(small + 8));
    else k_shadow<<<grid, kBlk, 0, st>>>(scene->d, sq, small + 1, Lacc, (unsigned long long*)(small + 8));
    k_vis_finish<<<(ni + 255) / 256, 256, 0, st>>>(ni, Lacc, d_visible);
    MCPT_CUDA(cudaGetLastError());
}

// ---- which wide BVH layout the trace kernels walk for this scene: measured, not guessed -----------------------------------
// The compressed 8-wide nodes need fewer and smaller loads per visit but more ALU work (byte decoding, no distance sort):
// they win where the 4-wide traversal is bound by the L1 pipeline (1 M-triangle soup +26 %, Cornell box +17 %) and lose
// where it is bound by the ALU pipe (veach-mis, interior: -5 %).  Both layouts return identical hits, so the choice is
// free: scene creation traces 512 Ki rays leaving random surface points -- half into the hemisphere of their shading normal,
// half towards random points on the lights, the two kinds of rays a render's bounces make -- through both kernels and keeps
// the faster.
namespace {
__global__ void k_calib_rays(DScene s, int n, float4* org, float4* dir)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Rng4 a = rng_draw4(0x5eedu, (uint32_t)i, 0u, 1u, 0u), b = rng_draw4(0x5eedu, (uint32_t)i, 0u, 1u, 1u);
    const int prim = min(s.P - 1, (int)(a.u[0] * (float)s.P));
    const float be = 1.f - sqrtf(1.f - a.u[1]), ga = (1.f - be) * a.u[2];
    const f3 p = to_f3(__ldg(s.pverts + 3 * prim)) * (1.f - be - ga) + to_f3(__ldg(s.pverts + 3 * prim + 1)) * be + to_f3(__ldg(s.pverts + 3 * prim + 2)) * ga;
    const f3 nn = to_f3(__ldg(s.pnorms + 3 * prim)) * (1.f - be - ga) + to_f3(__ldg(s.pnorms + 3 * prim + 1)) * be + to_f3(__ldg(s.pnorms + 3 * prim + 2)) * ga;
    const float z = 1.f - 2.f * b.u[0], r = sqrtf(fmaxf(0.f, 1.f - z * z)), phi = 6.2831853f * b.u[1];
    f3 w = mk<float>(r * cosf(phi), r * sinf(phi), z);
    if (dot(w, nn) < 0.f) w = -w;  // uniform over the hemisphere of the shading normal
    if ((i & 1) && s.L > 0) {      // every other ray: towards a random point of a random light triangle (light-sample / visibility rays)
        const int l = min(s.L - 1, (int)(a.u[3] * (float)s.L));
        const float lb = 1.f - sqrtf(1.f - b.u[2]), lg = (1.f - lb) * b.u[3];
        const f3 t = to_f3(__ldg(s.lights + 5 * l)) * (1.f - lb - lg) + to_f3(__ldg(s.lights + 5 * l + 1)) * lb + to_f3(__ldg(s.lights + 5 * l + 2)) * lg;
        const f3 dl = t - p;
        const float d2 = dot(dl, dl);
        if (d2 > 0.f) w = dl * rsqrtf(d2);
    }
    org[i] = make_float4(p.x, p.y, p.z, __int_as_float(prim));
    dir[i] = make_float4(w.x, w.y, w.z, __int_as_float(i));
}
}  // namespace

void calibrate_trace_layout(McptScene* scene)
{
    scene->calib_ms[0] = scene->calib_ms[1] = 0.f;
    const int n = 512 << 10;
    char* buf = nullptr;
    MCPT_CUDA(cudaMalloc((void**)&buf, (size_t)n * 48 + 64));
    float4 *org = (float4*)buf, *dir = org + n, *hit = dir + n;
    int* small = (int*)(hit + n);  // [0] n, [1] cursor
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    cudaError_t err = cudaEventCreate(&e0);
    if (err == cudaSuccess) err = cudaEventCreate(&e1);
    if (err == cudaSuccess) {
        k_calib_rays<<<(n + 255) / 256, 256>>>(scene->d, n, org, dir);
        const int grid = std::min(148 * 8, (n + kBlk - 1) / kBlk);
        for (int rep = 0; rep < 3 && err == cudaSuccess; rep++)  // repetition 0 warms the caches
            for (int w = 0; w < 2 && err == cudaSuccess; w++) {
                const int init[2] = {n, 0};
                err = cudaMemcpy(small, init, 8, cudaMemcpyHostToDevice);
                cudaEventRecord(e0);
                if (w) k_extend8<<<grid, kBlk>>>(scene->d, org, dir, small, small + 1, hit);
                else k_extend<<<grid, kBlk>>>(scene->d, org, dir, small, small + 1, hit);
                cudaEventRecord(e1);
                if (err == cudaSuccess) err = cudaEventSynchronize(e1);
                float ms = 0.f;
                cudaEventElapsedTime(&ms, e0, e1);
                if (rep) scene->calib_ms[w] = rep == 1 ? ms : std::min(scene->calib_ms[w], ms);
            }
        if (err == cudaSuccess) err = cudaGetLastError();
    }
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    cudaFree(buf);
    MCPT_CUDA(err);
    scene->d.use8 = scene->calib_ms[1] < scene->calib_ms[0] ? 1 : 0;
}

void workspace_free(Workspace& ws)
{
    if (ws.base) cudaFree(ws.base);
    if (ws.pinned) cudaFreeHost(ws.pinned);
    if (ws.d_image) cudaFree(ws.d_image);
    ws = Workspace();
}

void tonemap_device(McptScene* scene, const float* rgb, size_t n, double max_radiance, double gamma, uint8_t* out)
{
    (void)scene;
    float* d_in = nullptr;
    uint8_t* d_out = nullptr;
    size_t n3 = n * 3;
    MCPT_CUDA(cudaMalloc((void**)&d_in, n3 * sizeof(float)));
    MCPT_CUDA(cudaMalloc((void**)&d_out, n3));
    MCPT_CUDA(cudaMemcpy(d_in, rgb, n3 * sizeof(float), cudaMemcpyHostToDevice));
    if (n3) k_tonemap<<<(unsigned)((n3 + 255) / 256), 256>>>(d_in, n3, std::pow(max_radiance, -gamma), gamma, d_out);
    cudaError_t e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = cudaMemcpy(out, d_out, n3, cudaMemcpyDeviceToHost);
    cudaFree(d_in); cudaFree(d_out);
    MCPT_CUDA(e);
}

namespace {
struct QueueOverflow { int batch; };  // thrown by render_wavefront when the MIS path tree outgrows its queue

// One attempt at the whole call.  tree_factor: capacity of the MIS path-tree queues in entries per path slot (x 1/100).
void render_wavefront(McptScene* scene, const McptCamera* cam, const McptRenderParams* prm, float* d_out_rgb, cudaStream_t st,
                      McptStats* stats, int tree_factor, size_t pretend_cap)
{
    const bool dbg_timing = std::getenv("MCPT_DEBUG_TIMING") != nullptr;
    const auto dbg_t0 = std::chrono::steady_clock::now();
    auto dbg_ms = [&]() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - dbg_t0).count(); };
    double dbg_setup = 0, dbg_loop = 0;
    const int W = prm->width, H = prm->height;
    if (W <= 0 || H <= 0 || prm->spp <= 0) throw Error(MCPT_ERR_INVALID, "width, height and spp must be positive");
    if (prm->strategy < 0 || prm->strategy > 4) throw Error(MCPT_ERR_INVALID, "unknown strategy");
    if (prm->leaf_mode != MCPT_LEAF_FAST && prm->leaf_mode != MCPT_LEAF_EXACT) throw Error(MCPT_ERR_INVALID, "unknown leaf_mode");
    if (prm->reserved != 0) throw Error(MCPT_ERR_INVALID, "McptRenderParams.reserved must be 0");
    const bool exact = prm->leaf_mode == MCPT_LEAF_EXACT;
    if (prm->strategy != MCPT_BRDF_ONLY && scene->L == 0) throw Error(MCPT_ERR_INVALID, "light sampling needs at least one light");
    const size_t npix = (size_t)W * H;
    if (npix > (1u << 30)) throw Error(MCPT_ERR_INVALID, "image too large");
    const bool tree = prm->strategy == MCPT_MIS;
    const int spp_total = prm->spp_total > 0 ? prm->spp_total : prm->spp;
    const float p_rr = prm->p_rr > 0.f ? prm->p_rr : 0.6f;
    size_t target = kTargetSlots;
    if (const char* e = std::getenv("MCPT_TARGET_SLOTS_MI")) { const long v = std::atol(e); if (v > 0) target = (size_t)v << 20; }
    {
        // keep the workspace below half of the device memory that is free (or already ours).  cudaMemGetInfo goes to
        // the kernel driver and was measured to stall for 30-65 ms now and then when another thread of the process is
        // inside an NVML query (bench.py samples the clocks), so the answer is cached while the workspace lives.
        MCPT_CUDA(cudaSetDevice(scene->device));
        if (scene->ws.mem_budget == 0 || scene->ws.bytes == 0) {
            size_t free_b = 0, total_b = 0;
            if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) scene->ws.mem_budget = free_b + scene->ws.bytes;
            else cudaGetLastError();
        }
        if (scene->ws.mem_budget) {
            const size_t per_slot = (tree ? 224 + (size_t)155 * tree_factor / 100 : 224) + (scene->d.nlists > 0 ? (tree ? (size_t)8 * tree_factor / 100 : 8) : 0);  // bytes of queues + staging + hit + accumulators per path slot
            target = std::min(target, std::max<size_t>(npix, scene->ws.mem_budget / 2 / per_slot));
        }
    }
    int batch = prm->batch_spp > 0 ? prm->batch_spp : (int)std::max<size_t>(1, target / npix);
    batch = std::min(batch, prm->spp);
    const size_t slots = npix * (size_t)batch;
    if (slots > (1u << 30)) throw Error(MCPT_ERR_INVALID, "batch_spp * pixels too large");
    const size_t cap = tree ? std::max(slots * (size_t)tree_factor / 100, slots + 1) : slots;  // never below the first bounce's own queue
    if (cap > (1u << 30)) throw Error(MCPT_ERR_INVALID, "batch_spp * pixels too large for the path-tree queue");
    const size_t chunks = (cap + kChunk - 1) / kChunk + 1;
    const bool single = prm->strategy == MCPT_MIS_SINGLE;
    const bool sph = prm->strategy == MCPT_NEE_SPHERICAL || prm->strategy == MCPT_MIS || single;

    int dev = scene->device, sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int bps = prm->strategy == MCPT_NEE_SPHERICAL ? bounce_blocks_per_sm<MCPT_NEE_SPHERICAL>()
              : prm->strategy == MCPT_BRDF_ONLY   ? bounce_blocks_per_sm<MCPT_BRDF_ONLY>()
              : prm->strategy == MCPT_MIS         ? bounce_blocks_per_sm<MCPT_MIS>()
              : prm->strategy == MCPT_MIS_SINGLE  ? bounce_blocks_per_sm<MCPT_MIS_SINGLE>()
                                                  : bounce_blocks_per_sm<MCPT_NEE_UNIFORM>();
    const size_t sq_warps = (size_t)sms * bps * (kBlk / 32);
    // workspace carve-up
    size_t off = 0;
    auto place = [&](size_t bytes) { size_t at = off; off = (off + bytes + 255) & ~(size_t)255; return at; };
    // Q[0]: dense ray queue (cap entries).  Q[1]: staging queue, one window of `stride` entries per chunk of kChunk queue entries.
    const size_t stride = tree ? 2 * kChunk : kChunk;
    const size_t qn[2] = {cap, chunks * stride};
    size_t o_q[2][4];
    for (int q = 0; q < 2; q++) {
        o_q[q][0] = place(qn[q] * 16); o_q[q][1] = place(qn[q] * 16); o_q[q][2] = place(qn[q] * 16);
        o_q[q][3] = place(tree ? qn[q] * 8 : 16);
    }
    const size_t o_lacc = place(slots * 16), o_contrib = place(tree ? cap * 16 : 16), o_ph = place(npix * 16), o_img = place(npix * 16), o_comp = place(npix * 16);
    const size_t groups = (chunks + kGroup - 1) / kGroup;
    const size_t o_counts = place(chunks * 4), o_small = place(256 + groups * 4);
    const size_t o_sel = place(sph ? slots * 8 : 16), o_wsum = place(sph ? npix * 4 : 16), o_sel4 = place(sph ? cap * 16 : 16);
    const bool lists = sph && scene->d.nlists > 0 && scene->d.L > kSerialLights && !(prm->flags & MCPT_LIGHT_TREE);
    const size_t nl = lists ? (size_t)scene->d.nlists : 0;
    const size_t o_skey = place(lists ? cap * 4 : 16), o_sorted = place(lists ? (cap + kSelV * kSelSub * nl) * 4 : 16), o_lhist = place(nl * kSelSub * 4 + 16);
    const bool nee = prm->strategy == MCPT_NEE_SPHERICAL || prm->strategy == MCPT_NEE_UNIFORM || prm->strategy == MCPT_MIS_SINGLE;  // queues visibility rays
    const size_t o_hit = place(cap * 16);
    // every resident warp of k_shade may leave one partly used reservation behind
    const size_t sqn = nee ? cap + (size_t)sq_warps * kSqGrab : 1;
    const size_t o_sq[4] = {place(sqn * 16), place(sqn * 16), place(sqn * 16), place(sqn * 8)};
    Workspace& ws = scene->ws;
    if (ws.bytes < off) {
        if (ws.base) MCPT_CUDA(cudaFree(ws.base));
        ws.base = nullptr; ws.bytes = 0;
        if (cudaMalloc(&ws.base, off) != cudaSuccess) { cudaGetLastError(); throw Error(MCPT_ERR_OOM, "cannot allocate render workspace; lower batch_spp"); }
        ws.bytes = off;
    }
    if (!ws.pinned) MCPT_CUDA(cudaMallocHost(&ws.pinned, 256));
    char* base = (char*)ws.base;
    Queue Q[2];
    for (int q = 0; q < 2; q++) {
        Q[q].org = (float4*)(base + o_q[q][0]); Q[q].dir = (float4*)(base + o_q[q][1]); Q[q].thr = (float4*)(base + o_q[q][2]);
        Q[q].node = (uint2*)(base + o_q[q][3]);
    }
    float4* Lacc = (float4*)(base + o_lacc);
    float4* contrib = (float4*)(base + o_contrib);
    float4* primary = (float4*)(base + o_ph);
    float4* image = (float4*)(base + o_img);
    float4* comp = (float4*)(base + o_comp);
    int* counts = (int*)(base + o_counts);
    int2* sel = (int2*)(base + o_sel);
    float* wsum_pix = (float*)(base + o_wsum);
    float4* sel4 = (float4*)(base + o_sel4);
    float4* hit = (float4*)(base + o_hit);
    // small block: [0] out_count, [1] shade ticket, [2] shadow-queue length, [3] k_extend cursor, [4] k_shadow cursor,
    //              [5] k_lightsel cursor (ints; [1..8] are zeroed before every bounce),
    //              [64..] counters (u64 x CNT_N), [256..] group sums (int x groups)
    int* d_count = (int*)(base + o_small);
    int* d_ticket = d_count + 1;
    ShadowQueue SQ;
    SQ.org = (float4*)(base + o_sq[0]); SQ.dir = (float4*)(base + o_sq[1]); SQ.ctr = (float4*)(base + o_sq[2]);
    SQ.ids = (int2*)(base + o_sq[3]); SQ.count = d_count + 2;
    int* d_ext_cursor = d_count + 3;
    int* d_sh_cursor = d_count + 4;
    int* d_sel_cursor = d_count + 5;
    unsigned long long* d_counters = (unsigned long long*)(base + o_small + 64);
    int* d_groupsum = (int*)(base + o_small + 256);
    volatile int* h_count = (volatile int*)ws.pinned;
    unsigned long long* h_counters = (unsigned long long*)((char*)ws.pinned + 64);

    // camera (main.cpp:546-553), basis in FP64 on the host
    DCamera dc;
    {
        auto nrm = [](double* v) { double l = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); v[0] /= l; v[1] /= l; v[2] /= l; };
        auto crs = [](const double* a, const double* b, double* o) { o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0]; };
        double N[3] = {cam->w[0], cam->w[1], cam->w[2]}, V[3], U[3];
        double wl = std::sqrt(N[0] * N[0] + N[1] * N[1] + N[2] * N[2]);
        if (!(wl > 0)) throw Error(MCPT_ERR_INVALID, "camera w must be non-zero");
        nrm(N); crs(N, cam->up, V); nrm(V); crs(V, N, U); nrm(U);
        for (int k = 0; k < 3; k++) { dc.eye[k] = (float)cam->eye[k]; dc.U[k] = (float)U[k]; dc.V[k] = (float)V[k]; dc.N[k] = (float)N[k]; }
        dc.pixellen = (float)cam->pixellen; dc.wlen = (float)wl;
        dc.ci = (float)((H - 1) * 0.5); dc.cj = (float)((W - 1) * 0.5);
        dc.W = W; dc.H = H;
    }

    if (const char* e = std::getenv("MCPT_GRID_DIV")) g_grid_div = std::max(1, std::atoi(e));
    const int gdiv = g_grid_div;
    const int max_grid = sms * std::max(1, bps / gdiv);
    bool use8 = scene->d.use8 != 0;
    int ebps = 1, sbps = 1, ebps8 = 1, sbps8 = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ebps8, k_extend8, kBlk, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&sbps8, k_shadow8, kBlk, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ebps, k_extend, kBlk, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&sbps, k_shadow, kBlk, 0);
    const int ext_grid4 = sms * std::max(1, ebps / gdiv), sh_grid4 = sms * std::max(1, sbps / gdiv), ext_grid8 = sms * std::max(1, ebps8 / gdiv), sh_grid8 = sms * std::max(1, sbps8 / gdiv);
    int ext_grid = use8 ? ext_grid8 : ext_grid4, sh_grid = use8 ? sh_grid8 : sh_grid4;
    // The layout guess of scene creation (random rays) is settled on the first real bounce: the rays a render makes are
    // more coherent than the calibration batch, which shifts the balance towards the 4-wide nodes on some scenes (interior).
    cudaEvent_t evm[4] = {nullptr, nullptr, nullptr, nullptr};
    int lbps = 1;
    if (prm->strategy == MCPT_MIS) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&lbps, k_lightsel<MCPT_MIS>, kBlk, 0);
    else if (prm->strategy == MCPT_NEE_SPHERICAL) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&lbps, k_lightsel<MCPT_NEE_SPHERICAL>, kBlk, 0);
    else if (prm->strategy == MCPT_MIS_SINGLE) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&lbps, k_lightsel<MCPT_MIS_SINGLE>, kBlk, 0);
    const int sel_grid = sms * std::max(1, lbps / gdiv);

    McptStats S;
    std::memset(&S, 0, sizeof S);
    S.paths = (uint64_t)npix * (uint64_t)prm->spp;
    dbg_setup = dbg_ms();
    MCPT_CUDA(cudaEventRecord(scene->ev[0], st));
    MCPT_CUDA(cudaMemsetAsync(d_counters, 0, CNT_N * 8, st));
    const bool jitter = (prm->flags & MCPT_JITTER) != 0;
    if (!jitter) {
        if (exact) k_primary_exact<<<(unsigned)((npix + kBlk - 1) / kBlk), kBlk, 0, st>>>(scene->d, dc, (int)npix, primary);
        else k_primary<<<(unsigned)((npix + kBlk - 1) / kBlk), kBlk, 0, st>>>(scene->d, dc, (int)npix, primary);
        S.kernel_launches++;
        S.rays_closest += npix;
    }
    float ms_bounce = 0.f, ms_shade = 0.f, ms_sweep = 0.f, ms_prep = 0.f;
    bool prep_timed = false;
    bool overflow = false, measuring = false;
    int done = 0, pass = 0;
    while (done < prm->spp) {
        const int b = std::min(batch, prm->spp - done);
        BounceArgs a;
        a.s = scene->d; a.cam = dc;
        a.Lacc = Lacc; a.contrib = contrib; a.primary_hit = primary; a.sel = sel; a.wsum_pix = wsum_pix; a.sel4 = sel4; a.counts = counts; a.groupsum = d_groupsum; a.ticket = d_ticket;
        a.hit = hit; a.sq = SQ;
        a.skey = (int*)(base + o_skey); a.sorted = (int*)(base + o_sorted); a.lhist = (int*)(base + o_lhist); a.ngroups = a.lhist + nl * kSelSub;
        a.counters = d_counters; a.seed = prm->seed; a.capacity = (int)cap; a.npix = (int)npix; a.batch_spp = b;
        a.sample_base = prm->spp_offset + done; a.max_depth = prm->max_depth; a.flags = prm->flags; a.p_rr = p_rr;
        a.by_lists = lists ? 1 : 0;
        int count = (int)(npix * (size_t)b);
        a.in = Q[0]; a.out = Q[1];
        if (sph && !jitter) {
            // first-vertex light draws, weights computed once per pixel (F4: all samples share the camera ray)
            const size_t per_block = (size_t)(kBlk / 32) * (2 * (size_t)scene->d.Lpad + 64) * sizeof(float);
            const int use_cdf = per_block <= 200 * 1024;  // larger light tables: per-sample warp picks instead
            const size_t dyn = use_cdf ? per_block : 0;
            if (dyn > 48 * 1024) MCPT_CUDA(cudaFuncSetAttribute(k_lightprep, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
            int pbps = 1;
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&pbps, k_lightprep, kBlk, dyn);
            const size_t npairs = (npix + 1) / 2;
            const int pgrid = (int)std::min<size_t>((npairs + kBlk / 32 - 1) / (kBlk / 32), (size_t)sms * std::max(1, pbps / gdiv));
            MCPT_CUDA(cudaEventRecord(scene->ev[8], st));
            k_lightprep<<<pgrid, kBlk, dyn, st>>>(scene->d, dc, primary, (int)npix, b, a.sample_base, prm->seed, p_rr, (tree || single) ? 1 : 0,
                                                  prm->max_depth, use_cdf, sel, wsum_pix, d_counters);
            MCPT_CUDA(cudaEventRecord(scene->ev[9], st));
            prep_timed = true;
            MCPT_CUDA(cudaGetLastError());
            S.kernel_launches++;
        }
        if (jitter) {
            // jittered camera rays differ per sample: generate them into the dense queue and trace them like any bounce
            k_raygen<<<(unsigned)((count + 255) / 256), 256, 0, st>>>(dc, (int)npix, count, a.sample_base, prm->seed, Q[0], tree ? 1 : 0, Lacc, d_count);
            MCPT_CUDA(cudaMemsetAsync(d_ticket, 0, 32, st));
            if (exact) k_extend_exact<<<std::min(ext_grid, (count + kBlk - 1) / kBlk), kBlk, 0, st>>>(scene->d, Q[0].org, Q[0].dir, d_count, hit);
            else if (use8) k_extend8<<<std::min(ext_grid, (count + kBlk - 1) / kBlk), kBlk, 0, st>>>(scene->d, Q[0].org, Q[0].dir, d_count, d_ext_cursor, hit);
            else k_extend<<<std::min(ext_grid, (count + kBlk - 1) / kBlk), kBlk, 0, st>>>(scene->d, Q[0].org, Q[0].dir, d_count, d_ext_cursor, hit);
            MCPT_CUDA(cudaGetLastError());
            S.kernel_launches += 2;
            S.rays_closest += (uint64_t)count;
        }
        for (int bounce = 0; count > 0; bounce++) {
            if (prm->max_depth <= 0 && bounce >= 256) break;  // unlimited depth: RR has long emptied the queue
            a.in_count = count; a.bounce = bounce;
            const int nch = (count + kChunk - 1) / kChunk, ngroups = (nch + kGroup - 1) / kGroup;
            MCPT_CUDA(cudaMemsetAsync(d_ticket, 0, 32, st));
            MCPT_CUDA(cudaMemsetAsync(d_groupsum, 0, (size_t)ngroups * 4, st));
            MCPT_CUDA(cudaEventRecord(scene->ev[2], st));
            const int grid = std::min((nch + kBlk / 32 - 1) / (kBlk / 32), max_grid);
            const bool first = bounce == 0 && !jitter;
            int nl = 1;
            switch (prm->strategy) {
                case MCPT_NEE_SPHERICAL: nl = launch_bounce<MCPT_NEE_SPHERICAL>(first, exact, a, grid, sel_grid, d_sel_cursor, st, scene->ev[6], scene->ev[7]); break;
                case MCPT_BRDF_ONLY: nl = launch_bounce<MCPT_BRDF_ONLY>(first, exact, a, grid, sel_grid, d_sel_cursor, st, scene->ev[6], scene->ev[7]); break;
                case MCPT_MIS: nl = launch_bounce<MCPT_MIS>(first, exact, a, grid, sel_grid, d_sel_cursor, st, scene->ev[6], scene->ev[7]); break;
                case MCPT_MIS_SINGLE: nl = launch_bounce<MCPT_MIS_SINGLE>(first, exact, a, grid, sel_grid, d_sel_cursor, st, scene->ev[6], scene->ev[7]); break;
                default: nl = launch_bounce<MCPT_NEE_UNIFORM>(first, exact, a, grid, sel_grid, d_sel_cursor, st, scene->ev[6], scene->ev[7]); break;
            }
            MCPT_CUDA(cudaGetLastError());
            MCPT_CUDA(cudaEventRecord(scene->ev[4], st));
            S.kernel_launches += nl;
            S.elements += (uint64_t)count;
            const bool last = prm->max_depth > 0 && bounce + 1 >= prm->max_depth + 1;  // vertices deeper than max_depth never spawn
            if (!last) {
                k_gather<<<ngroups, 256, 0, st>>>(Q[1], Q[0], counts, d_groupsum, nch, (int)stride, (int)(pretend_cap ? std::min(cap, pretend_cap) : cap), tree ? 1 : 0, d_count, d_counters);
                const int eg = std::min(ext_grid, (count * (tree ? 2 : 1) + kBlk - 1) / kBlk);
                measuring = !exact && scene->layout_free && !scene->layout_final && count >= (1 << 18);
                if (measuring) {
                    // both layouts on the same queue (identical hits): time them, keep the faster from here on
                    for (auto& e : evm) if (!e) MCPT_CUDA(cudaEventCreate(&e));
                    const int need = (count * (tree ? 2 : 1) + kBlk - 1) / kBlk;
                    // ... and the 8-wide kernel at the two ends of its refill-threshold range (bvh.h: kRefill8Lo / kRefill8Hi)
                    DScene d_lo = scene->d, d_hi = scene->d;
                    d_lo.refill8 = kRefill8Lo; d_hi.refill8 = kRefill8Hi;
                    MCPT_CUDA(cudaEventRecord(evm[0], st));
                    k_extend<<<std::min(ext_grid4, need), kBlk, 0, st>>>(scene->d, Q[0].org, Q[0].dir, d_count, d_ext_cursor, hit);
                    MCPT_CUDA(cudaEventRecord(evm[1], st));
                    MCPT_CUDA(cudaMemsetAsync(d_ext_cursor, 0, 4, st));
                    k_extend8<<<std::min(ext_grid8, need), kBlk, 0, st>>>(d_lo, Q[0].org, Q[0].dir, d_count, d_ext_cursor, hit);
                    MCPT_CUDA(cudaEventRecord(evm[2], st));
                    MCPT_CUDA(cudaMemsetAsync(d_ext_cursor, 0, 4, st));
                    k_extend8<<<std::min(ext_grid8, need), kBlk, 0, st>>>(d_hi, Q[0].org, Q[0].dir, d_count, d_ext_cursor, hit);
                    MCPT_CUDA(cudaEventRecord(evm[3], st));
                    S.kernel_launches += 2;
                } else if (exact) k_extend_exact<<<eg, kBlk, 0, st>>>(scene->d, Q[0].org, Q[0].dir, d_count, hit);
                else if (use8) k_extend8<<<eg, kBlk, 0, st>>>(scene->d, Q[0].org, Q[0].dir, d_count, d_ext_cursor, hit);
                else k_extend<<<eg, kBlk, 0, st>>>(scene->d, Q[0].org, Q[0].dir, d_count, d_ext_cursor, hit);
                S.kernel_launches += 2;
            }
            if (nee) {
                if (exact) k_shadow_exact<<<std::min(sh_grid, (count + kBlk - 1) / kBlk), kBlk, 0, st>>>(scene->d, SQ, Lacc, d_counters);
                else if (use8) k_shadow8<<<std::min(sh_grid, (count + kBlk - 1) / kBlk), kBlk, 0, st>>>(scene->d, SQ, d_sh_cursor, Lacc, d_counters);
                else k_shadow<<<std::min(sh_grid, (count + kBlk - 1) / kBlk), kBlk, 0, st>>>(scene->d, SQ, d_sh_cursor, Lacc, d_counters);
                S.kernel_launches++;
            }
            if (tree && !first) {
                k_runsum<<<(unsigned)((count + 255) / 256), 256, 0, st>>>(contrib, count, Lacc);
                S.kernel_launches++;
            }
            MCPT_CUDA(cudaGetLastError());
            MCPT_CUDA(cudaEventRecord(scene->ev[3], st));
            if (!last) MCPT_CUDA(cudaMemcpyAsync((void*)h_count, d_count, 4, cudaMemcpyDeviceToHost, st));
            MCPT_CUDA(cudaStreamSynchronize(st));
            if (measuring) {
                float t4 = 0.f, t8lo = 0.f, t8hi = 0.f;
                cudaEventElapsedTime(&t4, evm[0], evm[1]);
                cudaEventElapsedTime(&t8lo, evm[1], evm[2]);
                cudaEventElapsedTime(&t8hi, evm[2], evm[3]);
                const float t8 = std::min(t8lo, t8hi);
                scene->render_ms[0] = t4; scene->render_ms[1] = t8;
                scene->d.refill8 = t8hi < t8lo ? kRefill8Hi : kRefill8Lo;
                use8 = t8 < t4;
                scene->d.use8 = use8 ? 1 : 0;
                a.s.use8 = scene->d.use8; a.s.refill8 = scene->d.refill8;
                scene->layout_final = true;
                ext_grid = use8 ? ext_grid8 : ext_grid4; sh_grid = use8 ? sh_grid8 : sh_grid4;
                measuring = false;
            }
            float ms = 0.f;
            cudaEventElapsedTime(&ms, scene->ev[2], scene->ev[3]);
            ms_bounce += ms;
            cudaEventElapsedTime(&ms, scene->ev[2], scene->ev[4]);
            ms_shade += ms;
            if (sph && !first) {  // the light-sweep kernel of this bounce alone
                float msw = 0.f;
                if (cudaEventElapsedTime(&msw, scene->ev[6], scene->ev[7]) == cudaSuccess) ms_sweep += msw; else cudaGetLastError();
            }
            if (prep_timed) {  // k_lightprep of this pass (it finished before the first bounce did)
                float mp = 0.f;
                if (cudaEventElapsedTime(&mp, scene->ev[8], scene->ev[9]) == cudaSuccess) ms_prep += mp; else cudaGetLastError();
                prep_timed = false;
            }
            if (dbg_timing) {
                float mb = 0.f;
                cudaEventElapsedTime(&mb, scene->ev[2], scene->ev[3]);
                std::fprintf(stderr, "  bounce %d: %d rays, %.3f ms (shade %.3f)\n", bounce, count, mb, ms);
            }
            S.queue_peak = std::max<uint32_t>(S.queue_peak, (uint32_t)count);
            S.bounces++;
            if (last) break;
            count = *h_count;
            if (count < 0) { for (auto& e : evm) if (e) cudaEventDestroy(e); throw QueueOverflow{batch}; }  // k_gather: the tree outgrew the queue; nothing of this call has reached d_out_rgb yet
            S.rays_closest += (uint64_t)count;
        }
        k_accumulate<<<(unsigned)((npix + 255) / 256), 256, 0, st>>>(Lacc, (int)npix, b, image, comp, pass == 0);
        S.kernel_launches++;
        done += b;
        pass++;
    }
    k_finalize<<<(unsigned)((npix + 255) / 256), 256, 0, st>>>(image, (int)npix, 1.0f / (float)spp_total, (prm->flags & MCPT_ACCUMULATE) ? 1 : 0, d_out_rgb);
    S.kernel_launches++;
    MCPT_CUDA(cudaMemcpyAsync(h_counters, d_counters, CNT_N * 8, cudaMemcpyDeviceToHost, st));
    MCPT_CUDA(cudaEventRecord(scene->ev[1], st));
    dbg_loop = dbg_ms();
    MCPT_CUDA(cudaStreamSynchronize(st));
    MCPT_CUDA(cudaGetLastError());
    if (dbg_timing) std::fprintf(stderr, "render_device host ms: setup %.3f, loop issued at %.3f, synced at %.3f\n", dbg_setup, dbg_loop, dbg_ms());
    float ms_total = 0.f;
    cudaEventElapsedTime(&ms_total, scene->ev[0], scene->ev[1]);
    S.passes = (uint32_t)pass;
    S.vertices = h_counters[CNT_VERTICES];
    S.rays_shadow = h_counters[CNT_SHADOW] - h_counters[CNT_SQ_HOLES];  // the queue length counts k_shade's unused reservations
    S.rays_light = h_counters[CNT_LIGHTRAYS];
    S.primary_hits = h_counters[CNT_PRIMARY_HITS];
    S.light_sweeps = h_counters[CNT_SWEEPS];
    S.light_evals = h_counters[CNT_EVALS];
    overflow = h_counters[CNT_OVERFLOW] != 0;
    S.ms_total = ms_total; S.ms_bounce = ms_bounce; S.ms_other = ms_total - ms_bounce;
    S.ms_shade = ms_shade; S.ms_trace = ms_bounce - ms_shade;
    S.ms_sweep = ms_sweep; S.ms_prep = ms_prep;
    for (auto& e : evm) if (e) cudaEventDestroy(e);
    if (stats) *stats = S;
    if (overflow) throw QueueOverflow{batch};
}
}  // namespace

void render_device(McptScene* scene, const McptCamera* cam, const McptRenderParams* prm, float* d_out_rgb, cudaStream_t st,
                   McptStats* stats)
{
    if (prm->flags & MCPT_PARITY_FP64) { render_parity_device(scene, cam, prm, d_out_rgb, st, stats); return; }
    // shade_with_mis grows a path TREE (F6).  Its queues hold 3 entries per path slot, which the reference scenes never
    // fill; a scene that does (light samples mostly occluded, so both children of most vertices live on) is retried here
    // with half the pass and twice the room per slot -- same memory, and a bit-identical film, because nothing depends on
    // the pass size (per-slot radiance is computed inside a pass, the film adds the samples one by one in sample order).  MCPT_TREE_FACTOR (percent, testing only) sets the first attempt's room.
    int factor = 300;
    if (const char* e = std::getenv("MCPT_TREE_FACTOR")) { const int v = std::atoi(e); if (v > 100) factor = v; }
    // testing only: pretend the queue of the FIRST attempt holds no more than this many entries, so that the retry path runs
    size_t pretend = 0;
    if (const char* e = std::getenv("MCPT_TEST_OVERFLOW_ABOVE")) { const long v = std::atol(e); if (v > 0) pretend = (size_t)v; }
    McptRenderParams p = *prm;
    for (;;) {
        try {
            render_wavefront(scene, cam, &p, d_out_rgb, st, stats, factor, pretend);
            return;
        } catch (const QueueOverflow& q) {
            if (factor >= 9600) throw Error(MCPT_ERR_QUEUE_OVERFLOW, "MIS path tree outgrew the ray queue at 96 entries per path slot");
            factor *= 2;
            pretend = 0;
            p.batch_spp = std::max(1, q.batch / 2);
        }
    }
}

}  // namespace mcpt
