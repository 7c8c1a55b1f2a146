// csrc/render_parity.cu -- MCPT_PARITY_FP64: mcpt_render's PARITY mode.
//
// The FP32 wavefront of render.cu is the throughput path; its hit ids and BRDF / light values agree with the reference
// only to FP32 rounding.  This translation unit runs the SAME loop -- main.cpp:557-588 with shade (main.cpp:269-344),
// shade_with_brdf (:348-399) or shade_with_mis (:402-494) -- entirely in FP64 and in the reference's own expression
// order: camera ray (:563-564), exhaustive-search closest hit through the FP64 Cramer test of Myobj.cpp:165-192
// (trace_exact), vertex and shading normal (:26-29,273-274), Phong eval / sample / pdf (BRDF.cpp:17-132, brdf.cuh
// T = double), the per-vertex loop over ALL light triangles with nine acos per survivor and libstdc++'s
// discrete_distribution pick in light order (Mylight.cpp:329-482, lights.cuh T = double), Arvo's SampleTriangle and
// the re-intersection that turns the direction into a point (:289-311), roulette (:429-437), and the film sum
// `sum += I * (1/spp)` in sample order (:575-576).  The file is compiled with -fmad=false, so no multiply-add is
// contracted (the g++ x86-64 build of the reference has none either): every value differs from the reference's only by
// the last-bit differences between CUDA's and glibc's acos / sin / cos / pow.
//
// Uniforms are the renderer's keyed Philox draws (rng.cuh) widened to double: (seed; pixel, sample, path-tree node,
// call) -> call 0 = (roulette, u_lobe, xi1, xi2) of BRDF.cpp:41-50, call 1 = (u_tri, xi1, xi2) of Mylight.cpp:435-450 or
// (u_class, u_tri, xi1, xi2) of :120-146.  The oracle reads the same table under ORACLE_RNG_KEYED, which makes a
// PARITY film comparable with the oracle's pixel by pixel (tests/test_gpu_parity_render.py), not just statistically.
//
// One thread walks the whole path tree of one (pixel, sample) depth first with an explicit stack -- the reference's
// recursion order.  It is orders of magnitude slower than the wavefront and is meant to be: it is the renderer's
// proof of parity, and the yardstick the FP32 wavefront is measured against at full size on the GPU.
#include <algorithm>
#include <cmath>
#include <cstring>
#include "brdf.cuh"
#include "lights.cuh"
#include "mcpt_internal.h"
#include "rng.cuh"
#include "traverse.cuh"

namespace mcpt {
namespace {

constexpr int kPB = 128;      // threads per block
constexpr int kPStack = 64;   // pending path-tree nodes per thread: one per level; P(depth >= 64) = 0.6^64 ~ 6e-15
constexpr double kDblMax = 1.7976931348623157e308;

enum { PC_VERTICES = 0, PC_RAYS, PC_SHADOW, PC_LIGHTRAYS, PC_PRIMARY_HITS, PC_OVERFLOW, PC_N };

struct PCam { double eye[3], U[3], V[3], N[3], pixellen, wlen, ci, cj; int W, H; };

// a hit waiting to be shaded: shade*(point, wo) of main.cpp:269/348/402 with the path throughput in front of it
struct PItem { d3 wo, thr; double b, g; uint64_t node; int prim, depth; };

struct PArgs {
    DScene s;
    PCam cam;
    double* Lacc;  // [slots][3]
    unsigned long long* counters;
    uint64_t seed;
    double p_rr;
    int npix, batch, sample_base, max_depth;
    uint32_t flags;
};

__device__ __forceinline__ PhongMat<double> p_load_mat(const DScene& s, int m)
{
    const float4 a = __ldg(s.mats + 2 * m), b = __ldg(s.mats + 2 * m + 1);
    PhongMat<double> r;
    r.kd = to_d3(a); r.ns = (double)a.w; r.ks = to_d3(b);
    return r;
}

// linear_interpolation, main.cpp:26-29: (v0 * a + v1 * b) + v2 * c
__device__ __forceinline__ d3 p_lerp(const float4* v, double a, double b, double c)
{
    return (to_d3(__ldg(v)) * a + to_d3(__ldg(v + 1)) * b) + to_d3(__ldg(v + 2)) * c;
}

__device__ __forceinline__ HitD p_closest(const DScene& s, d3 o, d3 d, int exclude, int* stk)
{
    return trace_exact<false>(s.nodes, s.tris, o, d, exclude, kDblMax, stk, kPB);
}

// prepared_for_lights_spherical_triangle_sampling, Mylight.cpp:329-421: weights_sum and the number of survivors
__device__ __forceinline__ void p_light_prepare(const DScene& s, d3 x, d3 n, double* wsum, int* cnt)
{
    double sum = 0;
    int c = 0;
    for (int l = 0; l < s.L; l++) {
        const SphTri<double> st = spherical_project<double>(load_light<double>(s, l), x, n);
        if (st.ok) { sum += st.w; c++; }
    }
    *wsum = sum; *cnt = c;
}

// lights_spherical_triangle_sampling, Mylight.cpp:424-482: discrete_distribution over the survivors in light order
// (first index whose cumulative normalised weight >= u, last forced to 1, no draw below two weights), Arvo sample,
// re-intersection (:311; a miss leaves t = 0).  Returns false for the dummy of :427-430.
struct PLightSample { int prim; d3 coord, Le; double prob; };
__device__ __forceinline__ bool p_light_sample(const DScene& s, d3 x, d3 n, double wsum, int cnt, double u, double xi1, double xi2,
                                               PLightSample* out)
{
    if (cnt == 0 || fabs(wsum) < 1e-8) return false;
    double cp = 0;
    int k = 0;
    for (int l = 0; l < s.L; l++) {
        const LightTri<double> lt = load_light<double>(s, l);
        const SphTri<double> st = spherical_project<double>(lt, x, n);
        if (!st.ok) continue;
        const double p = st.w / wsum;
        cp = (k == 0) ? p : cp + p;
        const double c = (k == cnt - 1) ? 1.0 : cp;
        k++;
        if (cnt < 2 || c >= u) {
            const d3 Pd = arvo_sample<double>(st, xi1, xi2);
            double be, ga, tt = 0;
            if (!tri_test_exact(x, Pd, lt.v0, lt.v1, lt.v2, &be, &ga, &tt)) tt = 0;
            out->prim = lt.prim; out->coord = x + Pd * tt; out->Le = lt.Le; out->prob = lt.lsum / wsum;
            return true;
        }
    }
    return false;  // unreachable: the last survivor's cumulative weight is forced to 1
}

// eval_spherical_triangle_sampling_pdf, Mylight.cpp:484-493, evaluated for THIS vertex (F10 is never reproduced)
__device__ __forceinline__ double p_light_pdf(const DScene& s, int prim, d3 x, d3 n, double wsum)
{
    const int l = __float_as_int(__ldg(s.pverts + 3 * prim + 1).w);
    if (l < 0) return 0;
    const LightTri<double> lt = load_light<double>(s, l);
    if (!spherical_project<double>(lt, x, n).ok) return 0;
    if (fabs(wsum) < 1e-8) return 0;
    return lt.lsum / wsum;
}

// select_a_point_from_lights, Mylight.cpp:102-160 (pdf in AREA measure)
__device__ __forceinline__ PLightSample p_light_uniform(const DScene& s, double u_class, double u_tri, double xi1, double xi2)
{
    auto pick = [](auto weight, int n, double u, double* prob) {
        if (n < 2) { *prob = 1.0; return 0; }
        double sum = 0;
        for (int i = 0; i < n; i++) sum += weight(i);
        double cp = 0;
        int idx = n - 1;
        for (int i = 0; i < n; i++) {
            const double p = weight(i) / sum;
            cp = (i == 0) ? p : cp + p;
            const double c = (i == n - 1) ? 1.0 : cp;
            if (c >= u) { idx = i; break; }
        }
        *prob = weight(idx) / sum;
        return idx;
    };
    double pc, pt;
    const int c = pick([&](int k) { return s.crad64[3 * k] + s.crad64[3 * k + 1] + s.crad64[3 * k + 2]; }, s.C, u_class, &pc);
    const int b = s.class_begin[c], e = s.class_begin[c + 1];
    const int a = pick([&](int k) { return s.larea64[b + k]; }, e - b, u_tri, &pt);
    const LightTri<double> lt = load_light<double>(s, b + a);
    PLightSample r;
    double proba = 1.0;
    proba *= pc; proba *= pt;
    r.coord = uniform_point<double>(lt, xi1, xi2);
    proba *= 1.0 / lt.area;
    r.prim = lt.prim; r.Le = lt.Le; r.prob = proba;
    return r;
}

__device__ __forceinline__ d3 p_u4(uint64_t seed, uint32_t pixel, uint32_t sample, uint64_t node, uint32_t call, double* u3)
{
    const Rng4 r = rng_draw4(seed, pixel, sample, node, call);
    *u3 = (double)r.u[3];
    return mk<double>((double)r.u[0], (double)r.u[1], (double)r.u[2]);
}

template <int STRAT>
__global__ void __launch_bounds__(kPB) k_parity(const PArgs a)
{
    constexpr bool kNee = (STRAT == MCPT_NEE_SPHERICAL || STRAT == MCPT_NEE_UNIFORM);
    constexpr bool kMis = (STRAT == MCPT_MIS);
    __shared__ int stack[kStackSm * kPB];
    int* stk = stack + threadIdx.x;
    const DScene& s = a.s;
    PItem items[kPStack];
    unsigned long long c_vert = 0, c_rays = 0, c_shadow = 0, c_lrays = 0, c_phits = 0, c_over = 0;
    const size_t slots = (size_t)a.npix * (size_t)a.batch;
    for (size_t slot = (size_t)blockIdx.x * kPB + threadIdx.x; slot < slots; slot += (size_t)gridDim.x * kPB) {
        const uint32_t pixel = (uint32_t)(slot % (size_t)a.npix);
        const uint32_t sample = (uint32_t)(a.sample_base + (int)(slot / (size_t)a.npix));
        // camera ray, main.cpp:563-564
        const int pi = (int)pixel / a.cam.W, pj = (int)pixel - pi * a.cam.W;
        double di = 0, dj = 0;
        if (a.flags & MCPT_JITTER) {  // extension of F4, same draw as k_raygen
            const Rng4 r = rng_draw4(a.seed, pixel, sample, 1, 2);
            di = (double)r.u[0] - 0.5; dj = (double)r.u[1] - 0.5;
        }
        const double dx = -a.cam.pixellen * (((double)pi + di) - a.cam.ci), dy = a.cam.pixellen * (((double)pj + dj) - a.cam.cj);
        const double dz = a.cam.wlen;
        const d3 eye = mk<double>(a.cam.eye[0], a.cam.eye[1], a.cam.eye[2]);
        const d3 dir = normalized(mk<double>((a.cam.U[0] * dx + a.cam.V[0] * dy) + a.cam.N[0] * dz,
                                             (a.cam.U[1] * dx + a.cam.V[1] * dy) + a.cam.N[1] * dz,
                                             (a.cam.U[2] * dx + a.cam.V[2] * dy) + a.cam.N[2] * dz));
        d3 L = mk<double>(0, 0, 0);
        int sp = 0;
        {
            const HitD h = p_closest(s, eye, dir, -1, stk);  // main.cpp:572
            c_rays++;
            if (h.prim >= 0) {
                c_phits++;
                PItem& it = items[sp++];
                it.wo = dir * -1.0; it.thr = mk<double>(1, 1, 1); it.b = h.b; it.g = h.g; it.node = 1; it.prim = h.prim; it.depth = 0;
            }
        }
        while (sp > 0) {
            const PItem it = items[--sp];
            // ---- shade prologue, main.cpp:273-288 / :351-366 / :406-421 ------------------------------------------------
            const double al = 1.0 - it.b - it.g;
            const d3 p = p_lerp(s.pverts + 3 * it.prim, al, it.b, it.g);
            const d3 N = normalized(p_lerp(s.pnorms + 3 * it.prim, al, it.b, it.g));
            const d3 wo = it.wo;
            if (dot(N, wo) < 0) continue;
            const int li = __float_as_int(__ldg(s.pverts + 3 * it.prim + 1).w);
            if (li >= 0) {
                const int cls = __float_as_int(__ldg(s.lights + 5 * li + 1).w);
                L = L + it.thr * mk<double>(s.crad64[3 * cls], s.crad64[3 * cls + 1], s.crad64[3 * cls + 2]);
                continue;
            }
            const PhongMat<double> mat = p_load_mat(s, __float_as_int(__ldg(s.pverts + 3 * it.prim).w));
            double r0w, r1w;
            const d3 r0 = p_u4(a.seed, pixel, sample, it.node, 0, &r0w);  // roulette | u_lobe | xi1 ; r0w = xi2
            d3 r1 = mk<double>(0.5, 0.5, 0.5);
            r1w = 0.5;
            if (STRAT != MCPT_BRDF_ONLY) r1 = p_u4(a.seed, pixel, sample, it.node, 1, &r1w);
            const int depth = it.depth + 1;
            const bool rr_forced = a.max_depth > 0 && depth >= a.max_depth;
            const bool rr_ok = !rr_forced && !(r0.x > a.p_rr);

            if (kNee) {
                // ---- direct term of shade(), main.cpp:290-316 ------------------------------------------------------------
                PLightSample lp;
                bool have;
                if (STRAT == MCPT_NEE_UNIFORM) { lp = p_light_uniform(s, r1.x, r1.y, r1.z, r1w); have = true; }
                else {
                    double wsum; int cnt;
                    p_light_prepare(s, p, N, &wsum, &cnt);
                    have = p_light_sample(s, p, N, wsum, cnt, r1.x, r1.y, r1.z, &lp);
                }
                if (have) {
                    const int ll = __float_as_int(__ldg(s.pverts + 3 * lp.prim + 1).w);
                    const d3 n1 = mk<double>(s.lng64[3 * ll], s.lng64[3 * ll + 1], s.lng64[3 * ll + 2]);
                    const d3 wl = normalized(lp.coord - p);
                    if (dot(wl, N) > 0 && dot(wl * -1.0, n1) > 0) {
                        const HitD hs = p_closest(s, p, wl, it.prim, stk);  // main.cpp:309
                        c_shadow++;
                        if (hs.prim == lp.prim) {
                            const d3 f = phong_eval<double>(N, wl, wo, mat);
                            double g;
                            if (STRAT == MCPT_NEE_SPHERICAL && (a.flags & MCPT_FIX_NEE_MEASURE)) g = dot(wl, N) / lp.prob;
                            else g = dot(wl, N) * dot(wl * -1.0, n1) / dot(lp.coord - p, lp.coord - p) / lp.prob;  // :314 (F9)
                            L = L + it.thr * ((lp.Le * f) * g);
                        }
                    }
                }
                c_vert++;
                if (!rr_ok) continue;  // main.cpp:321-329
            } else {
                c_vert++;
                if (!rr_ok) continue;  // main.cpp:375-383 / :429-437: roulette before any sampling
            }

            // ---- light-sample branch of shade_with_mis(), main.cpp:439-466 ------------------------------------------------
            double wsum = 0;
            PItem child_l;
            bool have_l = false;
            if (kMis) {
                int cnt;
                p_light_prepare(s, p, N, &wsum, &cnt);
                PLightSample lp;
                if (p_light_sample(s, p, N, wsum, cnt, r1.x, r1.y, r1.z, &lp)) {
                    const d3 wl = normalized(lp.coord - p);
                    if (dot(wl, N) > 0) {
                        const HitD hl = p_closest(s, p, wl, it.prim, stk);  // :455
                        c_rays++;
                        if (hl.prim >= 0) {
                            const d3 f = phong_eval<double>(N, wl, wo, mat);
                            const double pp = phong_pdf<double>(N, wl, wo, mat);
                            child_l.wo = wl * -1.0; child_l.thr = (it.thr * f) * (dot(wl, N) / (lp.prob + pp) / a.p_rr);
                            child_l.b = hl.b; child_l.g = hl.g; child_l.node = it.node * 2; child_l.prim = hl.prim; child_l.depth = depth;
                            have_l = true;
                        }
                    }
                }
            }

            // ---- BRDF-sample branch, main.cpp:331-341 / :385-395 / :468-491 ------------------------------------------------
            PItem child_b;
            bool have_b = false;
            {
                double pdf;
                const d3 wi = phong_sample<double>(N, wo, mat, r0.y, r0.z, r0w, &pdf);
                if (!(dot(wi, N) < 0)) {
                    if (a.flags & MCPT_FIX_MIXTURE_PDF) pdf = phong_pdf<double>(N, wi, wo, mat);  // F8
                    const HitD hb = p_closest(s, p, wi, it.prim, stk);
                    c_rays++;
                    bool recurse = hb.prim >= 0;
                    if (kNee && recurse && __float_as_int(__ldg(s.pverts + 3 * hb.prim + 1).w) >= 0) recurse = false;  // main.cpp:338
                    if (recurse) {
                        const d3 f = phong_eval<double>(N, wi, wo, mat);
                        double lpdf = 0;
                        if (kMis) {  // :484-489: nearest EMISSIVE triangle along wi, occluders ignored
                            const HitD hq = trace_exact<false>(s.lnodes, s.ltris, p, wi, it.prim, kDblMax, stk, kPB);
                            c_lrays++;
                            if (hq.prim >= 0) lpdf = p_light_pdf(s, hq.prim, p, N, wsum);
                        }
                        child_b.wo = wi * -1.0; child_b.thr = (it.thr * f) * (dot(wi, N) / (pdf + lpdf) / a.p_rr);
                        child_b.b = hb.b; child_b.g = hb.g; child_b.node = kMis ? it.node * 2 + 1 : (uint64_t)depth + 1;
                        child_b.prim = hb.prim; child_b.depth = depth;
                        have_b = true;
                    }
                }
            }
            // depth first, light-sample child first (recursion #1 of main.cpp:464 runs before #2 of :491)
            if (have_b) { if (sp < kPStack) items[sp++] = child_b; else c_over++; }
            if (have_l) { if (sp < kPStack) items[sp++] = child_l; else c_over++; }
        }
        a.Lacc[3 * slot] = L.x; a.Lacc[3 * slot + 1] = L.y; a.Lacc[3 * slot + 2] = L.z;
    }
    // one atomic per counter per warp
    auto flush = [&](unsigned long long v, int which) {
#pragma unroll
        for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0 && v) atomicAdd(a.counters + which, v);
    };
    flush(c_vert, PC_VERTICES); flush(c_rays, PC_RAYS); flush(c_shadow, PC_SHADOW); flush(c_lrays, PC_LIGHTRAYS);
    flush(c_phits, PC_PRIMARY_HITS); flush(c_over, PC_OVERFLOW);
}

// film, main.cpp:566,575-576,582: sum = sum + I * (1/spp) over the samples of this pass, in sample order
__global__ void k_parity_accumulate(const double* __restrict__ Lacc, int npix, int batch, double inv, double* film, int first_pass)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npix) return;
    double sx = first_pass ? 0.0 : film[3 * p], sy = first_pass ? 0.0 : film[3 * p + 1], sz = first_pass ? 0.0 : film[3 * p + 2];
    for (int k = 0; k < batch; k++) {
        const double* v = Lacc + 3 * ((size_t)k * npix + p);
        sx = sx + v[0] * inv; sy = sy + v[1] * inv; sz = sz + v[2] * inv;
    }
    film[3 * p] = sx; film[3 * p + 1] = sy; film[3 * p + 2] = sz;
}

__global__ void k_parity_finalize(const double* __restrict__ film, int npix, int accumulate, float* out_rgb)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 3 * npix) return;
    const float v = (float)film[i];
    out_rgb[i] = accumulate ? out_rgb[i] + v : v;
}

}  // namespace

void render_parity_device(McptScene* scene, const McptCamera* cam, const McptRenderParams* prm, float* d_out_rgb, cudaStream_t st,
                          McptStats* stats)
{
    const int W = prm->width, H = prm->height;
    if (W <= 0 || H <= 0 || prm->spp <= 0) throw Error(MCPT_ERR_INVALID, "width, height and spp must be positive");
    if (prm->strategy < 0 || prm->strategy > 3) throw Error(MCPT_ERR_INVALID, "MCPT_PARITY_FP64 runs the reference's estimators only (strategy 0..3)");
    if (prm->flags & MCPT_LIGHT_TREE) throw Error(MCPT_ERR_INVALID, "MCPT_LIGHT_TREE is not a reference estimator: no PARITY mode");
    if (prm->strategy != MCPT_BRDF_ONLY && scene->L == 0) throw Error(MCPT_ERR_INVALID, "light sampling needs at least one light");
    const size_t npix = (size_t)W * H;
    if (npix > (1u << 30)) throw Error(MCPT_ERR_INVALID, "image too large");
    const int spp_total = prm->spp_total > 0 ? prm->spp_total : prm->spp;
    MCPT_CUDA(cudaSetDevice(scene->device));
    // passes of at most 8 Mi path slots (24 B each)
    int batch = prm->batch_spp > 0 ? prm->batch_spp : (int)std::max<size_t>(1, ((size_t)8 << 20) / npix);
    batch = std::min(batch, prm->spp);
    const size_t slots = npix * (size_t)batch;
    size_t off = 0;
    auto place = [&](size_t bytes) { size_t at = off; off = (off + bytes + 255) & ~(size_t)255; return at; };
    const size_t o_lacc = place(slots * 24), o_film = place(npix * 24), o_cnt = place(PC_N * 8);
    Workspace& ws = scene->ws;
    if (ws.bytes < off) {
        if (ws.base) MCPT_CUDA(cudaFree(ws.base));
        ws.base = nullptr; ws.bytes = 0;
        if (cudaMalloc(&ws.base, off) != cudaSuccess) { cudaGetLastError(); throw Error(MCPT_ERR_OOM, "cannot allocate render workspace; lower batch_spp"); }
        ws.bytes = off;
    }
    if (!ws.pinned) MCPT_CUDA(cudaMallocHost(&ws.pinned, 256));
    char* base = (char*)ws.base;
    double* Lacc = (double*)(base + o_lacc);
    double* film = (double*)(base + o_film);
    unsigned long long* d_counters = (unsigned long long*)(base + o_cnt);
    unsigned long long* h_counters = (unsigned long long*)((char*)ws.pinned + 64);

    PArgs a;
    a.s = scene->d;
    {   // camera basis, main.cpp:546-553, FP64 in the reference's order
        auto nrm = [](double* v) { double l = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); v[0] /= l; v[1] /= l; v[2] /= l; };
        auto crs = [](const double* x, const double* y, double* o) { o[0] = x[1] * y[2] - x[2] * y[1]; o[1] = x[2] * y[0] - x[0] * y[2]; o[2] = x[0] * y[1] - x[1] * y[0]; };
        double N[3] = {cam->w[0], cam->w[1], cam->w[2]}, V[3], U[3];
        const double wl = std::sqrt(N[0] * N[0] + N[1] * N[1] + N[2] * N[2]);
        if (!(wl > 0)) throw Error(MCPT_ERR_INVALID, "camera w must be non-zero");
        nrm(N); crs(N, cam->up, V); nrm(V); crs(V, N, U); nrm(U);
        for (int k = 0; k < 3; k++) { a.cam.eye[k] = cam->eye[k]; a.cam.U[k] = U[k]; a.cam.V[k] = V[k]; a.cam.N[k] = N[k]; }
        a.cam.pixellen = cam->pixellen; a.cam.wlen = wl;
        a.cam.ci = (H - 1) * 0.5; a.cam.cj = (W - 1) * 0.5;
        a.cam.W = W; a.cam.H = H;
    }
    a.Lacc = Lacc; a.counters = d_counters; a.seed = prm->seed;
    a.p_rr = prm->p_rr > 0.f ? (double)prm->p_rr : 0.6;  // main.cpp:321,375,429: the literal 0.6, as a double
    a.npix = (int)npix; a.max_depth = prm->max_depth; a.flags = prm->flags;

    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, scene->device);
    McptStats S;
    std::memset(&S, 0, sizeof S);
    S.paths = (uint64_t)npix * (uint64_t)prm->spp;
    MCPT_CUDA(cudaEventRecord(scene->ev[0], st));
    MCPT_CUDA(cudaMemsetAsync(d_counters, 0, PC_N * 8, st));
    int done = 0, pass = 0;
    while (done < prm->spp) {
        const int b = std::min(batch, prm->spp - done);
        a.batch = b; a.sample_base = prm->spp_offset + done;
        const size_t n = npix * (size_t)b;
        const int grid = (int)std::min<size_t>((n + kPB - 1) / kPB, (size_t)sms * 16);
        switch (prm->strategy) {
            case MCPT_NEE_SPHERICAL: k_parity<MCPT_NEE_SPHERICAL><<<grid, kPB, 0, st>>>(a); break;
            case MCPT_BRDF_ONLY: k_parity<MCPT_BRDF_ONLY><<<grid, kPB, 0, st>>>(a); break;
            case MCPT_MIS: k_parity<MCPT_MIS><<<grid, kPB, 0, st>>>(a); break;
            default: k_parity<MCPT_NEE_UNIFORM><<<grid, kPB, 0, st>>>(a); break;
        }
        MCPT_CUDA(cudaGetLastError());
        k_parity_accumulate<<<(unsigned)((npix + 255) / 256), 256, 0, st>>>(Lacc, (int)npix, b, 1.0 / spp_total, film, pass == 0);
        S.kernel_launches += 2;
        done += b;
        pass++;
    }
    k_parity_finalize<<<(unsigned)((3 * npix + 255) / 256), 256, 0, st>>>(film, (int)npix, (prm->flags & MCPT_ACCUMULATE) ? 1 : 0, d_out_rgb);
    S.kernel_launches++;
    MCPT_CUDA(cudaMemcpyAsync(h_counters, d_counters, PC_N * 8, cudaMemcpyDeviceToHost, st));
    MCPT_CUDA(cudaEventRecord(scene->ev[1], st));
    MCPT_CUDA(cudaStreamSynchronize(st));
    MCPT_CUDA(cudaGetLastError());
    float ms_total = 0.f;
    cudaEventElapsedTime(&ms_total, scene->ev[0], scene->ev[1]);
    S.passes = (uint32_t)pass;
    S.vertices = h_counters[PC_VERTICES];
    S.rays_closest = h_counters[PC_RAYS];
    S.rays_shadow = h_counters[PC_SHADOW];
    S.rays_light = h_counters[PC_LIGHTRAYS];
    S.primary_hits = h_counters[PC_PRIMARY_HITS];
    S.ms_total = ms_total; S.ms_bounce = ms_total;
    S.elements = S.vertices;
    if (stats) *stats = S;
    if (h_counters[PC_OVERFLOW]) throw Error(MCPT_ERR_QUEUE_OVERFLOW, "PARITY mode: a path tree outgrew the per-thread stack of 64 pending nodes");
}

}  // namespace mcpt
