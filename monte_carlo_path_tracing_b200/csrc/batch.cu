// csrc/batch.cu -- deterministic parity entry points: explicit inputs/uniforms in, values out.
// Each runs the same device functions the renderer uses (precision 1 = FP32 instantiation) or their
// FP64 instantiation in the reference's expression order (precision 0); tests compare both with the
// oracle (SURVEY.md 8c "Deterministic parity targets").
#include <algorithm>
#include <cstring>
#include "brdf.cuh"
#include "lights.cuh"
#include "mcpt_internal.h"
#include "traverse.cuh"

namespace mcpt {
namespace {

constexpr int kBlock = 128;

template <class T> struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    explicit DevBuf(size_t n_) : n(n_) { if (n) MCPT_CUDA(cudaMalloc((void**)&p, n * sizeof(T))); }
    DevBuf(const T* host, size_t n_) : DevBuf(n_) { if (n) MCPT_CUDA(cudaMemcpy(p, host, n * sizeof(T), cudaMemcpyHostToDevice)); }
    ~DevBuf() { if (p) cudaFree(p); }
    void to_host(T* host) const { if (n && host) MCPT_CUDA(cudaMemcpy(host, p, n * sizeof(T), cudaMemcpyDeviceToHost)); }
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
};

inline unsigned grid_for(size_t n) { return (unsigned)((n + kBlock - 1) / kBlock); }
inline void finish() { MCPT_CUDA(cudaGetLastError()); MCPT_CUDA(cudaDeviceSynchronize()); }

__device__ __forceinline__ d3 ld3(const double* p, size_t i) { return mk<double>(p[3 * i], p[3 * i + 1], p[3 * i + 2]); }
__device__ __forceinline__ void st3(double* p, size_t i, d3 v) { p[3 * i] = v.x; p[3 * i + 1] = v.y; p[3 * i + 2] = v.z; }
template <class T> __device__ __forceinline__ V3<T> as(d3 v) { return mk<T>((T)v.x, (T)v.y, (T)v.z); }
template <class T> __device__ __forceinline__ d3 wide(V3<T> v) { return mk<double>((double)v.x, (double)v.y, (double)v.z); }

// ---- intersection ------------------------------------------------------------------------------------
__global__ void k_intersect_tri(DScene s, size_t n, const double* ro, const double* rd, const int* prim, int precision,
                                int* hit, double* bgt)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    int p = prim[i];
    const float4 a = s.pverts[3 * p], b = s.pverts[3 * p + 1], c = s.pverts[3 * p + 2];
    d3 o = ld3(ro, i), d = ld3(rd, i);
    double be = 0, ga = 0, tt = 0;
    bool ok;
    if (precision == 0) ok = tri_test_exact(o, d, to_d3(a), to_d3(b), to_d3(c), &be, &ga, &tt);
    else {
        float fb, fg, ft;
        ok = tri_test_fast(to_f3(o), to_f3(d), to_f3(a), to_f3(b), to_f3(c), 1e-8f, &fb, &fg, &ft);
        if (ok) { be = fb; ga = fg; tt = ft; }
    }
    hit[i] = ok ? 1 : 0;
    bgt[3 * i] = ok ? be : 0; bgt[3 * i + 1] = ok ? ga : 0; bgt[3 * i + 2] = ok ? tt : 0;
}

__global__ void __launch_bounds__(kBlock) k_intersect(DScene s, size_t n, const double* ro, const double* rd, const int* exclude,
                                                      int light_only, int precision, int* prim, double* bgt)
{
    __shared__ int stack[kStackSm * kBlock];
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4* nodes = light_only ? s.lnodes : s.nodes;
    const float4* tris = light_only ? s.ltris : s.tris;
    d3 o = ld3(ro, i), d = ld3(rd, i);
    int ex = exclude ? exclude[i] : -1;
    if (precision == 0) {
        HitD h = trace_exact<false>(nodes, tris, o, d, ex, 1.7976931348623157e308, stack + threadIdx.x, kBlock);
        prim[i] = h.prim;
        bgt[3 * i] = h.prim >= 0 ? h.b : 0; bgt[3 * i + 1] = h.prim >= 0 ? h.g : 0; bgt[3 * i + 2] = h.prim >= 0 ? h.t : 0;
    } else {
        Hit h = trace_fast<false>(nodes, tris, to_f3(o), to_f3(d), ex, s.tmin, 3.0e38f, stack + threadIdx.x, kBlock);
        prim[i] = h.prim;
        bgt[3 * i] = h.prim >= 0 ? h.b : 0; bgt[3 * i + 1] = h.prim >= 0 ? h.g : 0; bgt[3 * i + 2] = h.prim >= 0 ? h.t : 0;
    }
}

// ---- Phong -----------------------------------------------------------------------------------------------
template <class T> __device__ __forceinline__ PhongMat<T> ldmat(const double* kd, const double* ks, const double* ns, size_t i)
{
    PhongMat<T> m;
    m.kd = as<T>(ld3(kd, i)); m.ks = as<T>(ld3(ks, i)); m.ns = (T)ns[i];
    return m;
}

template <class T>
__global__ void k_phong_eval(size_t n, const double* N, const double* wi, const double* wr, const double* kd, const double* ks,
                             const double* ns, double* rgb)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    st3(rgb, i, wide(phong_eval<T>(as<T>(ld3(N, i)), as<T>(ld3(wi, i)), as<T>(ld3(wr, i)), ldmat<T>(kd, ks, ns, i))));
}

// mode 0: phong_sample<T>; mode 1 (T = float only): phong_sample_fast, the renderer's variant
template <class T>
__global__ void k_phong_sample(size_t n, const double* N, const double* wr, const double* kd, const double* ks, const double* ns,
                               const double* u3, int fast, double* dir, double* pdf)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    T p;
    V3<T> d;
    PhongMat<T> m = ldmat<T>(kd, ks, ns, i);
    d = phong_sample<T>(as<T>(ld3(N, i)), as<T>(ld3(wr, i)), m, (T)u3[3 * i], (T)u3[3 * i + 1], (T)u3[3 * i + 2], &p);
    st3(dir, i, wide(d));
    pdf[i] = (double)p;
}
__global__ void k_phong_sample_fast(size_t n, const double* N, const double* wr, const double* kd, const double* ks,
                                    const double* ns, const double* u3, double* dir, double* pdf)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    float p;
    PhongMat<float> m = ldmat<float>(kd, ks, ns, i);
    f3 d = phong_sample_fast(as<float>(ld3(N, i)), as<float>(ld3(wr, i)), m, (float)u3[3 * i], (float)u3[3 * i + 1],
                             (float)u3[3 * i + 2], &p);
    st3(dir, i, wide(d));
    pdf[i] = (double)p;
}

template <class T>
__global__ void k_phong_pdf(size_t n, const double* N, const double* wi, const double* wr, const double* kd, const double* ks,
                            const double* ns, double* pdf)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    pdf[i] = (double)phong_pdf<T>(as<T>(ld3(N, i)), as<T>(ld3(wi, i)), as<T>(ld3(wr, i)), ldmat<T>(kd, ks, ns, i));
}

// ---- lights ----------------------------------------------------------------------------------------------
// libstdc++ discrete_distribution over weights w[0..n): one draw u, first index whose cumulative
// normalised weight >= u (last forced to 1); <2 weights: no draw, index 0 (bits/random.tcc:2657-2714).
template <class T, class W>
__device__ __forceinline__ int discrete_pick(W weight, int n, const double* u, int& ui, T* prob)
{
    if (n < 2) { *prob = T(1); return 0; }
    T sum = T(0);
    for (int i = 0; i < n; i++) sum += weight(i);
    T uu = (T)u[ui++];
    T cp = T(0);
    int idx = n - 1;
    for (int i = 0; i < n; i++) {
        T p = weight(i) / sum;
        cp = (i == 0) ? p : cp + p;
        T c = (i == n - 1) ? T(1) : cp;
        if (c >= uu) { idx = i; break; }
    }
    *prob = weight(idx) / sum;
    return idx;
}

template <class T>
__global__ void k_light_uniform(DScene s, size_t n, const double* u4, int* prim, double* point, double* Le, double* pdf)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double* u = u4 + 4 * i;
    int ui = 0;
    T pc, pt;
    int c = discrete_pick<T>([&](int k) { return (T)(s.crad64[3 * k] + s.crad64[3 * k + 1] + s.crad64[3 * k + 2]); }, s.C, u, ui, &pc);
    int b = s.class_begin[c], e = s.class_begin[c + 1];
    int a = discrete_pick<T>([&](int k) { return (T)s.larea64[b + k]; }, e - b, u, ui, &pt);
    T xi1 = (T)u[ui++];
    T xi2 = (T)u[ui++];
    LightTri<T> lt = load_light<T>(s, b + a);
    V3<T> p = uniform_point<T>(lt, xi1, xi2);
    prim[i] = lt.prim;
    st3(point, i, wide(p));
    st3(Le, i, wide(lt.Le));
    pdf[i] = (double)(pc * pt * (T(1) / lt.area));
}

// staged spherical-triangle sampling of one vertex: prepare (all lights), pick, Arvo sample, re-intersect,
// pdf of the query prim (Mylight.cpp:322-493).  One thread per vertex, two passes over the lights.
template <class T, bool DIRECT = false>
__global__ void k_light_spherical(DScene s, size_t n, const double* x1, const double* N, const double* u3, const int* query,
                                  int* prim, double* point, double* Le, double* pdf, double* wsum_out, int* nsurv,
                                  double* query_pdf)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    V3<T> x = as<T>(ld3(x1, i)), nn = as<T>(ld3(N, i));
    T wsum = T(0);
    int cnt = 0;
    bool query_in = false;
    int q = query ? query[i] : -1;
    for (int l = 0; l < s.L; l++) {
        LightTri<T> lt = load_light<T>(s, l);
        SphTri<T> st = spherical_project<T>(lt, x, nn);
        if (st.ok) { wsum += st.w; cnt++; if (lt.prim == q) query_in = true; }
    }
    if (wsum_out) wsum_out[i] = (double)wsum;
    if (nsurv) nsurv[i] = cnt;
    if (query_pdf) {
        double qp = 0;
        if (query_in && !(M<T>::abs_(wsum) < M<T>::eps)) {  // Mylight.cpp:484-493
            int ql = __float_as_int(s.pverts[3 * q + 1].w);
            LightTri<T> lt = load_light<T>(s, ql);
            qp = (double)(lt.lsum / wsum);
        }
        query_pdf[i] = qp;
    }
    if (cnt == 0 || M<T>::abs_(wsum) < M<T>::eps) {  // Mylight.cpp:263-266 / :427-430 dummy
        prim[i] = 0;
        st3(point, i, wide(nn * T(-1) + x));
        st3(Le, i, mk<double>(0, 0, 0));
        pdf[i] = 1.0;
        return;
    }
    const double* u = u3 + 3 * i;
    int ui = 0;
    // discrete pick over the survivors in light order
    int chosen = -1;
    SphTri<T> cst;
    LightTri<T> clt;
    T uu = (cnt >= 2) ? (T)u[ui++] : T(0);
    T cp = T(0);
    int k = 0;
    for (int l = 0; l < s.L; l++) {
        LightTri<T> lt = load_light<T>(s, l);
        SphTri<T> st = spherical_project<T>(lt, x, nn);
        if (!st.ok) continue;
        T p = st.w / wsum;
        cp = (k == 0) ? p : cp + p;
        T c = (k == cnt - 1) ? T(1) : cp;
        k++;
        if (cnt < 2 || c >= uu) { chosen = l; cst = st; clt = lt; break; }
    }
    T xi1 = (T)u[ui++];
    T xi2 = (T)u[ui++];
    V3<T> Pd = arvo_sample<T>(cst, xi1, xi2);
    if (DIRECT) {
        // what the renderer runs (render.cu): solid angle from the FP32 Van Oosterom-Strackee weight, direction from
        // arvo_direct<double>
        const f3 xf = as<float>(ld3(x1, i)), nf = as<float>(ld3(N, i));
        const float4 l0 = s.lights[5 * chosen], l1 = s.lights[5 * chosen + 1], l2 = s.lights[5 * chosen + 2];
        const float sA = light_weight_fast(s, chosen, xf, nf) / s.lights[5 * chosen + 3].w;
        d3 w = to_d3(nf);
        arvo_direct<double>(to_d3(l0), to_d3(l1), to_d3(l2), to_d3(xf), to_d3(nf), (double)sA, (double)xi1, (double)xi2, &w);
        Pd = as<T>(w);
    }
    // Mylight.cpp:311,317: point = x1 + P * t with t from intersect_with_triangle (0 on a miss)
    double be, ga, tt = 0;
    bool ok = tri_test_exact(wide(x), wide(Pd), wide(clt.v0), wide(clt.v1), wide(clt.v2), &be, &ga, &tt);
    if (!ok) tt = 0;
    prim[i] = clt.prim;
    st3(point, i, wide(x + Pd * (T)tt));
    st3(Le, i, wide(clt.Le));
    pdf[i] = (double)(clt.lsum / wsum);
    (void)chosen;
}

// FP32 renderer weights (Van Oosterom-Strackee) for the same vertices: wsum only
__global__ void k_light_wsum_fast(DScene s, size_t n, const double* x1, const double* N, double* wsum_out, int* nsurv)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    f3 x = as<float>(ld3(x1, i)), nn = as<float>(ld3(N, i));
    float wsum = 0.f;
    int cnt = 0;
    for (int l = 0; l < s.L; l++) {
        float w = light_weight_fast(s, l, x, nn);
        if (w > 0.f) { wsum += w; cnt++; }
    }
    wsum_out[i] = wsum;
    if (nsurv) nsurv[i] = cnt;
}

// light hierarchy (MCPT_LIGHT_TREE, lights.cuh): one draw + the selection probability / FP32 weight of a queried light
__global__ void k_light_tree(DScene s, size_t n, const double* __restrict__ x1, const double* __restrict__ N, const double* __restrict__ u,
                             const int* __restrict__ query, int* light, double* prob, double* qprob, double* qweight)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const f3 x = mk<float>((float)x1[3 * i], (float)x1[3 * i + 1], (float)x1[3 * i + 2]);
    const f3 nn = mk<float>((float)N[3 * i], (float)N[3 * i + 1], (float)N[3 * i + 2]);
    float p = 0.f;
    light[i] = lt_pick(s, x, nn, (float)u[i], &p);
    prob[i] = (double)p;
    const int q = query ? query[i] : -1;
    qprob[i] = (q >= 0 && q < s.L) ? (double)lt_prob(s, q, x, nn) : 0.0;
    qweight[i] = (q >= 0 && q < s.L) ? (double)light_weight_fast(s, q, x, nn) : 0.0;
}

// explicit FP64 rays -> the FP32 ray records of the wavefront queues (o | exclude, d | index)
__global__ void k_pack_rays(size_t n, const double* ro, const double* rd, const int* exclude, float4* org, float4* dir)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    org[i] = make_float4((float)ro[3 * i], (float)ro[3 * i + 1], (float)ro[3 * i + 2], __int_as_float(exclude ? exclude[i] : -1));
    dir[i] = make_float4((float)rd[3 * i], (float)rd[3 * i + 1], (float)rd[3 * i + 2], __int_as_float((int)i));
}
__global__ void k_unpack_hits(size_t n, const float4* hit, int* prim, double* bgt)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 h = hit[i];
    const int p = __float_as_int(h.x);
    prim[i] = p;
    bgt[3 * i] = p >= 0 ? h.y : 0; bgt[3 * i + 1] = p >= 0 ? h.z : 0; bgt[3 * i + 2] = p >= 0 ? h.w : 0;
}
__global__ void k_pack_ids(size_t n, const int* exclude, const int* light_prim, int2* ids)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < n) ids[i] = make_int2(exclude ? exclude[i] : -1, light_prim[i]);
}

// ---- FP32 ceiling of this device, measured (bench.py's roofline denominator for the shading kernels) ----------------------
// Register-resident multiply-add chains, 8 independent per thread, no memory traffic.  PACKED = the FFMA2 form the light
// sweeps use (two FMAs per instruction on a 64-bit register pair); scalar = three-register FFMA.
template <bool PACKED>
__global__ void __launch_bounds__(256) k_fp32_peak(float* out, int iters, float b, float c)
{
    float2 a[8];
#pragma unroll
    for (int k = 0; k < 8; k++) a[k] = make_float2(threadIdx.x * 1e-3f + k, threadIdx.x * 2e-3f - k);
    const float2 bb = make_float2(b, b + 1e-7f), cc = make_float2(c, c - 1e-7f);
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 8; k++) {
            if (PACKED) a[k] = __ffma2_rn(a[k], bb, cc);
            else { a[k].x = __fmaf_rn(a[k].x, bb.x, cc.x); a[k].y = __fmaf_rn(a[k].y, bb.y, cc.y); }
        }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; k++) s += a[k].x + a[k].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

void need(const McptScene* s) { if (!s) throw Error(MCPT_ERR_INVALID, "null scene"); }

}  // namespace
}  // namespace mcpt

using namespace mcpt;

extern "C" {

int mcpt_intersect_tri_batch(McptScene* s, size_t n, const double* ro, const double* rd, const int32_t* prim, int precision,
                             int32_t* hit, double* bgt)
{
    return guarded([&] {
        need(s);
        MCPT_CUDA(cudaSetDevice(s->device));
        for (size_t i = 0; i < n; i++)
            if (prim[i] < 0 || (uint32_t)prim[i] >= s->P) throw Error(MCPT_ERR_INVALID, "prim index out of range");
        DevBuf<double> dro(ro, 3 * n), drd(rd, 3 * n), dbgt(3 * n);
        DevBuf<int> dprim(prim, n), dhit(n);
        if (n) k_intersect_tri<<<grid_for(n), kBlock>>>(s->d, n, dro.p, drd.p, dprim.p, precision, dhit.p, dbgt.p);
        finish();
        dhit.to_host(hit); dbgt.to_host(bgt);
    });
}

int mcpt_intersect_batch(McptScene* s, size_t n, const double* ro, const double* rd, const int32_t* exclude, int light_only,
                         int precision, int32_t* prim, double* bgt)
{
    return guarded([&] {
        need(s);
        MCPT_CUDA(cudaSetDevice(s->device));
        DevBuf<double> dro(ro, 3 * n), drd(rd, 3 * n), dbgt(3 * n);
        DevBuf<int> dex(exclude, exclude ? n : 0), dprim(n);
        if (n && precision == 2) {
            // the wavefront's MCPT_LEAF_EXACT kernel itself (render.cu: k_extend_exact) on the FP32 roundings of the rays
            if (light_only) throw Error(MCPT_ERR_INVALID, "precision 2 (k_extend_exact) traces the scene BVH only");
            if (n > (size_t)1 << 30) throw Error(MCPT_ERR_INVALID, "too many rays in one batch");
            DevBuf<float4> org(n), dir(n), hit(n);
            k_pack_rays<<<grid_for(n), kBlock>>>(n, dro.p, drd.p, exclude ? dex.p : nullptr, org.p, dir.p);
            MCPT_CUDA(cudaGetLastError());
            trace_rays_device(s, (int)n, org.p, dir.p, MCPT_LEAF_EXACT, hit.p);
            k_unpack_hits<<<grid_for(n), kBlock>>>(n, hit.p, dprim.p, dbgt.p);
        } else if (n && precision == 1 && !light_only) {
            // the renderer's own trace kernel (render.cu: k_extend over the 4-wide BVH, or k_extend8 over the compressed 8-wide
            // BVH, whichever this scene uses) on the FP32 roundings of the rays
            if (n > (size_t)1 << 30) throw Error(MCPT_ERR_INVALID, "too many rays in one batch");
            DevBuf<float4> org(n), dir(n), hit(n);
            k_pack_rays<<<grid_for(n), kBlock>>>(n, dro.p, drd.p, exclude ? dex.p : nullptr, org.p, dir.p);
            MCPT_CUDA(cudaGetLastError());
            trace_rays_device(s, (int)n, org.p, dir.p, MCPT_LEAF_FAST, hit.p);
            k_unpack_hits<<<grid_for(n), kBlock>>>(n, hit.p, dprim.p, dbgt.p);
        } else if (n) k_intersect<<<grid_for(n), kBlock>>>(s->d, n, dro.p, drd.p, exclude ? dex.p : nullptr, light_only, precision, dprim.p, dbgt.p);
        finish();
        dprim.to_host(prim); dbgt.to_host(bgt);
    });
}

int mcpt_measure_fp32(int device, double* tflops_scalar, double* tflops_packed)
{
    return guarded([&] {
        int ndev = 0;
        if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); throw Error(MCPT_ERR_NO_DEVICE, "no CUDA device available"); }
        if (device < 0 || device >= ndev) throw Error(MCPT_ERR_INVALID, "device index out of range");
        MCPT_CUDA(cudaSetDevice(device));
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        const int grid = sms * 8, iters = 1 << 14;
        DevBuf<float> out((size_t)grid * 256);
        cudaEvent_t e0, e1;
        MCPT_CUDA(cudaEventCreate(&e0)); MCPT_CUDA(cudaEventCreate(&e1));
        double best[2] = {0, 0};
        for (int packed = 0; packed < 2; packed++)
            for (int rep = 0; rep < 4; rep++) {  // first repetition warms up
                MCPT_CUDA(cudaEventRecord(e0));
                if (packed) k_fp32_peak<true><<<grid, 256>>>(out.p, iters, 0.999f, 1e-3f);
                else k_fp32_peak<false><<<grid, 256>>>(out.p, iters, 0.999f, 1e-3f);
                MCPT_CUDA(cudaEventRecord(e1));
                MCPT_CUDA(cudaEventSynchronize(e1));
                float ms = 0.f;
                cudaEventElapsedTime(&ms, e0, e1);
                const double flops = 2.0 * 16.0 * (double)iters * (double)grid * 256.0;  // 16 FMAs per thread per iteration
                if (rep) best[packed] = std::max(best[packed], flops / (ms * 1e-3) / 1e12);
            }
        cudaEventDestroy(e0); cudaEventDestroy(e1);
        MCPT_CUDA(cudaGetLastError());
        if (tflops_scalar) *tflops_scalar = best[0];
        if (tflops_packed) *tflops_packed = best[1];
    });
}

int mcpt_visibility_batch(McptScene* s, size_t n, const double* ro, const double* rd, const int32_t* exclude, const int32_t* light_prim,
                          int leaf_mode, int32_t* visible)
{
    return guarded([&] {
        need(s);
        if (!ro || !rd || !light_prim || !visible) throw Error(MCPT_ERR_INVALID, "null argument");
        if (leaf_mode != MCPT_LEAF_FAST && leaf_mode != MCPT_LEAF_EXACT) throw Error(MCPT_ERR_INVALID, "unknown leaf_mode");
        if (n > (size_t)1 << 30) throw Error(MCPT_ERR_INVALID, "too many rays in one batch");
        for (size_t i = 0; i < n; i++)
            if (light_prim[i] < 0 || (uint32_t)light_prim[i] >= s->P) throw Error(MCPT_ERR_INVALID, "light prim index out of range");
        MCPT_CUDA(cudaSetDevice(s->device));
        DevBuf<double> dro(ro, 3 * n), drd(rd, 3 * n);
        DevBuf<int> dex(exclude, exclude ? n : 0), dlp(light_prim, n), dvis(n);
        DevBuf<float4> org(n), dir(n);
        DevBuf<int2> ids(n);
        if (n) {
            k_pack_rays<<<grid_for(n), kBlock>>>(n, dro.p, drd.p, exclude ? dex.p : nullptr, org.p, dir.p);
            k_pack_ids<<<grid_for(n), kBlock>>>(n, exclude ? dex.p : nullptr, dlp.p, ids.p);
            MCPT_CUDA(cudaGetLastError());
            visibility_device(s, (int)n, org.p, dir.p, ids.p, leaf_mode, dvis.p);
        }
        finish();
        dvis.to_host(visible);
    });
}

int mcpt_visibility_noalloc(McptScene* s, size_t n, const void* d_org, const void* d_dir, const void* d_ids, int leaf_mode, void* d_visible,
                            void* stream)
{
    return guarded([&] {
        need(s);
        if (leaf_mode != MCPT_LEAF_FAST && leaf_mode != MCPT_LEAF_EXACT) throw Error(MCPT_ERR_INVALID, "unknown leaf_mode");
        if (n && (!d_org || !d_dir || !d_ids || !d_visible)) throw Error(MCPT_ERR_INVALID, "null argument");
        visibility_noalloc_device(s, n, (const float4*)d_org, (const float4*)d_dir, (const int2*)d_ids, leaf_mode, (int*)d_visible, (cudaStream_t)stream);
    });
}

int mcpt_phong_eval_batch(McptScene* s, size_t n, const double* N, const double* wi, const double* wr, const double* kd,
                          const double* ks, const double* ns, int precision, double* rgb)
{
    return guarded([&] {
        need(s);
        MCPT_CUDA(cudaSetDevice(s->device));
        DevBuf<double> a(N, 3 * n), b(wi, 3 * n), c(wr, 3 * n), d(kd, 3 * n), e(ks, 3 * n), f(ns, n), o(3 * n);
        if (n) {
            if (precision == 0) k_phong_eval<double><<<grid_for(n), kBlock>>>(n, a.p, b.p, c.p, d.p, e.p, f.p, o.p);
            else k_phong_eval<float><<<grid_for(n), kBlock>>>(n, a.p, b.p, c.p, d.p, e.p, f.p, o.p);
        }
        finish();
        o.to_host(rgb);
    });
}

int mcpt_phong_sample_batch(McptScene* s, size_t n, const double* N, const double* wr, const double* kd, const double* ks,
                            const double* ns, const double* u3, int precision, double* dir, double* pdf)
{
    return guarded([&] {
        need(s);
        MCPT_CUDA(cudaSetDevice(s->device));
        DevBuf<double> a(N, 3 * n), c(wr, 3 * n), d(kd, 3 * n), e(ks, 3 * n), f(ns, n), u(u3, 3 * n), od(3 * n), op(n);
        if (n) {
            if (precision == 0) k_phong_sample<double><<<grid_for(n), kBlock>>>(n, a.p, c.p, d.p, e.p, f.p, u.p, 0, od.p, op.p);
            else if (precision == 1) k_phong_sample_fast<<<grid_for(n), kBlock>>>(n, a.p, c.p, d.p, e.p, f.p, u.p, od.p, op.p);
            else k_phong_sample<float><<<grid_for(n), kBlock>>>(n, a.p, c.p, d.p, e.p, f.p, u.p, 0, od.p, op.p);
        }
        finish();
        od.to_host(dir); op.to_host(pdf);
    });
}

int mcpt_phong_pdf_batch(McptScene* s, size_t n, const double* N, const double* wi, const double* wr, const double* kd,
                         const double* ks, const double* ns, int precision, double* pdf)
{
    return guarded([&] {
        need(s);
        MCPT_CUDA(cudaSetDevice(s->device));
        DevBuf<double> a(N, 3 * n), b(wi, 3 * n), c(wr, 3 * n), d(kd, 3 * n), e(ks, 3 * n), f(ns, n), o(n);
        if (n) {
            if (precision == 0) k_phong_pdf<double><<<grid_for(n), kBlock>>>(n, a.p, b.p, c.p, d.p, e.p, f.p, o.p);
            else k_phong_pdf<float><<<grid_for(n), kBlock>>>(n, a.p, b.p, c.p, d.p, e.p, f.p, o.p);
        }
        finish();
        o.to_host(pdf);
    });
}

int mcpt_light_sample_uniform_batch(McptScene* s, size_t n, const double* u4, int precision, int32_t* prim, double* point,
                                    double* Le, double* pdf)
{
    return guarded([&] {
        need(s);
        if (s->L == 0) throw Error(MCPT_ERR_INVALID, "scene has no lights");
        MCPT_CUDA(cudaSetDevice(s->device));
        DevBuf<double> u(u4, 4 * n), pt(3 * n), le(3 * n), pd(n);
        DevBuf<int> pr(n);
        if (n) {
            if (precision == 0) k_light_uniform<double><<<grid_for(n), kBlock>>>(s->d, n, u.p, pr.p, pt.p, le.p, pd.p);
            else k_light_uniform<float><<<grid_for(n), kBlock>>>(s->d, n, u.p, pr.p, pt.p, le.p, pd.p);
        }
        finish();
        pr.to_host(prim); pt.to_host(point); le.to_host(Le); pd.to_host(pdf);
    });
}

int mcpt_light_sample_spherical_batch(McptScene* s, size_t n, const double* x1, const double* N, const double* u3,
                                      const int32_t* query, int precision, int32_t* prim, double* point, double* Le, double* pdf,
                                      double* wsum, int32_t* nsurv, double* query_pdf)
{
    return guarded([&] {
        need(s);
        if (s->L == 0) throw Error(MCPT_ERR_INVALID, "scene has no lights");
        MCPT_CUDA(cudaSetDevice(s->device));
        DevBuf<double> x(x1, 3 * n), nn(N, 3 * n), u(u3, 3 * n), pt(3 * n), le(3 * n), pd(n), ws(n), qp(n);
        DevBuf<int> q(query, query ? n : 0), pr(n), ns(n);
        if (n) {
            if (precision == 0)
                k_light_spherical<double><<<grid_for(n), kBlock>>>(s->d, n, x.p, nn.p, u.p, query ? q.p : nullptr, pr.p, pt.p, le.p, pd.p, ws.p, ns.p, qp.p);
            else if (precision == 3)  // reference pick (FP64), direction from the renderer's sampler arvo_direct
                k_light_spherical<double, true><<<grid_for(n), kBlock>>>(s->d, n, x.p, nn.p, u.p, query ? q.p : nullptr, pr.p, pt.p, le.p, pd.p, ws.p, ns.p, qp.p);
            else if (precision == 2)  // renderer weights only (Van Oosterom-Strackee): wsum, nsurv
                k_light_wsum_fast<<<grid_for(n), kBlock>>>(s->d, n, x.p, nn.p, ws.p, ns.p);
            else
                k_light_spherical<float><<<grid_for(n), kBlock>>>(s->d, n, x.p, nn.p, u.p, query ? q.p : nullptr, pr.p, pt.p, le.p, pd.p, ws.p, ns.p, qp.p);
        }
        finish();
        if (precision != 2) { pr.to_host(prim); pt.to_host(point); le.to_host(Le); pd.to_host(pdf); qp.to_host(query_pdf); }
        ws.to_host(wsum); ns.to_host(nsurv);
    });
}

int mcpt_light_tree_batch(McptScene* s, size_t n, const double* x1, const double* N, const double* u, const int32_t* query_light,
                          int32_t* light, double* prob, double* query_prob, double* query_weight)
{
    return guarded([&] {
        need(s);
        if (s->L == 0) throw Error(MCPT_ERR_INVALID, "scene has no lights");
        MCPT_CUDA(cudaSetDevice(s->device));
        DevBuf<double> x(x1, 3 * n), nn(N, 3 * n), uu(u, n), pr(n), qp(n), qw(n);
        DevBuf<int> q(query_light, query_light ? n : 0), li(n);
        if (n) k_light_tree<<<grid_for(n), kBlock>>>(s->d, n, x.p, nn.p, uu.p, query_light ? q.p : nullptr, li.p, pr.p, qp.p, qw.p);
        finish();
        li.to_host(light); pr.to_host(prob); qp.to_host(query_prob); qw.to_host(query_weight);
    });
}

}  // extern "C"
