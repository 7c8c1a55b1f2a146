// csrc/lights.cuh -- area-triangle light sampling: uniform-area and Arvo spherical-triangle.
// Device restatement of Mylight.cpp:102-160 (select_a_point_from_lights) and :163-493 (spherical
// triangle sampling, one-shot and staged), templated on the real type like brdf.cuh.
#pragma once
#include "device_math.cuh"
#include "scene_dev.cuh"
#include "light_lists.h"

namespace mcpt {

template <class T> struct LightTri { V3<T> v0, v1, v2, ng, Le; T lsum, area; int prim, cls; };

template <class T> __device__ __forceinline__ LightTri<T> load_light(const DScene& s, int l);
template <> __device__ __forceinline__ LightTri<float> load_light<float>(const DScene& s, int l)
{
    const float4 a = __ldg(s.lights + 5 * l + 0), b = __ldg(s.lights + 5 * l + 1), c = __ldg(s.lights + 5 * l + 2);
    const float4 n = __ldg(s.lights + 5 * l + 3), e = __ldg(s.lights + 5 * l + 4);
    LightTri<float> t;
    t.v0 = to_f3(a); t.v1 = to_f3(b); t.v2 = to_f3(c); t.ng = to_f3(n); t.Le = to_f3(e);
    t.lsum = n.w; t.area = c.w; t.prim = __float_as_int(a.w); t.cls = __float_as_int(b.w);
    return t;
}
template <> __device__ __forceinline__ LightTri<double> load_light<double>(const DScene& s, int l)
{
    const float4 a = __ldg(s.lights + 5 * l + 0), b = __ldg(s.lights + 5 * l + 1), c = __ldg(s.lights + 5 * l + 2);
    LightTri<double> t;
    t.v0 = to_d3(a); t.v1 = to_d3(b); t.v2 = to_d3(c);
    t.ng = mk<double>(s.lng64[3 * l], s.lng64[3 * l + 1], s.lng64[3 * l + 2]);
    t.prim = __float_as_int(a.w); t.cls = __float_as_int(b.w);
    t.Le = mk<double>(s.crad64[3 * t.cls], s.crad64[3 * t.cls + 1], s.crad64[3 * t.cls + 2]);
    t.lsum = t.Le.x + t.Le.y + t.Le.z;  // RadianceRGB::sum, RadianceRGB.cpp:70-73
    t.area = s.larea64[l];
    return t;
}

// ---- spherical triangle of one light as seen from x1 (Mylight.cpp:175-253 == :335-413) ----------
template <class T> struct SphTri { bool ok; V3<T> A, B, C; T alpha, c, sA, w; };

template <class T>
__device__ __forceinline__ SphTri<T> spherical_project(const LightTri<T>& lt, V3<T> x1, V3<T> n)
{
    const T eps = M<T>::eps;
    SphTri<T> r;
    r.ok = false; r.w = T(0); r.sA = T(0); r.alpha = T(0); r.c = T(0);
    T tmp = dot(lt.ng, x1 - lt.v0);
    if (tmp < T(0) || M<T>::abs_(tmp) < eps) return r;  // :180-185 x1 behind the light
    T t0 = dot(n, lt.v0 - x1), t1 = dot(n, lt.v1 - x1), t2 = dot(n, lt.v2 - x1);
    if ((t0 < T(0) || M<T>::abs_(t0) < eps) && (t1 < T(0) || M<T>::abs_(t1) < eps) && (t2 < T(0) || M<T>::abs_(t2) < eps))
        return r;  // :194 light entirely below the horizon
    V3<T> A = normalized(lt.v0 - x1), B = normalized(lt.v1 - x1), C = normalized(lt.v2 - x1);
    if (dot(cross(normalized(C - A), normalized(B - A)), n) < T(0)) { V3<T> t = B; B = C; C = t; }  // :206-211
    T a = M<T>::acos_(clamp1(dot(B, C)));
    T b = M<T>::acos_(clamp1(dot(A, C)));
    T c = M<T>::acos_(clamp1(dot(A, B)));
    if (a < eps || b < eps || c < eps) return r;  // :219
    T alpha = M<T>::acos_(clamp1(dot(-normalized(cross(B, A)), normalized(cross(A, C)))));
    T beta = M<T>::acos_(clamp1(dot(-normalized(cross(C, B)), normalized(cross(B, A)))));
    T gamma = M<T>::acos_(clamp1(dot(-normalized(cross(A, C)), normalized(cross(C, B)))));
    if (alpha < eps || beta < eps || gamma < eps) return r;  // :230
    T sA = alpha + beta + gamma - M<T>::pi;                   // :236
    if (sA < T(0)) return r;
    T w = sA * lt.lsum;                                       // :244
    if (w < T(0) || M<T>::bad_(w)) return r;
    r.ok = true; r.A = A; r.B = B; r.C = C; r.alpha = alpha; r.c = c; r.sA = sA; r.w = w;
    return r;
}

// Arvo SampleTriangle(xi1, xi2), Mylight.cpp:289-297: direction uniformly distributed in the
// spherical triangle (A,B,C).
template <class T>
__device__ __forceinline__ V3<T> arvo_sample(const SphTri<T>& st, T xi1, T xi2)
{
    T sA1 = xi1 * st.sA;
    T s, t, sa, ca;
    M<T>::sincos_(sA1 - st.alpha, &s, &t);
    M<T>::sincos_(st.alpha, &sa, &ca);
    T u = t - ca;
    T v = s + sa * M<T>::cos_(st.c);
    T q = ((v * t - u * s) * ca - v) / ((v * s + u * t) * sa);
    V3<T> C1 = normalized(st.A * q + normalized(st.C - st.A * dot(st.C, st.A)) * M<T>::sqrt_(T(1) - q * q));
    T z = T(1) - xi2 * (T(1) - dot(C1, st.B));
    return normalized(st.B * z + normalized(C1 - st.B * dot(C1, st.B)) * M<T>::sqrt_(T(1) - z * z));
}

// ---- FP32 weight of one light for the renderer's O(L) loop -------------------------------------------
// Same culls (:180-197) and the same quantity sA * (R+G+B), with the solid angle from the
// Van Oosterom-Strackee form  tan(sA/2) = |a.(b x c)| / (|a||b||c| + (a.b)|c| + (a.c)|b| + (b.c)|a|)
// where a,b,c = vertices - x1.  a.(b x c) = (v0-x1).((v1-v0) x (v2-v0)) = 2 area * ng.(v0-x1) has no
// cancellation, so tiny lights keep full FP32 relative accuracy where alpha+beta+gamma-pi has none.
//
// The loop reads the SoA table `lsoa` (scene_dev.cuh, light_lists.h): five float4 planes per block of 32 lights, so a
// warp's 32 lanes fetch 32 consecutive 16-byte vectors per plane (5 x 512 B, fully coalesced, one pointer + constant offsets):
//   plane0 = v0, plane1 = e1 = v1 - v0, plane2 = e2 = v2 - v0, plane3 = -Ns = -2 area * ng | 2 Lsum,
//   plane4 = e1.e1 | e2.e2 | e1.e2 | cull threshold 1e-8 * 2 area;  padding entries have threshold = +inf and evaluate to 0.
// Branch-free on purpose.

// atan(q) for q in [0,1]: Abramowitz & Stegun 4.4.49, |error| <= 2e-8; exact to first order as q -> 0
#define MCPT_ATAN_C0 0.0028662257f
#define MCPT_ATAN_C1 -0.0161657367f
#define MCPT_ATAN_C2 0.0429096138f
#define MCPT_ATAN_C3 -0.0752896400f
#define MCPT_ATAN_C4 0.1065626393f
#define MCPT_ATAN_C5 -0.1420889944f
#define MCPT_ATAN_C6 0.1999355085f
#define MCPT_ATAN_C7 -0.3333314528f
__device__ __forceinline__ float atan_unit(float q)
{
    const float s = __fmul_rn(q, q);
    float p = MCPT_ATAN_C0;
    p = __fmaf_rn(p, s, MCPT_ATAN_C1);
    p = __fmaf_rn(p, s, MCPT_ATAN_C2);
    p = __fmaf_rn(p, s, MCPT_ATAN_C3);
    p = __fmaf_rn(p, s, MCPT_ATAN_C4);
    p = __fmaf_rn(p, s, MCPT_ATAN_C5);
    p = __fmaf_rn(p, s, MCPT_ATAN_C6);
    p = __fmaf_rn(p, s, MCPT_ATAN_C7);
    return __fmaf_rn(q, __fmul_rn(p, s), q);
}
__device__ __forceinline__ float rcp_fast(float x)  // MUFU.RCP
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
// |v| from v.v as x * MUFU.RSQ(x).  MUFU.SQRT would save the (packed) multiply but measured slower: veach-mis 1663 instead
// of 1697 Msamples/s -- the sweep is co-limited by the special-function unit (8 MUFU per light and vertex pair).
__device__ __forceinline__ float len_fast(float x) { return __fmul_rn(x, rsqrtf(x)); }
// the octant fix-ups of atan2(y, x) for y > 0 (result in (0, pi)), given r = atan(min/max) of (|x|, y)
__device__ __forceinline__ float atan2_fix(float r, float y, float x)
{
    r = (y > fabsf(x)) ? 1.57079632679489662f - r : r;
    return (x < 0.f) ? 3.14159265358979324f - r : r;
}

// One light, one vertex.  Written with explicit round-to-nearest intrinsics in exactly the operation order of the
// packed version below (light_weight_pair), so the short second pass of warp_light_pick reproduces the sweep's
// weights bit for bit.  p3 = -2 area ng | 2 Lsum.
__device__ __forceinline__ float dot_rn(float ax, float ay, float az, float bx, float by, float bz)
{
    return __fmaf_rn(az, bz, __fmaf_rn(ay, by, __fmul_rn(ax, bx)));
}
// Edge form: with e1 = v1 - v0, e2 = v2 - v0 (precomputed per light, with e1.e1, e2.e2, e1.e2) and a = v0 - x1, the
// other two corner vectors are b = a + e1, c = a + e2, so
//   a.b = a.a + a.e1,  a.c = a.a + a.e2,  b.c = a.a + a.e1 + a.e2 + e1.e2,  b.b = a.a + 2 a.e1 + e1.e1,  c.c likewise:
// three dot products (a.a, a.e1, a.e2) and eight additions stand for the six subtractions and five dot products of the
// direct form -- 47 instead of 54 packed FP32 operations per light and vertex pair.  The horizon cull needs
// n.b = n.a + n.e1 and n.c = n.a + n.e2 only when the first vertex does not settle it.  p4 = e1.e1 | e2.e2 | e1.e2 | threshold.
__device__ __forceinline__ float light_weight_soa(float4 p0, float4 p1, float4 p2, float4 p3, float4 p4, f3 x1, f3 n)
{
    const float ax = __fadd_rn(p0.x, -x1.x), ay = __fadd_rn(p0.y, -x1.y), az = __fadd_rn(p0.z, -x1.z);
    const float num = dot_rn(p3.x, p3.y, p3.z, ax, ay, az);            // 2 area * (ng . (x1 - v0)), :178-185
    const float ha = dot_rn(n.x, n.y, n.z, ax, ay, az);
    const float hb = __fadd_rn(ha, dot_rn(n.x, n.y, n.z, p1.x, p1.y, p1.z)), hc = __fadd_rn(ha, dot_rn(n.x, n.y, n.z, p2.x, p2.y, p2.z));
    const bool up = ha >= 1e-8f || fmaxf(hb, hc) >= 1e-8f;             // :187-197
    const float aa = dot_rn(ax, ay, az, ax, ay, az);
    const float ae1 = dot_rn(ax, ay, az, p1.x, p1.y, p1.z), ae2 = dot_rn(ax, ay, az, p2.x, p2.y, p2.z);
    const float ab = __fadd_rn(aa, ae1), ac = __fadd_rn(aa, ae2);
    const float bb = __fmaf_rn(2.f, ae1, __fadd_rn(aa, p4.x)), cc = __fmaf_rn(2.f, ae2, __fadd_rn(aa, p4.y));
    const float bc = __fadd_rn(__fadd_rn(ab, ae2), p4.z);
    const float la = len_fast(aa), lb = len_fast(bb), lc = len_fast(cc);
    float den = __fmul_rn(__fmul_rn(la, lb), lc);
    den = __fmaf_rn(ab, lc, den);
    den = __fmaf_rn(ac, lb, den);
    den = __fmaf_rn(bc, la, den);
    const float ad = fabsf(den);
    const float q = __fmul_rn(fminf(ad, num), rcp_fast(fmaxf(ad, num)));
    const float h = atan2_fix(atan_unit(q), num, den);        // half the solid angle
    const bool ok = num >= p4.w && up && h < CUDART_INF_F;    // h < inf also rejects NaN; 0 adds nothing
    return ok ? __fmul_rn(h, p3.w) : 0.f;
}

// lsoa is blocked by 32 lights: [Lpad / 32][kLsoaPlanes][32] float4, so one pointer with constant offsets serves a sweep
__host__ __device__ __forceinline__ int lsoa_index(int l) { return (l >> 5) * kLsoaBlock + (l & 31); }

__host__ __device__ __forceinline__ int lsoa_light(int off) { return (off / kLsoaBlock) * 32 + (off % kLsoaBlock); }  // inverse of lsoa_index

// weight of the light at lsoa offset `off` (what the survivor lists of light_lists.h store)
__device__ __forceinline__ float light_weight_at(const DScene& s, int off, f3 x1, f3 n)
{
    const float4* P = s.lsoa + off;
    return light_weight_soa(__ldg(P), __ldg(P + 32), __ldg(P + 64), __ldg(P + 96), __ldg(P + 128), x1, n);
}
__device__ __forceinline__ float light_weight_fast(const DScene& s, int l, f3 x1, f3 n) { return light_weight_at(s, lsoa_index(l), x1, n); }

// ---- packed FP32 (sm_100 FFMA2 / FADD2 / FMUL2) --------------------------------------------------------------
// Blackwell issues ONE instruction for two independent FP32 operations on a 64-bit register pair, and any operand
// may instead be a 32-bit register broadcast to both halves (SASS `R4.F32`) or an immediate.  The O(L) loop is
// issue-bound (ncu: 77 % issue-active at 29 threads/instruction in its scalar form), so a sweep evaluates one
// light against TWO VERTICES per instruction: the vertex pair lives in the two halves, the light's scalars and the
// polynomial constants are broadcast operands.  Only min/max, the three reciprocal square roots, the reciprocal
// (MUFU) and the final selects stay scalar.  Same operations in the same order as light_weight_soa.
__device__ __forceinline__ float2 bc2(float a) { return make_float2(a, a); }
__device__ __forceinline__ float2 len2_fast(float2 x) { return __fmul2_rn(x, make_float2(rsqrtf(x.x), rsqrtf(x.y))); }
__device__ __forceinline__ float2 dot2_rn(float2 ax, float2 ay, float2 az, float2 bx, float2 by, float2 bz)
{
    return __ffma2_rn(az, bz, __ffma2_rn(ay, by, __fmul2_rn(ax, bx)));
}
// two vertices side by side: -x1 and n per component, (vertex 0, vertex 1)
struct VertexPair2 { float2 mx, my, mz, nx, ny, nz; };
__device__ __forceinline__ VertexPair2 make_vertex2(f3 xa, f3 na, f3 xb, f3 nb)
{
    VertexPair2 v;
    v.mx = make_float2(-xa.x, -xb.x); v.my = make_float2(-xa.y, -xb.y); v.mz = make_float2(-xa.z, -xb.z);
    v.nx = make_float2(na.x, nb.x); v.ny = make_float2(na.y, nb.y); v.nz = make_float2(na.z, nb.z);
    return v;
}
struct LightRec { float4 p0, p1, p2, p3, p4; };
__device__ __forceinline__ LightRec load_light_rec(const float4* __restrict__ P)
{
    LightRec r;
    r.p0 = __ldg(P); r.p1 = __ldg(P + 32); r.p2 = __ldg(P + 64); r.p3 = __ldg(P + 96); r.p4 = __ldg(P + 128);
    return r;
}
__device__ __forceinline__ float2 light_weight_pair(const LightRec& L, const VertexPair2& v)
{
    const float4 p1 = L.p1, p2 = L.p2, p3 = L.p3, p4 = L.p4;
    const float2 ax = __fadd2_rn(bc2(L.p0.x), v.mx), ay = __fadd2_rn(bc2(L.p0.y), v.my), az = __fadd2_rn(bc2(L.p0.z), v.mz);
    const float2 num = dot2_rn(bc2(p3.x), bc2(p3.y), bc2(p3.z), ax, ay, az);
    // horizon cull (:187-197): the light counts iff ANY of its vertices is above the vertex's tangent plane.  The first
    // vertex alone settles that for nearly every light; the other two are only looked at when some lane needs them
    // (all 32 lanes are here together: the sweeps have warp-uniform trip counts).
    const float2 ha = dot2_rn(v.nx, v.ny, v.nz, ax, ay, az);
    bool up0 = ha.x >= 1e-8f, up1 = ha.y >= 1e-8f;
    if (__any_sync(0xffffffffu, !(up0 && up1))) {
        const float2 hb = __fadd2_rn(ha, dot2_rn(v.nx, v.ny, v.nz, bc2(p1.x), bc2(p1.y), bc2(p1.z)));
        const float2 hc = __fadd2_rn(ha, dot2_rn(v.nx, v.ny, v.nz, bc2(p2.x), bc2(p2.y), bc2(p2.z)));
        up0 = up0 || fmaxf(hb.x, hc.x) >= 1e-8f;
        up1 = up1 || fmaxf(hb.y, hc.y) >= 1e-8f;
    }
    const float2 aa = dot2_rn(ax, ay, az, ax, ay, az);
    const float2 ae1 = dot2_rn(ax, ay, az, bc2(p1.x), bc2(p1.y), bc2(p1.z)), ae2 = dot2_rn(ax, ay, az, bc2(p2.x), bc2(p2.y), bc2(p2.z));
    const float2 ab = __fadd2_rn(aa, ae1), ac = __fadd2_rn(aa, ae2);
    const float2 bb = __ffma2_rn(bc2(2.f), ae1, __fadd2_rn(aa, bc2(p4.x))), cc = __ffma2_rn(bc2(2.f), ae2, __fadd2_rn(aa, bc2(p4.y)));
    const float2 bcd = __fadd2_rn(__fadd2_rn(ab, ae2), bc2(p4.z));
    const float2 la = len2_fast(aa), lb = len2_fast(bb), lc = len2_fast(cc);
    float2 den = __fmul2_rn(__fmul2_rn(la, lb), lc);
    den = __ffma2_rn(ab, lc, den);
    den = __ffma2_rn(ac, lb, den);
    den = __ffma2_rn(bcd, la, den);
    const float adx = fabsf(den.x), ady = fabsf(den.y);
    const float2 q = __fmul2_rn(make_float2(fminf(adx, num.x), fminf(ady, num.y)),
                                make_float2(rcp_fast(fmaxf(adx, num.x)), rcp_fast(fmaxf(ady, num.y))));
    const float2 s = __fmul2_rn(q, q);
    float2 p = bc2(MCPT_ATAN_C0);
    p = __ffma2_rn(p, s, bc2(MCPT_ATAN_C1));
    p = __ffma2_rn(p, s, bc2(MCPT_ATAN_C2));
    p = __ffma2_rn(p, s, bc2(MCPT_ATAN_C3));
    p = __ffma2_rn(p, s, bc2(MCPT_ATAN_C4));
    p = __ffma2_rn(p, s, bc2(MCPT_ATAN_C5));
    p = __ffma2_rn(p, s, bc2(MCPT_ATAN_C6));
    p = __ffma2_rn(p, s, bc2(MCPT_ATAN_C7));
    const float2 r = __ffma2_rn(q, __fmul2_rn(p, s), q);
    // half the solid angle, 0 where the light is culled; the caller multiplies by 2 Lsum and accumulates in one FFMA2
    const float hx = atan2_fix(r.x, num.x, den.x), hy = atan2_fix(r.y, num.y, den.y);
    const bool ok0 = num.x >= p4.w && up0 && hx < CUDART_INF_F;
    const bool ok1 = num.y >= p4.w && up1 && hy < CUDART_INF_F;
    return make_float2(ok0 ? hx : 0.f, ok1 ? hy : 0.f);
}
// acc += weight of light L seen from the two vertices of v (weight = half solid angle x 2 Lsum)
__device__ __forceinline__ float2 light_weight_pair_acc(const LightRec& L, const VertexPair2& v, float2 acc)
{
    return __ffma2_rn(light_weight_pair(L, v), bc2(L.p3.w), acc);
}

// The transposed packing for scenes with few lights (k_lightsel's per-lane loop): TWO LIGHTS against one vertex per
// instruction.  L = 17 float2 (light 2j | light 2j+1) in the order v0.xyz, e1.xyz, e2.xyz, q.xyz, 2 Lsum, e1.e1, e2.e2, e1.e2,
// threshold (kLightPairFields; staged in shared memory by the kernel, every lane reads the same pair: broadcast).  The same
// operations in the same order as light_weight_soa, so the two weights equal light_weight_fast's bit for bit.  The horizon
// branch is per lane here (the lanes of a warp hold different vertices).
constexpr int kLightPairFields = 17;
__device__ __forceinline__ float2 light_weight_2lights(const float2* __restrict__ L, f3 x1, f3 n)
{
    const float2 mx = bc2(-x1.x), my = bc2(-x1.y), mz = bc2(-x1.z), nx = bc2(n.x), ny = bc2(n.y), nz = bc2(n.z);
    const float2 e1x = L[3], e1y = L[4], e1z = L[5], e2x = L[6], e2y = L[7], e2z = L[8];
    const float2 ax = __fadd2_rn(L[0], mx), ay = __fadd2_rn(L[1], my), az = __fadd2_rn(L[2], mz);
    const float2 num = dot2_rn(L[9], L[10], L[11], ax, ay, az);
    const float2 ha = dot2_rn(nx, ny, nz, ax, ay, az);
    bool up0 = ha.x >= 1e-8f, up1 = ha.y >= 1e-8f;
    if (!(up0 && up1)) {
        const float2 hb = __fadd2_rn(ha, dot2_rn(nx, ny, nz, e1x, e1y, e1z)), hc = __fadd2_rn(ha, dot2_rn(nx, ny, nz, e2x, e2y, e2z));
        up0 = up0 || fmaxf(hb.x, hc.x) >= 1e-8f;
        up1 = up1 || fmaxf(hb.y, hc.y) >= 1e-8f;
    }
    const float2 aa = dot2_rn(ax, ay, az, ax, ay, az);
    const float2 ae1 = dot2_rn(ax, ay, az, e1x, e1y, e1z), ae2 = dot2_rn(ax, ay, az, e2x, e2y, e2z);
    const float2 ab = __fadd2_rn(aa, ae1), ac = __fadd2_rn(aa, ae2);
    const float2 bb = __ffma2_rn(bc2(2.f), ae1, __fadd2_rn(aa, L[13])), cc = __ffma2_rn(bc2(2.f), ae2, __fadd2_rn(aa, L[14]));
    const float2 bcd = __fadd2_rn(__fadd2_rn(ab, ae2), L[15]);
    const float2 la = len2_fast(aa), lb = len2_fast(bb), lc = len2_fast(cc);
    float2 den = __fmul2_rn(__fmul2_rn(la, lb), lc);
    den = __ffma2_rn(ab, lc, den);
    den = __ffma2_rn(ac, lb, den);
    den = __ffma2_rn(bcd, la, den);
    const float adx = fabsf(den.x), ady = fabsf(den.y);
    const float2 q = __fmul2_rn(make_float2(fminf(adx, num.x), fminf(ady, num.y)),
                                make_float2(rcp_fast(fmaxf(adx, num.x)), rcp_fast(fmaxf(ady, num.y))));
    const float2 s = __fmul2_rn(q, q);
    float2 p = bc2(MCPT_ATAN_C0);
    p = __ffma2_rn(p, s, bc2(MCPT_ATAN_C1));
    p = __ffma2_rn(p, s, bc2(MCPT_ATAN_C2));
    p = __ffma2_rn(p, s, bc2(MCPT_ATAN_C3));
    p = __ffma2_rn(p, s, bc2(MCPT_ATAN_C4));
    p = __ffma2_rn(p, s, bc2(MCPT_ATAN_C5));
    p = __ffma2_rn(p, s, bc2(MCPT_ATAN_C6));
    p = __ffma2_rn(p, s, bc2(MCPT_ATAN_C7));
    const float2 r = __ffma2_rn(q, __fmul2_rn(p, s), q);
    const float hx = atan2_fix(r.x, num.x, den.x), hy = atan2_fix(r.y, num.y, den.y);
    const bool ok0 = num.x >= L[16].x && up0 && hx < CUDART_INF_F;
    const bool ok1 = num.y >= L[16].y && up1 && hy < CUDART_INF_F;
    const float2 w = __fmul2_rn(make_float2(hx, hy), L[12]);
    return make_float2(ok0 ? w.x : 0.f, ok1 ? w.y : 0.f);
}

// Per-lane partial sums, over this lane's lights (l = lane, lane + 32, ...), of the weights seen from V = 2 or 4
// vertices (one packed evaluation per light and vertex pair: the four 16-byte loads of a light serve all).  With `cdf` (shared memory,
// [2][Lpad] floats) every running partial is also stored: cdf[v][l] = sum of the lane's weights up to and
// including light l -- the per-lane columns k_lightprep inverts for its picks.  All 32 lanes call.
#ifndef MCPT_SWEEP_UNROLL
#define MCPT_SWEEP_UNROLL 2
#endif
constexpr int kSweepUnroll = MCPT_SWEEP_UNROLL;
template <int V>
__device__ __forceinline__ void warp_light_partial(const DScene& s, const f3 (&x1)[V], const f3 (&n)[V], float (&part)[V],
                                                   float* cdf = nullptr)
{
    static_assert(V == 2 || V == 4, "the sweep is packed over vertex pairs");
    const int lane = threadIdx.x & 31;
    VertexPair2 vp[V / 2];
    float2 acc[V / 2];
#pragma unroll
    for (int h = 0; h < V / 2; h++) { vp[h] = make_vertex2(x1[2 * h], n[2 * h], x1[2 * h + 1], n[2 * h + 1]); acc[h] = make_float2(0.f, 0.f); }
    const float4* __restrict__ P = s.lsoa + lane;
#pragma unroll kSweepUnroll
    for (int l = lane; l < s.Lpad; l += 32, P += kLsoaBlock) {  // Lpad is a multiple of 64
        const LightRec rec = load_light_rec(P);
#pragma unroll
        for (int h = 0; h < V / 2; h++) acc[h] = light_weight_pair_acc(rec, vp[h], acc[h]);
        if (V == 2 && cdf) { cdf[l] = acc[0].x; cdf[s.Lpad + l] = acc[0].y; }
    }
#pragma unroll
    for (int h = 0; h < V / 2; h++) { part[2 * h] = acc[h].x; part[2 * h + 1] = acc[h].y; }
}

// The same sweep over a survivor list (light_lists.h) instead of the whole table: `list` holds n lsoa offsets (n a multiple
// of 32, padded with a weight-0 light), lane k takes entries k, k + 32, ...  The lights a list leaves out are culled at
// every vertex of the primitive it belongs to, i.e. they would have added exact zeros.  cdf (optional): running
// sums by list POSITION, cdf[v * cdf_stride + k].
// SUB > 1: only every SUB-th running sum of a lane's column is kept (after list positions SUB-1, 2 SUB-1, ... of the lane), at
// cdf[v * cdf_stride + 32 m + lane] for the m-th checkpoint: the picker bisects the checkpoints and re-evaluates the <= SUB
// weights of one group (k_lightsel_sorted).  A quarter of the shared memory leaves the light table room in L1.
template <int V, int SUB = 1>
__device__ __forceinline__ void warp_light_partial_list(const DScene& s, const uint32_t* __restrict__ list, int n, const f3 (&x1)[V],
                                                        const f3 (&nn)[V], float (&part)[V], float* cdf = nullptr, int cdf_stride = 0)
{
    static_assert(V == 2 || V == 4, "the sweep is packed over vertex pairs");
    const int lane = threadIdx.x & 31;
    VertexPair2 vp[V / 2];
    float2 acc[V / 2];
#pragma unroll
    for (int h = 0; h < V / 2; h++) { vp[h] = make_vertex2(x1[2 * h], nn[2 * h], x1[2 * h + 1], nn[2 * h + 1]); acc[h] = make_float2(0.f, 0.f); }
    const uint32_t* __restrict__ E = list + lane;
    // the entry of the NEXT iteration is fetched one iteration ahead: the light's four loads depend on it (veach-mis: +2.4 %)
    uint32_t e_next = n > 0 ? __ldg(E) : 0u;
#pragma unroll kSweepUnroll
    for (int k = lane; k < n; k += 32, E += 32) {
        const float4* __restrict__ P = s.lsoa + e_next;
        if (k + 32 < n) e_next = __ldg(E + 32);
        const LightRec rec = load_light_rec(P);
#pragma unroll
        for (int h = 0; h < V / 2; h++) acc[h] = light_weight_pair_acc(rec, vp[h], acc[h]);
        if (cdf) {
            if (SUB == 1) {
#pragma unroll
                for (int h = 0; h < V / 2; h++) { cdf[(2 * h) * cdf_stride + k] = acc[h].x; cdf[(2 * h + 1) * cdf_stride + k] = acc[h].y; }
            } else if (((k >> 5) & (SUB - 1)) == SUB - 1) {
                const int at = ((k >> 5) / SUB << 5) + lane;
#pragma unroll
                for (int h = 0; h < V / 2; h++) { cdf[(2 * h) * cdf_stride + at] = acc[h].x; cdf[(2 * h + 1) * cdf_stride + at] = acc[h].y; }
            }
        }
    }
#pragma unroll
    for (int h = 0; h < V / 2; h++) { part[2 * h] = acc[h].x; part[2 * h + 1] = acc[h].y; }
}

// ---- Arvo SampleTriangle without inverse trigonometry ------------------------------------------------------
// Same map (xi1, xi2) -> direction as Mylight.cpp:200-211,289-297, but fed with cos/sin of the vertex angle alpha
// straight from the edge-plane normals and with the solid angle sA the weight loop already has (Van Oosterom-
// Strackee), instead of nine acos() and cos/sin of their results.  s = sin(xi1 sA - alpha) and t = cos(...) come
// from the angle-difference identities.  Instantiated in FP64 for small lights (u = t - cos(alpha) and
// v = s + sin(alpha) cos(c) cancel to ~sA^2) and in FP32 for comfortably large ones.
template <class T>
__device__ __forceinline__ bool arvo_direct(V3<T> v0, V3<T> v1, V3<T> v2, V3<T> x1, V3<T> n, T sA, T xi1, T xi2, V3<T>* dir)
{
    V3<T> A = v0 - x1, B = v1 - x1, C = v2 - x1;
    A = A * M<T>::rsqrt_(dot(A, A)); B = B * M<T>::rsqrt_(dot(B, B)); C = C * M<T>::rsqrt_(dot(C, C));
    // winding (:206-211): the sign of ((C-A)^ x (B-A)^).n is that of ((C-A) x (B-A)).n
    if (dot(cross(C - A, B - A), n) < T(0)) { V3<T> t = B; B = C; C = t; }
    const V3<T> nab = cross(A, B), nac = cross(A, C);
    const T lab = dot(nab, nab), lac = dot(nac, nac);
    if (!(lab > T(0)) || !(lac > T(0)) || !(sA > T(0))) return false;
    const T ca = clamp1(dot(nab, nac) * M<T>::rsqrt_(lab * lac));     // cos(alpha) = (AxB)^.(AxC)^   (:225)
    const T sa = M<T>::sqrt_(M<T>::max_(T(0), T(1) - ca * ca));
    const T cc = clamp1(dot(A, B));                                   // cos(c)                       (:217)
    T sd, cd;
    M<T>::sincos_(xi1 * sA, &sd, &cd);
    const T s = sd * ca - cd * sa, t = cd * ca + sd * sa;             // sin / cos(xi1 sA - alpha)    (:290)
    const T u = t - ca, v = s + sa * cc;                              // (:291)
    const T q = clamp1(((v * t - u * s) * ca - v) / ((v * s + u * t) * sa));  // (:292)
    V3<T> Ct = C - A * dot(C, A);
    Ct = Ct * M<T>::rsqrt_(dot(Ct, Ct));
    V3<T> C1 = A * q + Ct * M<T>::sqrt_(M<T>::max_(T(0), T(1) - q * q));      // (:293)
    C1 = C1 * M<T>::rsqrt_(dot(C1, C1));
    const T z = T(1) - xi2 * (T(1) - dot(C1, B));                     // (:295)
    V3<T> Bt = C1 - B * dot(C1, B);
    const T lbt = dot(Bt, Bt);
    V3<T> P = B * z;
    if (lbt > T(0)) P = P + Bt * (M<T>::rsqrt_(lbt) * M<T>::sqrt_(M<T>::max_(T(0), T(1) - z * z)));  // (:296)
    P = P * M<T>::rsqrt_(dot(P, P));
    *dir = P;
    return !M<T>::bad_(P.x + P.y + P.z);
}

// ---- light hierarchy (MCPT_LIGHT_TREE, light_tree.h) ---------------------------------------------------------
// O(log L) stand-in for the O(L) selection loop of Mylight.cpp:169-276: descend a binary tree over the light
// triangles, choosing a child with probability proportional to a conservative estimate of what it can send to
// the vertex.  The estimate is zero only where the culls of Mylight.cpp:180-197 remove every light below the node
// (the vertex is behind all of them, or all of them are below its horizon), so every light with a non-zero
// spherical-triangle weight keeps a non-zero probability and the estimator stays unbiased.
//
// cos(max(0, a - b)) and sin(max(0, a - b)) for angles in [0, pi] given by their sines and cosines
__device__ __forceinline__ float cos_sub_clamped(float sa, float ca, float sb, float cb) { return ca > cb ? 1.f : ca * cb + sa * sb; }
__device__ __forceinline__ float sin_sub_clamped(float sa, float ca, float sb, float cb) { return ca > cb ? 0.f : sa * cb - ca * sb; }

// t0 = centre | radius, t1 = cone axis | cos(theta_o)
__device__ __forceinline__ float lt_importance(float4 t0, float4 t1, float power, f3 x, f3 n)
{
    const f3 dv = mk<float>(x.x - t0.x, x.y - t0.y, x.z - t0.z);  // centre -> vertex
    const float d2 = dot(dv, dv), r2 = t0.w * t0.w;
    const f3 wi = dv * rsqrtf(fmaxf(d2, 1e-30f));
    // half-angle under which the vertex sees the node's bounding sphere; pi when it is inside
    const bool inside = !(d2 > r2);
    const float sin2_b = inside ? 0.f : r2 / d2;
    const float sin_b = sqrtf(sin2_b), cos_b = inside ? -1.f : sqrtf(fmaxf(0.f, 1.f - sin2_b));
    // emitter side: angle between the cone axis and the direction to the vertex, less the cone, less the sphere
    const float cos_w = dot(mk<float>(t1.x, t1.y, t1.z), wi), sin_w = sqrtf(fmaxf(0.f, 1.f - cos_w * cos_w));
    const float cos_o = t1.w, sin_o = sqrtf(fmaxf(0.f, 1.f - cos_o * cos_o));
    const float cos_x = cos_sub_clamped(sin_w, cos_w, sin_o, cos_o), sin_x = sin_sub_clamped(sin_w, cos_w, sin_o, cos_o);
    const float cos_e = cos_sub_clamped(sin_x, cos_x, sin_b, cos_b);
    // receiver side: angle between the vertex normal and the direction to the centre, less the sphere
    const float cos_i = -dot(n, wi), sin_i = sqrtf(fmaxf(0.f, 1.f - cos_i * cos_i));
    const float cos_r = cos_sub_clamped(sin_i, cos_i, sin_b, cos_b);
    if (!(cos_e > 0.f) || !(cos_r > 0.f)) return 0.f;  // every light below faces away (:180-185) or is below the horizon (:187-197)
    return power * cos_e * cos_r / fmaxf(d2, r2);
}

// probability of descending into the LEFT child of inner node `node`; writes the two children
__device__ __forceinline__ float lt_left_prob(const DScene& s, int left, int right, f3 x, f3 n)
{
    const float4* L = s.ltree + 3 * left;
    const float4* R = s.ltree + 3 * right;
    const float pl = __ldg(L + 2).x, pr = __ldg(R + 2).x;
    float il = lt_importance(__ldg(L), __ldg(L + 1), pl, x, n), ir = lt_importance(__ldg(R), __ldg(R + 1), pr, x, n);
    if (!(il + ir > 0.f) || !(il + ir < CUDART_INF_F)) { il = pl; ir = pr; }  // nothing below can contribute: any choice will do
    if (!(il + ir > 0.f)) return 0.5f;
    return fminf(il / (il + ir), 1.f);  // the approximate division may land one ulp above 1
}

// draws a light with one uniform (rescaled on the way down); *prob = its selection probability
__device__ __forceinline__ int lt_pick(const DScene& s, f3 x, f3 n, float u, float* prob)
{
    int node = 0;
    float p = 1.f;
    for (;;) {
        const float4 t2 = __ldg(s.ltree + 3 * node + 2);
        const int left = __float_as_int(t2.y), right = __float_as_int(t2.z);
        if (left < 0) { *prob = p; return ~left; }
        const float pl = lt_left_prob(s, left, right, x, n);
        if (u < pl) { node = left; p *= pl; u = u / pl; }
        else { node = right; p *= 1.f - pl; u = (u - pl) / (1.f - pl); }
        u = fminf(fmaxf(u, 0.f), 0.99999994f);
    }
}

// selection probability of a given light at this vertex (the same arithmetic as lt_pick, bottom-up)
__device__ __forceinline__ float lt_prob(const DScene& s, int light, f3 x, f3 n)
{
    int node = __ldg(s.ltree_leaf + light);
    float p = 1.f;
    for (;;) {
        const int parent = __float_as_int(__ldg(s.ltree + 3 * node + 2).w);
        if (parent < 0) return p;
        const float4 t2 = __ldg(s.ltree + 3 * parent + 2);
        const int left = __float_as_int(t2.y), right = __float_as_int(t2.z);
        const float pl = lt_left_prob(s, left, right, x, n);
        p *= (node == left) ? pl : 1.f - pl;
        node = parent;
    }
}

// ---- uniform-area sampling, Mylight.cpp:102-160 --------------------------------------------------------
template <class T> struct LightPoint { int prim; int light; V3<T> coord, Le; T prob; };

// barycentric point of :141-154 on light l
template <class T>
__device__ __forceinline__ V3<T> uniform_point(const LightTri<T>& lt, T xi1, T xi2)
{
    T beta = T(1) - M<T>::sqrt_(T(1) - xi1);
    T gamma = (T(1) - beta) * xi2;
    T alpha = T(1) - beta - gamma;
    return lt.v0 * alpha + lt.v1 * beta + lt.v2 * gamma;
}

// FP32 renderer: inverse-CDF lookups in the prebuilt tables (pdf in AREA measure)
__device__ __forceinline__ LightPoint<float> light_sample_uniform_fast(const DScene& s, float u_class, float u_tri, float xi1, float xi2)
{
    int c = 0;
    float pc = 1.f;
    if (s.C > 1) {
        while (c < s.C - 1 && __ldg(s.class_cdf + c) < u_class) c++;
        pc = __ldg(s.class_cdf + c) - (c ? __ldg(s.class_cdf + c - 1) : 0.f);
    }
    int b = __ldg(s.class_begin + c), e = __ldg(s.class_begin + c + 1);
    int l = b;
    float pt = 1.f;
    if (e - b > 1) {
        int lo = b, hi = e - 1;  // first l with tri_cdf[l] >= u
        while (lo < hi) {
            int mid = (lo + hi) >> 1;
            if (__ldg(s.tri_cdf + mid) < u_tri) lo = mid + 1; else hi = mid;
        }
        l = lo;
        pt = __ldg(s.tri_cdf + l) - (l > b ? __ldg(s.tri_cdf + l - 1) : 0.f);
    }
    LightTri<float> lt = load_light<float>(s, l);
    LightPoint<float> r;
    r.prim = lt.prim; r.light = l; r.coord = uniform_point<float>(lt, xi1, xi2); r.Le = lt.Le;
    r.prob = pc * pt / lt.area;
    return r;
}

}  // namespace mcpt
