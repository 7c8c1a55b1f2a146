// csrc/brdf.cuh -- Phong diffuse+specular BRDF: evaluation, importance sampling, sampling pdf.
// Device restatement of BRDF.cpp:17-132, templated on the real type: T = double is what the parity
// entry points run (same expression order as the reference), T = float is what the renderer runs.
#pragma once
#include "device_math.cuh"

namespace mcpt {

template <class T> struct PhongMat { V3<T> kd, ks; T ns; };

// BRDF::get_brdf_phong, BRDF.cpp:17-25.  n normal, wi incident, wr outgoing; no cosine inside.
template <class T>
__device__ __forceinline__ V3<T> phong_eval(V3<T> n, V3<T> wi, V3<T> wr, const PhongMat<T>& m)
{
    V3<T> R = wi * T(-1) + n * (T(2) * dot(wi, n));
    V3<T> ans = m.kd * (T(1) / M<T>::pi);
    T c = dot(wr, R);
    if (c > T(0)) ans = ans + m.ks * ((m.ns + T(1)) * M<T>::pow_(c, m.ns) / (T(2) * M<T>::pi));
    return ans;
}

// The tangent frame of BRDF.cpp:58-71 / :86-99: nx = normalize(axis x e_x) unless |axis.e_x - 1| <= eps,
// then axis x e_y; ny = normalize(axis x nx); dir = normalize([nx ny axis] * local).  The reference
// divides 0/0 when axis == (-1,0,0) (F11); here a vanishing axis x e_x also falls back to e_y, which is
// the only deviation (a measure-zero set of axes).
template <class T>
__device__ __forceinline__ V3<T> phong_frame_dir(V3<T> axis, T sin_t, T cos_t, T phi)
{
    V3<T> nx;
    bool use_x = M<T>::abs_(axis.x - T(1)) > M<T>::eps;
    if (use_x) {
        nx = cross(axis, mk<T>(T(1), T(0), T(0)));
        if (dot(nx, nx) < T(1e-24)) use_x = false;
    }
    if (!use_x) nx = cross(axis, mk<T>(T(0), T(1), T(0)));
    nx = normalized(nx);
    V3<T> ny = normalized(cross(axis, nx));
    T sp, cp;
    M<T>::sincos_(phi, &sp, &cp);
    T lx = sin_t * cp, ly = sin_t * sp, lz = cos_t;
    V3<T> d = mk<T>(nx.x * lx + ny.x * ly + axis.x * lz, nx.y * lx + ny.y * ly + axis.y * lz,
                    nx.z * lx + ny.z * ly + axis.z * lz);
    return normalized(d);
}

// BRDF::sample_from_phong, BRDF.cpp:28-104.  Uniforms in the reference's order: u_lobe (libstdc++
// discrete_distribution over {mean kd, mean ks}: first index whose cumulative probability >= u),
// xi1, xi2.  Returns the SINGLE-LOBE pdf P(lobe) * p_lobe (F8).  The direction may lie below the
// surface; callers reject dir.n < 0 (main.cpp:332,386,472).
template <class T>
__device__ __forceinline__ V3<T> phong_sample(V3<T> n, V3<T> wr, const PhongMat<T>& m, T u_lobe, T xi1, T xi2, T* pdf_out)
{
    T d = sum3(m.kd) / T(3);
    T s = sum3(m.ks) / T(3);
    T pd = d / (d + s);
    // cumulative {pd, 1.0}; index 0 iff pd >= u
    bool diffuse = pd >= u_lobe;
    T phi = T(2) * M<T>::pi * xi2;
    if (diffuse) {
        T theta = T(0.5) * M<T>::acos_(clamp1(T(1) - T(2) * xi1));
        T st, ct;
        M<T>::sincos_(theta, &st, &ct);
        *pdf_out = pd * (ct / M<T>::pi);
        return phong_frame_dir(n, st, ct, phi);
    }
    T ps = s / (d + s);
    T ct = clamp1(M<T>::pow_(xi1, T(1) / (m.ns + T(1))));
    T theta = M<T>::acos_(ct);
    T st, ct2;
    M<T>::sincos_(theta, &st, &ct2);
    *pdf_out = ps * ((m.ns + T(1)) / (T(2) * M<T>::pi) * M<T>::pow_(xi1, m.ns / (m.ns + T(1))));
    V3<T> R = normalized(wr * T(-1) + n * (T(2) * dot(wr, n)));
    return phong_frame_dir(R, st, ct2, phi);
}

// FP32 renderer variant of the specular lobe: cos/sin of the polar angle without the acos round trip
// and with 1 - cos^2 evaluated as -expm1(2 ln(xi1)/(ns+1)), which keeps sin(theta) accurate when
// ns is large (cos(theta) -> 1).  Same distribution and pdf as phong_sample<float>.
__device__ __forceinline__ f3 phong_sample_fast(f3 n, f3 wr, const PhongMat<float>& m, float u_lobe, float xi1, float xi2,
                                                float* pdf_out)
{
    float d = sum3(m.kd) * (1.0f / 3.0f);
    float s = sum3(m.ks) * (1.0f / 3.0f);
    float pd = d / (d + s);
    float phi = 2.0f * M<float>::pi * xi2;
    if (pd >= u_lobe) {
        // theta = acos(1-2 xi1)/2  ->  cos^2 = 1 - xi1, sin^2 = xi1
        float ct = sqrtf(fmaxf(0.f, 1.0f - xi1)), st = sqrtf(xi1);
        *pdf_out = pd * ct * (1.0f / M<float>::pi);
        return phong_frame_dir<float>(n, st, ct, phi);
    }
    float ps = 1.0f - pd;
    float l = __logf(fmaxf(xi1, 1e-30f)) / (m.ns + 1.0f);
    float ct = __expf(l);
    float st = sqrtf(fmaxf(0.f, -expm1f(2.0f * l)));
    *pdf_out = ps * (m.ns + 1.0f) * (0.5f / M<float>::pi) * __expf(l * m.ns);
    f3 R = normalize_fast(wr * -1.0f + n * (2.0f * dot(wr, n)));
    return phong_frame_dir<float>(R, st, ct, phi);
}

// BRDF::eval_sample_from_phong_pdf, BRDF.cpp:107-132: the MIXTURE pdf of sample_from_phong.
template <class T>
__device__ __forceinline__ T phong_pdf(V3<T> n, V3<T> wi, V3<T> wr, const PhongMat<T>& m)
{
    T d = sum3(m.kd) / T(3);
    T s = sum3(m.ks) / T(3);
    T sum = d + s;
    T prob_d = d / sum, prob_s = s / sum;
    T ct = dot(wi, n);
    prob_d = (ct < T(0)) ? T(0) : prob_d * (ct / M<T>::pi);
    V3<T> R = normalized(wr * T(-1) + n * (T(2) * dot(wr, n)));
    T cs = dot(wi, R);
    prob_s = (cs < T(0)) ? T(0) : prob_s * ((m.ns + T(1)) / (T(2) * M<T>::pi) * M<T>::pow_(cs, m.ns));
    return prob_d + prob_s;
}

}  // namespace mcpt
