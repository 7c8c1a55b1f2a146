// csrc/device_math.cuh -- 3-vector helpers templated on the real type (vec / matrix3d of the
// reference: vec.h:3-28, vec.cpp:1-103, matrix3d.cpp:8-40), plus the scalar math shims that let one
// body of shading code instantiate in FP64 (parity entry points) and FP32 (the renderer).
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>

namespace mcpt {

template <class T> struct V3 { T x, y, z; };
typedef V3<float> f3;
typedef V3<double> d3;

template <class T> __host__ __device__ __forceinline__ V3<T> mk(T x, T y, T z) { V3<T> r; r.x = x; r.y = y; r.z = z; return r; }
template <class T> __host__ __device__ __forceinline__ V3<T> operator+(V3<T> a, V3<T> b) { return mk<T>(a.x + b.x, a.y + b.y, a.z + b.z); }
template <class T> __host__ __device__ __forceinline__ V3<T> operator-(V3<T> a, V3<T> b) { return mk<T>(a.x - b.x, a.y - b.y, a.z - b.z); }
template <class T> __host__ __device__ __forceinline__ V3<T> operator*(V3<T> a, T c) { return mk<T>(a.x * c, a.y * c, a.z * c); }
template <class T> __host__ __device__ __forceinline__ V3<T> operator*(V3<T> a, V3<T> b) { return mk<T>(a.x * b.x, a.y * b.y, a.z * b.z); }
template <class T> __host__ __device__ __forceinline__ V3<T> operator-(V3<T> a) { return mk<T>(-a.x, -a.y, -a.z); }
template <class T> __host__ __device__ __forceinline__ T dot(V3<T> a, V3<T> b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
template <class T> __host__ __device__ __forceinline__ V3<T> cross(V3<T> a, V3<T> b)
{
    return mk<T>(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
template <class T> __host__ __device__ __forceinline__ T sum3(V3<T> a) { return a.x + a.y + a.z; }

// scalar shims -------------------------------------------------------------------------------------
template <class T> struct M;
template <> struct M<double> {
    static constexpr double pi = 3.14159265358979323846;
    static constexpr double eps = 1e-8;  // vec.h:7, Myobj.h:69, Mylight.h:101, BRDF.cpp:8
    static __device__ __forceinline__ double sqrt_(double x) { return ::sqrt(x); }
    static __device__ __forceinline__ double rsqrt_(double x) { return ::rsqrt(x); }
    static __device__ __forceinline__ double sin_(double x) { return ::sin(x); }
    static __device__ __forceinline__ double cos_(double x) { return ::cos(x); }
    static __device__ __forceinline__ void sincos_(double x, double* s, double* c) { ::sincos(x, s, c); }
    static __device__ __forceinline__ double acos_(double x) { return ::acos(x); }
    static __device__ __forceinline__ double atan2_(double y, double x) { return ::atan2(y, x); }
    static __device__ __forceinline__ double pow_(double x, double y) { return ::pow(x, y); }
    static __device__ __forceinline__ double abs_(double x) { return ::fabs(x); }
    static __device__ __forceinline__ double min_(double a, double b) { return ::fmin(a, b); }
    static __device__ __forceinline__ double max_(double a, double b) { return ::fmax(a, b); }
    static __device__ __forceinline__ bool bad_(double x) { return isinf(x) || isnan(x); }
};
template <> struct M<float> {
    static constexpr float pi = 3.14159265358979323846f;
    static constexpr float eps = 1e-8f;
    static __device__ __forceinline__ float sqrt_(float x) { return ::sqrtf(x); }
    static __device__ __forceinline__ float rsqrt_(float x) { return ::rsqrtf(x); }
    static __device__ __forceinline__ float sin_(float x) { return ::sinf(x); }
    static __device__ __forceinline__ float cos_(float x) { return ::cosf(x); }
    static __device__ __forceinline__ void sincos_(float x, float* s, float* c) { ::sincosf(x, s, c); }
    static __device__ __forceinline__ float acos_(float x) { return ::acosf(x); }
    static __device__ __forceinline__ float atan2_(float y, float x) { return ::atan2f(y, x); }
    static __device__ __forceinline__ float pow_(float x, float y) { return ::powf(x, y); }
    static __device__ __forceinline__ float abs_(float x) { return ::fabsf(x); }
    static __device__ __forceinline__ float min_(float a, float b) { return ::fminf(a, b); }
    static __device__ __forceinline__ float max_(float a, float b) { return ::fmaxf(a, b); }
    static __device__ __forceinline__ bool bad_(float x) { return isinf(x) || isnan(x); }
};

template <class T> __device__ __forceinline__ T norm2(V3<T> a) { return M<T>::sqrt_(dot(a, a)); }                 // vec.cpp:55
template <class T> __device__ __forceinline__ V3<T> normalized(V3<T> a) { T l = norm2(a); return mk<T>(a.x / l, a.y / l, a.z / l); }  // vec.cpp:99
// fast normalise for the FP32 renderer
__device__ __forceinline__ f3 normalize_fast(f3 a) { float r = rsqrtf(dot(a, a)); return a * r; }
template <class T> __device__ __forceinline__ T clamp1(T x) { return M<T>::max_(T(-1), M<T>::min_(T(1), x)); }    // fmax(-1,fmin(1,x))

__device__ __forceinline__ f3 to_f3(float4 v) { return mk<float>(v.x, v.y, v.z); }
__device__ __forceinline__ d3 to_d3(float4 v) { return mk<double>((double)v.x, (double)v.y, (double)v.z); }
__device__ __forceinline__ d3 to_d3(f3 v) { return mk<double>((double)v.x, (double)v.y, (double)v.z); }
__device__ __forceinline__ f3 to_f3(d3 v) { return mk<float>((float)v.x, (float)v.y, (float)v.z); }
template <class T> __device__ __forceinline__ V3<T> cvt(float4 v);
template <> __device__ __forceinline__ f3 cvt<float>(float4 v) { return to_f3(v); }
template <> __device__ __forceinline__ d3 cvt<double>(float4 v) { return to_d3(v); }

}  // namespace mcpt
