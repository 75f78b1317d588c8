// Shared device/host helpers for the sm_100a tile kernels of the Cholesky path.
// Element types follow the reference's MatrixElementTypes (test/include/dlaf_test/util_types.h:20-23):
// float, double, complex<float>, complex<double>; on the device the complex types are float2/double2.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

namespace dlafb {

// Number of kernels this library has launched (reported by bench.py as gpu_launches).
extern std::atomic<unsigned long long> g_launches;

template <class T>
struct Tr;

template <>
struct Tr<float> {
  using Real = float;
  static constexpr bool is_complex = false;
  static constexpr char code = 's';
  __host__ __device__ static float zero() { return 0.f; }
  __host__ __device__ static float from_real(float r) { return r; }
  __host__ __device__ static float real(float a) { return a; }
  __host__ __device__ static float conj(float a) { return a; }
  __host__ __device__ static float mul(float a, float b) { return a * b; }
  __host__ __device__ static float scale(float a, float r) { return a * r; }
  __host__ __device__ static float add(float a, float b) { return a + b; }
  // c + a * b
  __host__ __device__ static float fma(float a, float b, float c) { return fmaf(a, b, c); }
  // c - a * conj(b)
  __host__ __device__ static float fnma_conj(float a, float b, float c) { return fmaf(-a, b, c); }
  __host__ __device__ static float fnma(float a, float b, float c) { return fmaf(-a, b, c); }
  __host__ __device__ static float rsqrt_(float a) { return rsqrtf(a); }
  __host__ __device__ static float abs(float a) { return fabsf(a); }
};

template <>
struct Tr<double> {
  using Real = double;
  static constexpr bool is_complex = false;
  static constexpr char code = 'd';
  __host__ __device__ static double zero() { return 0.; }
  __host__ __device__ static double from_real(double r) { return r; }
  __host__ __device__ static double real(double a) { return a; }
  __host__ __device__ static double conj(double a) { return a; }
  __host__ __device__ static double mul(double a, double b) { return a * b; }
  __host__ __device__ static double scale(double a, double r) { return a * r; }
  __host__ __device__ static double add(double a, double b) { return a + b; }
  __host__ __device__ static double fma(double a, double b, double c) { return ::fma(a, b, c); }
  __host__ __device__ static double fnma_conj(double a, double b, double c) { return ::fma(-a, b, c); }
  __host__ __device__ static double fnma(double a, double b, double c) { return ::fma(-a, b, c); }
  __host__ __device__ static double rsqrt_(double a) { return ::rsqrt(a); }
  __host__ __device__ static double abs(double a) { return fabs(a); }
};

template <class C, class R, char CODE>
struct TrComplex {
  using Real = R;
  static constexpr bool is_complex = true;
  static constexpr char code = CODE;
  __host__ __device__ static C make(R x, R y) {
    C c;
    c.x = x;
    c.y = y;
    return c;
  }
  __host__ __device__ static C zero() { return make(R(0), R(0)); }
  __host__ __device__ static C from_real(R r) { return make(r, R(0)); }
  __host__ __device__ static R real(C a) { return a.x; }
  __host__ __device__ static C conj(C a) { return make(a.x, -a.y); }
  __host__ __device__ static C mul(C a, C b) { return make(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
  __host__ __device__ static C scale(C a, R r) { return make(a.x * r, a.y * r); }
  __host__ __device__ static C add(C a, C b) { return make(a.x + b.x, a.y + b.y); }
  __host__ __device__ static C fma(C a, C b, C c) {
    return make(c.x + a.x * b.x - a.y * b.y, c.y + a.x * b.y + a.y * b.x);
  }
  // c - a * b
  __host__ __device__ static C fnma(C a, C b, C c) {
    return make(c.x - a.x * b.x + a.y * b.y, c.y - a.x * b.y - a.y * b.x);
  }
  __host__ __device__ static R rsqrt_(R a) { return Tr<R>::rsqrt_(a); }
  // c - a * conj(b)
  __host__ __device__ static C fnma_conj(C a, C b, C c) {
    return make(c.x - a.x * b.x - a.y * b.y, c.y + a.x * b.y - a.y * b.x);
  }
  __host__ __device__ static R abs(C a) { return hypot(a.x, a.y); }
};

template <>
struct Tr<float2> : TrComplex<float2, float, 'c'> {};
template <>
struct Tr<double2> : TrComplex<double2, double, 'z'> {};

__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// Warp shuffle for any of the four element types.
template <class T>
__device__ inline T shfl(T v, int src) {
  return __shfl_sync(0xffffffffu, v, src);
}
template <>
__device__ inline float2 shfl<float2>(float2 v, int src) {
  float2 r;
  r.x = __shfl_sync(0xffffffffu, v.x, src);
  r.y = __shfl_sync(0xffffffffu, v.y, src);
  return r;
}
template <>
__device__ inline double2 shfl<double2>(double2 v, int src) {
  double2 r;
  r.x = __shfl_sync(0xffffffffu, v.x, src);
  r.y = __shfl_sync(0xffffffffu, v.y, src);
  return r;
}

}  // namespace dlafb
