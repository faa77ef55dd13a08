// scalar.cuh -- element types of the engine: double and complex<double> (as double2 = re, im).
#pragma once
#include <cuda_runtime.h>

namespace hmb200 {

template <typename T>
struct Scalar;

template <>
struct Scalar<double> {
    using type = double;
    static __device__ __forceinline__ double zero() { return 0.0; }
    static __device__ __forceinline__ void fma(double &acc, double a, double b) { acc = ::fma(a, b, acc); }
    static __device__ __forceinline__ double add(double a, double b) { return a + b; }
    static __device__ __forceinline__ double mul(double a, double b) { return a * b; }
    static __device__ __forceinline__ bool is_zero(double a) { return a == 0.0; }
    static __device__ __forceinline__ double make(double re, double) { return re; }
    static __device__ __forceinline__ double shfl(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }
    static __device__ __forceinline__ double shfl_xor(double v, int m) { return __shfl_xor_sync(0xffffffffu, v, m); }
};

template <>
struct Scalar<double2> {
    using type = double2;
    static __device__ __forceinline__ double2 zero() { return make_double2(0.0, 0.0); }
    static __device__ __forceinline__ void fma(double2 &acc, double2 a, double2 b) {
        acc.x = ::fma(a.x, b.x, acc.x);
        acc.x = ::fma(-a.y, b.y, acc.x);
        acc.y = ::fma(a.x, b.y, acc.y);
        acc.y = ::fma(a.y, b.x, acc.y);
    }
    static __device__ __forceinline__ double2 add(double2 a, double2 b) { return make_double2(a.x + b.x, a.y + b.y); }
    static __device__ __forceinline__ double2 mul(double2 a, double2 b) { return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
    static __device__ __forceinline__ bool is_zero(double2 a) { return a.x == 0.0 && a.y == 0.0; }
    static __device__ __forceinline__ double2 make(double re, double im) { return make_double2(re, im); }
    static __device__ __forceinline__ double2 shfl(double2 v, int src) {
        return make_double2(__shfl_sync(0xffffffffu, v.x, src), __shfl_sync(0xffffffffu, v.y, src));
    }
    static __device__ __forceinline__ double2 shfl_xor(double2 v, int m) {
        return make_double2(__shfl_xor_sync(0xffffffffu, v.x, m), __shfl_xor_sync(0xffffffffu, v.y, m));
    }
};

} // namespace hmb200
