// krylov_kernels.cuh -- the vector kernels of the device-resident GMRES (sm_100a).
//
// They replace the BLAS-1/2 calls of HPDDM's Arnoldi process on the Krylov basis
// (/root/reference/lib/hpddm/include/HPDDM_iterative.hpp:512-560 orthogonalization: gemv 'C' + gemv 'N';
//  :722-747 Arnoldi: dot, scaling; HPDDM_GMRES.hpp:62-64 residual axpby; HPDDM_iterative.hpp:316-326 addSol gemv + axpy)
// for basis vectors that never leave HBM.  All of them are bandwidth-bound streams over the local rows of the basis:
//   dots    : h[j] = <V_j, w>            reads (k+1) n elements        one warp per (vector, 2048-row chunk)
//   combine : w += sign * sum_j h[j] V_j reads (k+1) n, writes n       + ||w||^2 partials
//   axpby   : out = alpha x + beta y     reads 2n, writes n            + ||out||^2 partials
//   scale   : w /= sqrt(*norm2)
// Reductions are two-stage (per-chunk partials, then one fixed-order pass): no atomics, bit-reproducible run to run.
#pragma once
#include <cstdint>

#include "scalar.cuh"

namespace hmb200 {

constexpr int kKrylovChunk   = 2048; // rows per CTA
constexpr int kKrylovThreads = 256;

template <typename T>
struct KOps;
template <>
struct KOps<double> {
    static __device__ __forceinline__ void conj_fma(double &acc, double v, double w) { acc = ::fma(v, w, acc); }
    static __device__ __forceinline__ double norm2(double a) { return a * a; }
    static __device__ __forceinline__ double scale(double a, double s) { return a * s; }
    static __device__ __forceinline__ double div(double a, double s) { return a / s; }
};
template <>
struct KOps<double2> {
    // acc += conj(v) * w
    static __device__ __forceinline__ void conj_fma(double2 &acc, double2 v, double2 w) {
        acc.x = ::fma(v.x, w.x, acc.x);
        acc.x = ::fma(v.y, w.y, acc.x);
        acc.y = ::fma(v.x, w.y, acc.y);
        acc.y = ::fma(-v.y, w.x, acc.y);
    }
    static __device__ __forceinline__ double norm2(double2 a) { return ::fma(a.x, a.x, a.y * a.y); }
    static __device__ __forceinline__ double2 scale(double2 a, double s) { return make_double2(a.x * s, a.y * s); }
    static __device__ __forceinline__ double2 div(double2 a, double s) { return make_double2(a.x / s, a.y / s); }
};

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) v = Scalar<T>::add(v, Scalar<T>::shfl_xor(v, m));
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
    return v;
}

// sum of the squared moduli held by the threads of a CTA -> partial[blockIdx.x] (fixed order: lanes, then warps)
__device__ __forceinline__ void block_norm2_store(double local, double *partial) {
    __shared__ double red[kKrylovThreads / 32];
    local = warp_sum_d(local);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = local;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
#pragma unroll
        for (int i = 0; i < kKrylovThreads / 32; i++) s += red[i];
        partial[blockIdx.x] = s;
    }
}

// partial[chunk * k + j] = sum over the chunk's rows of conj(V[j][row]) * w[row]
template <typename T>
__global__ void __launch_bounds__(kKrylovThreads) krylov_dots_kernel(const T *__restrict__ V, int64_t ldv, int k, const T *__restrict__ w, int64_t n, T *__restrict__ partial) {
    __shared__ T ws[kKrylovChunk];
    const int64_t c0 = (int64_t)blockIdx.x * kKrylovChunk;
    const int len    = (int)min((int64_t)kKrylovChunk, n - c0);
    for (int e = threadIdx.x; e < len; e += kKrylovThreads) ws[e] = w[c0 + e];
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int j = warp; j < k; j += kKrylovThreads / 32) {
        const T *v = V + (int64_t)j * ldv + c0;
        T a0 = Scalar<T>::zero(), a1 = a0, a2 = a0, a3 = a0;
        int e = lane;
        for (; e + 96 < len; e += 128) { // 4 independent loads in flight per lane
            const T v0 = v[e], v1 = v[e + 32], v2 = v[e + 64], v3 = v[e + 96];
            KOps<T>::conj_fma(a0, v0, ws[e]);
            KOps<T>::conj_fma(a1, v1, ws[e + 32]);
            KOps<T>::conj_fma(a2, v2, ws[e + 64]);
            KOps<T>::conj_fma(a3, v3, ws[e + 96]);
        }
        for (; e < len; e += 32) KOps<T>::conj_fma(a0, v[e], ws[e]);
        T acc = Scalar<T>::add(Scalar<T>::add(a0, a1), Scalar<T>::add(a2, a3));
        acc   = warp_sum<T>(acc);
        if (lane == 0) partial[(int64_t)blockIdx.x * k + j] = acc;
    }
}

// out[j] = sum_c partial[c * k + j]: one warp per j, lane l sums the chunks l, l+32, ... in ascending order, then a fixed
// xor-shuffle tree (deterministic; the partials are a few hundred entries per j, so this is pure latency: 1 dependent load
// chain of n_chunks/32 instead of n_chunks)
template <typename T>
__global__ void __launch_bounds__(256) krylov_reduce_kernel(const T *__restrict__ partial, int n_chunks, int k, T *__restrict__ out) {
    const int j = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (j >= k) return;
    T a0 = Scalar<T>::zero(), a1 = a0;
    int c = lane;
    for (; c + 32 < n_chunks; c += 64) {
        const T p0 = partial[(int64_t)c * k + j], p1 = partial[(int64_t)(c + 32) * k + j];
        a0 = Scalar<T>::add(a0, p0);
        a1 = Scalar<T>::add(a1, p1);
    }
    if (c < n_chunks) a0 = Scalar<T>::add(a0, partial[(int64_t)c * k + j]);
    const T acc = warp_sum<T>(Scalar<T>::add(a0, a1));
    if (lane == 0) out[j] = acc;
}

// w[row] += sign * sum_j h[j] * V[j][row]; optional ||w||^2 partial per CTA
template <typename T>
__global__ void __launch_bounds__(kKrylovThreads) krylov_combine_kernel(const T *__restrict__ V, int64_t ldv, int k, const T *__restrict__ h, double sign, T *__restrict__ w, int64_t n, double *__restrict__ partial) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T *hs = reinterpret_cast<T *>(smem_raw);
    for (int j = threadIdx.x; j < k; j += kKrylovThreads) hs[j] = KOps<T>::scale(h[j], sign);
    __syncthreads();
    constexpr int Q   = kKrylovChunk / kKrylovThreads; // rows per thread
    const int64_t c0 = (int64_t)blockIdx.x * kKrylovChunk + threadIdx.x;
    T acc[Q];
#pragma unroll
    for (int q = 0; q < Q; q++) {
        const int64_t e = c0 + (int64_t)q * kKrylovThreads;
        acc[q] = e < n ? w[e] : Scalar<T>::zero();
    }
    for (int j = 0; j < k; j++) {
        const T hj   = hs[j];
        const T *v   = V + (int64_t)j * ldv;
        T vv[Q];
#pragma unroll
        for (int q = 0; q < Q; q++) {
            const int64_t e = c0 + (int64_t)q * kKrylovThreads;
            vv[q] = e < n ? v[e] : Scalar<T>::zero();
        }
#pragma unroll
        for (int q = 0; q < Q; q++) Scalar<T>::fma(acc[q], hj, vv[q]);
    }
    double local = 0.0;
#pragma unroll
    for (int q = 0; q < Q; q++) {
        const int64_t e = c0 + (int64_t)q * kKrylovThreads;
        if (e < n) {
            w[e] = acc[q];
            local += KOps<T>::norm2(acc[q]);
        }
    }
    if (partial) block_norm2_store(local, partial);
}

// out = alpha * x + beta * y; optional ||out||^2 partial per CTA
template <typename T>
__global__ void __launch_bounds__(kKrylovThreads) krylov_axpby_kernel(T alpha, const T *__restrict__ x, T beta, const T *__restrict__ y, T *__restrict__ out, int64_t n, double *__restrict__ partial) {
    constexpr int Q   = kKrylovChunk / kKrylovThreads;
    const int64_t c0 = (int64_t)blockIdx.x * kKrylovChunk + threadIdx.x;
    double local     = 0.0;
#pragma unroll
    for (int q = 0; q < Q; q++) {
        const int64_t e = c0 + (int64_t)q * kKrylovThreads;
        if (e < n) {
            T r = Scalar<T>::mul(alpha, x[e]);
            Scalar<T>::fma(r, beta, y[e]);
            out[e] = r;
            local += KOps<T>::norm2(r);
        }
    }
    if (partial) block_norm2_store(local, partial);
}

// w /= sqrt(*norm2)   (HPDDM divides by H(i+1,i) = sqrt(sn), HPDDM_iterative.hpp:743-745)
template <typename T>
__global__ void krylov_scale_kernel(T *__restrict__ w, int64_t n, const double *__restrict__ norm2) {
    const double d = sqrt(*norm2);
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (int64_t)gridDim.x * blockDim.x) w[e] = KOps<T>::div(w[e], d);
}

} // namespace hmb200
