// krylov.cu -- C ABI of the device-resident Krylov vector kernels (include/hmb200.h, "Krylov" section).
#include "hmb200.h"
#include "internal.h"
#include "../kernels/krylov_kernels.cuh"

#include <algorithm>
#include <stdexcept>
#include <string>

using namespace hmb200;

namespace {

#define KCHECK(expr)                                                                                                   \
    do {                                                                                                               \
        cudaError_t _e = (expr);                                                                                       \
        if (_e != cudaSuccess) throw std::runtime_error(std::string(#expr) + " failed: " + cudaGetErrorString(_e));   \
    } while (0)

template <typename F>
int kguard(F &&f) {
    try {
        return f();
    } catch (const std::exception &e) {
        hmb200i_set_error(e.what());
        return 1;
    }
}

inline int n_chunks(int64_t n) { return (int)((n + kKrylovChunk - 1) / kKrylovChunk); }

void check_common(int64_t n, int elem_bytes) {
    if (elem_bytes != 8 && elem_bytes != 16) throw std::runtime_error("elem_bytes must be 8 or 16");
    if (n < 0) throw std::runtime_error("negative size");
}

template <typename T>
void reduce_partials(const T *partial, int chunks, int k, T *out, cudaStream_t s) {
    krylov_reduce_kernel<T><<<(k + 7) / 8, 256, 0, s>>>(partial, chunks, k, out); // one warp per j
}

template <typename T>
int dots(const void *V, int64_t ldv, int k, const void *w, int64_t n, void *h, void *scratch, cudaStream_t s) {
    if (k <= 0) return 0;
    if (n == 0) { KCHECK(cudaMemsetAsync(h, 0, sizeof(T) * (size_t)k, s)); return 0; }
    const int chunks = n_chunks(n);
    krylov_dots_kernel<T><<<chunks, kKrylovThreads, 0, s>>>((const T *)V, ldv, k, (const T *)w, n, (T *)scratch);
    reduce_partials<T>((const T *)scratch, chunks, k, (T *)h, s);
    KCHECK(cudaGetLastError());
    return 0;
}

template <typename T>
int combine(const void *V, int64_t ldv, int k, const void *h, double sign, void *w, int64_t n, double *norm2, void *scratch, cudaStream_t s) {
    if (k > 2048) throw std::runtime_error("at most 2048 basis vectors per combine");
    if (n == 0) { if (norm2) KCHECK(cudaMemsetAsync(norm2, 0, sizeof(double), s)); return 0; }
    const int chunks = n_chunks(n);
    krylov_combine_kernel<T><<<chunks, kKrylovThreads, sizeof(T) * (size_t)std::max(k, 1), s>>>((const T *)V, ldv, std::max(k, 0), (const T *)h, sign, (T *)w, n, norm2 ? (double *)scratch : nullptr);
    if (norm2) reduce_partials<double>((const double *)scratch, chunks, 1, norm2, s);
    KCHECK(cudaGetLastError());
    return 0;
}

template <typename T>
int axpby(T alpha, const void *x, T beta, const void *y, void *out, int64_t n, double *norm2, void *scratch, cudaStream_t s) {
    if (n == 0) { if (norm2) KCHECK(cudaMemsetAsync(norm2, 0, sizeof(double), s)); return 0; }
    const int chunks = n_chunks(n);
    krylov_axpby_kernel<T><<<chunks, kKrylovThreads, 0, s>>>(alpha, (const T *)x, beta, (const T *)y, (T *)out, n, norm2 ? (double *)scratch : nullptr);
    if (norm2) reduce_partials<double>((const double *)scratch, chunks, 1, norm2, s);
    KCHECK(cudaGetLastError());
    return 0;
}

} // namespace

extern "C" {

int64_t hmb200_krylov_scratch_elems(int64_t n, int k) { return (int64_t)std::max(n_chunks(std::max<int64_t>(n, 0)), 1) * std::max(k, 1); }

int hmb200_krylov_dots(const void *V, int64_t ldv, int k, const void *w, int64_t n, void *h, void *scratch, int elem_bytes, void *stream) {
    return kguard([&] {
        check_common(n, elem_bytes);
        return elem_bytes == 8 ? dots<double>(V, ldv, k, w, n, h, scratch, (cudaStream_t)stream) : dots<double2>(V, ldv, k, w, n, h, scratch, (cudaStream_t)stream);
    });
}

int hmb200_krylov_combine(const void *V, int64_t ldv, int k, const void *h, double sign, void *w, int64_t n, double *norm2, void *scratch, int elem_bytes, void *stream) {
    return kguard([&] {
        check_common(n, elem_bytes);
        return elem_bytes == 8 ? combine<double>(V, ldv, k, h, sign, w, n, norm2, scratch, (cudaStream_t)stream)
                               : combine<double2>(V, ldv, k, h, sign, w, n, norm2, scratch, (cudaStream_t)stream);
    });
}

int hmb200_krylov_axpby(const double *alpha, const void *x, const double *beta, const void *y, void *out, int64_t n, double *norm2, void *scratch, int elem_bytes, void *stream) {
    return kguard([&] {
        check_common(n, elem_bytes);
        if (elem_bytes == 8) return axpby<double>(alpha[0], x, beta[0], y, out, n, norm2, scratch, (cudaStream_t)stream);
        return axpby<double2>(make_double2(alpha[0], alpha[1]), x, make_double2(beta[0], beta[1]), y, out, n, norm2, scratch, (cudaStream_t)stream);
    });
}

int hmb200_krylov_scale(void *w, int64_t n, const double *norm2, int elem_bytes, void *stream) {
    return kguard([&] {
        check_common(n, elem_bytes);
        if (n == 0) return 0;
        const unsigned g = (unsigned)std::min<int64_t>((n + 255) / 256, 148 * 8);
        if (elem_bytes == 8) krylov_scale_kernel<double><<<g, 256, 0, (cudaStream_t)stream>>>((double *)w, n, norm2);
        else krylov_scale_kernel<double2><<<g, 256, 0, (cudaStream_t)stream>>>((double2 *)w, n, norm2);
        KCHECK(cudaGetLastError());
        return 0;
    });
}
}
