// krylov_static.cu -- hmb200_krylov_dots_static: the Gram-Schmidt dot products without a caller-supplied workspace.
//
// Who this is for.  hmb200_krylov_dots needs a device workspace for its per-chunk partial sums and leaves it to the caller
// to provide one.  That is a problem for code that is not in a position to allocate device memory when it wants the
// product: an operator callback that a foreign solver invokes in the middle of its iteration (the reference's solver calls
// its operator with raw pointers and nothing else, core/wrappers/wrapper_hpddm.hpp:102-145), a plugin that has been handed
// a stream it must not synchronise (cudaMalloc and cudaFree synchronise the device), a host that runs with a fixed memory
// plan where every byte of HBM is already spoken for, or a C / Fortran / ctypes binding with no device allocator at hand.
// For those callers the partial sums live in ONE array that is part of this library's device image.
//
// What that costs and what to watch for:
//   * Every GPU on which the library's kernels get loaded holds HMB200_STATIC_SCRATCH_BYTES (4 MiB) for this array until
//     the process exits, used or not.
//   * A request that needs more than that is refused with an error (see hmb200_krylov_dots_static below); nothing is
//     allocated behind the caller's back, here or in any routine this file calls (no cudaMalloc / cudaFree / cudaMallocAsync,
//     no library that allocates).
//   * The first call on each device looks up where the array lives on that device (cudaGetSymbolAddress, which may make the
//     driver load the library's device code).  Do that first call during set-up, outside any stream capture.
//   * All calls on a device share the array, and it keeps whatever the previous call left.  The kernels below write every
//     partial sum they later read, so nothing has to be zeroed between calls -- but two calls whose kernels can run at the
//     same time on one GPU (different streams, different host threads) overwrite each other's partial sums.  Serialise them.
#include "hmb200.h"
#include "internal.h"
#include "../kernels/krylov_kernels.cuh"

#include <cstdio>
#include <mutex>
#include <string>

using namespace hmb200;

namespace {

// The workspace itself.  In HOST code this name is only a handle for the CUDA runtime, not a device address: it must never
// be passed to a kernel as is.  Host code asks the runtime for the address on the current device (static_scratch below).
static __device__ __align__(256) unsigned char g_static_scratch[HMB200_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
std::mutex g_addr_mutex;
void *g_scratch_addr[kMaxDevices]; // device address of g_static_scratch on device i, 0 until looked up

int fail(const std::string &msg) {
    hmb200i_set_error(msg.c_str());
    return 1;
}

// Address of the array on the CURRENT device; looked up once per device, with that device current.
int static_scratch(void **ptr) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail(std::string("hmb200_krylov_dots_static: cudaGetDevice failed: ") + cudaGetErrorString(e));
    if (dev < 0 || dev >= kMaxDevices) return fail("hmb200_krylov_dots_static: device index " + std::to_string(dev) + " is outside the supported range 0.." + std::to_string(kMaxDevices - 1));
    std::lock_guard<std::mutex> lock(g_addr_mutex);
    if (!g_scratch_addr[dev]) {
        void *p = nullptr;
        e       = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p) return fail(std::string("hmb200_krylov_dots_static: cudaGetSymbolAddress failed on device ") + std::to_string(dev) + ": " + cudaGetErrorString(e));
        g_scratch_addr[dev] = p;
    }
    *ptr = g_scratch_addr[dev];
    return 0;
}

template <typename T>
int dots_static(const void *V, int64_t ldv, int k, const void *w, int64_t n, void *h, int chunks, cudaStream_t s) {
    void *scratch = nullptr;
    if (int rc = static_scratch(&scratch)) return rc;
    // same two launches as hmb200_krylov_dots: partial[chunk*k + j] for every (chunk, j), then one fixed-order sum per j
    krylov_dots_kernel<T><<<chunks, kKrylovThreads, 0, s>>>((const T *)V, ldv, k, (const T *)w, n, (T *)scratch);
    krylov_reduce_kernel<T><<<(k + 7) / 8, 256, 0, s>>>((const T *)scratch, chunks, k, (T *)h);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("hmb200_krylov_dots_static: launch failed: ") + cudaGetErrorString(e));
    return 0;
}

} // namespace

extern "C" int hmb200_krylov_dots_static(const void *V, int64_t ldv, int k, const void *w, int64_t n, void *h, int elem_bytes, void *stream) {
    if (elem_bytes != 8 && elem_bytes != 16) return fail("elem_bytes must be 8 or 16");
    if (n < 0) return fail("negative size");
    if (k <= 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    if (n == 0) { // empty vectors: h = 0, nothing to reduce (a memset node, no allocation)
        const cudaError_t e = cudaMemsetAsync(h, 0, (size_t)elem_bytes * (size_t)k, s);
        return e == cudaSuccess ? 0 : fail(std::string("hmb200_krylov_dots_static: cudaMemsetAsync failed: ") + cudaGetErrorString(e));
    }
    const int64_t chunks = (n + kKrylovChunk - 1) / kKrylovChunk;
    const int64_t need   = chunks * (int64_t)k * elem_bytes; // one partial sum per (chunk of 2048 rows, basis vector)
    if (chunks > INT32_MAX || need > (int64_t)HMB200_STATIC_SCRATCH_BYTES) {
        char msg[320];
        std::snprintf(msg, sizeof msg,
                      "hmb200_krylov_dots_static: n = %lld rows x k = %d vectors needs %lld bytes of partial sums, the built-in workspace holds %lld "
                      "(ceil(n/2048) * k * elem_bytes must not exceed it); use hmb200_krylov_dots with a workspace of your own",
                      (long long)n, k, (long long)need, (long long)HMB200_STATIC_SCRATCH_BYTES);
        return fail(msg);
    }
    return elem_bytes == 8 ? dots_static<double>(V, ldv, k, w, n, h, (int)chunks, s) : dots_static<double2>(V, ldv, k, w, n, h, (int)chunks, s);
}
