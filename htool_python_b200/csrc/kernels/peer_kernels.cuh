// peer_kernels.cuh -- the two small collectives of the distributed solver over peer-mapped memory (CUDA IPC, NVLink), sm_100a.
//
// They replace, on one node, the MPI calls of the reference's solver iteration:
//   peer_gather_kernel    : all-gather of the local row slabs into the replicated operator input
//                           (MPI_Allgatherv of add_distributed_operator_vector_product_local_to_local.hpp:44)
//   peer_allreduce_kernel : sum over the ranks of the k dot products / of the squared norm of one Arnoldi step
//                           (MPI_Allreduce of lib/hpddm/include/HPDDM_iterative.hpp:549,740), summed in RANK ORDER on
//                           every rank, so all ranks hold bit-identical values
//   peer_publish_kernel + peer_sum_slabs_kernel : the exchange of the TRANSPOSED local-to-local product (MPI_Alltoallv + one
//                           axpy per rank, add_distributed_operator_vector_product_local_to_local.hpp:60-87): every rank leaves
//                           its full-length contribution in ITS OWN peer-mapped buffer, publishes the epoch, and pulls its slab
//                           of every rank's contribution over NVLink, added in rank order
// Protocol: the same epoch flags as the fused product epilogue (stream_kernels.cuh): data stores, fence.sys, one
// st.release.sys of the new epoch into every rank's flag word; consumers spin with ld.acquire.sys (bounded).
#pragma once
#include "ptx.cuh"
#include "stream_kernels.cuh"

namespace hmb200 {

struct PeerGatherParams {
    const void *src;        // local slab (n_local elements)
    long long n_local;
    int n_targets;          // all ranks, own included
    void *dst[kMaxPeers];   // replicated buffers, already offset to this rank's slab
    unsigned long long *flags[kMaxPeers]; // word "rank <me> finished epoch e" in every rank's flag block
    unsigned long long *epoch;
    uint32_t *done;
};

template <typename T>
__global__ void __launch_bounds__(256) peer_gather_kernel(const PeerGatherParams p) {
    const T *src = reinterpret_cast<const T *>(p.src);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < p.n_local; i += (long long)gridDim.x * blockDim.x) {
        const T v = src[i];
        for (int q = 0; q < p.n_targets; q++) reinterpret_cast<T *>(p.dst[q])[i] = v;
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t old = atomicAdd(p.done, 1u);
        if (old == gridDim.x - 1) {
            *p.done = 0;
            const unsigned long long e = *p.epoch + 1ull;
            *p.epoch = e;
            __threadfence_system();
            for (int q = 0; q < p.n_targets; q++) ptx::st_release_sys_u64(p.flags[q], e);
        }
    }
}

// epoch publication on its own: the data of this epoch was written into the rank's own region by the kernels in front of this
// one on the stream (one thread; the fence orders those writes, visible device-wide at the kernel boundary, before the flags)
__global__ void peer_publish_kernel(const PeerGatherParams p) { // uses p.epoch, p.n_targets, p.flags
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const unsigned long long e = *p.epoch + 1ull;
    *p.epoch = e;
    __threadfence_system();
    for (int q = 0; q < p.n_targets; q++) ptx::st_release_sys_u64(p.flags[q], e);
}

struct PeerSumParams {
    const void *src[kMaxPeers]; // rank q's full-length contribution, already offset to THIS rank's slab (peer-mapped)
    int n_ranks;
    long long n;                // elements of this rank's slab
    void *out;                  // local slab: out = beta * out + sum_q src[q]   (q ascending: the same bits on every run)
    double beta[2];
};

template <typename T>
__device__ __forceinline__ T peer_load(const T *p);
template <>
__device__ __forceinline__ double peer_load<double>(const double *p) { // ld.volatile: never served from a stale L1 line
    double v;
    asm volatile("ld.volatile.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}
template <>
__device__ __forceinline__ double2 peer_load<double2>(const double2 *p) {
    double2 v;
    asm volatile("ld.volatile.global.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p) : "memory");
    return v;
}

template <typename T>
__global__ void __launch_bounds__(256) peer_sum_slabs_kernel(const PeerSumParams p) {
    using S = Scalar<T>;
    T *out = reinterpret_cast<T *>(p.out);
    const T beta = S::make(p.beta[0], p.beta[1]);
    const bool beta_zero = S::is_zero(beta);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < p.n; i += (long long)gridDim.x * blockDim.x) {
        T v[kMaxPeers];
#pragma unroll
        for (int q = 0; q < kMaxPeers; q++)
            if (q < p.n_ranks) v[q] = peer_load<T>(reinterpret_cast<const T *>(p.src[q]) + i); // all loads in flight before the sum
        T s = beta_zero ? S::zero() : S::mul(beta, out[i]);
#pragma unroll
        for (int q = 0; q < kMaxPeers; q++)
            if (q < p.n_ranks) s = S::add(s, v[q]);
        out[i] = s;
    }
}

constexpr int kPeerReduceMax = 512; // doubles per all-reduce (a Hessenberg column of a 255-vector complex basis)

struct PeerReduceParams {
    double *vals;           // in: this rank's partial sums (n doubles); out: the sums over all ranks
    int n;                  // doubles (<= kPeerReduceMax)
    int n_ranks, rank;
    double *slots[kMaxPeers];             // every rank's slot array for the current parity: [n_ranks][kPeerReduceMax]
    unsigned long long *flags[kMaxPeers]; // word [rank <me>] of every rank's reduce-flag block for the current parity
    const unsigned long long *my_flags;   // this rank's block (n_ranks words)
    unsigned long long *epoch;
    int *err;
    long long max_cycles;
};

// one CTA: publish my partials to every rank, signal, wait for everybody's, sum in rank order
__global__ void __launch_bounds__(256) peer_allreduce_kernel(const PeerReduceParams p) {
    for (int i = threadIdx.x; i < p.n; i += blockDim.x) {
        const double v = p.vals[i];
        for (int q = 0; q < p.n_ranks; q++) p.slots[q][(size_t)p.rank * kPeerReduceMax + i] = v;
    }
    __threadfence_system();
    __syncthreads();
    __shared__ unsigned long long target;
    if (threadIdx.x == 0) {
        const unsigned long long e = *p.epoch + 1ull;
        *p.epoch = e;
        target   = e;
        for (int q = 0; q < p.n_ranks; q++) ptx::st_release_sys_u64(p.flags[q], e);
    }
    __syncthreads();
    if (threadIdx.x < p.n_ranks) {
        const long long t0 = clock64();
        while (ptx::ld_acquire_sys_u64(p.my_flags + threadIdx.x) < target) {
            if (clock64() - t0 > p.max_cycles) {
                *(volatile int *)p.err = 101 + (int)threadIdx.x; // mapped host memory (see peer_wait_kernel)
                __threadfence_system();
                break;
            }
            __nanosleep(100);
        }
    }
    __syncthreads();
    const volatile double *mine = p.slots[p.rank]; // written by the peers during this kernel: volatile loads bypass the (non-coherent) L1
    for (int i = threadIdx.x; i < p.n; i += blockDim.x) {
        double s = 0.0;
        for (int q = 0; q < p.n_ranks; q++) s += mine[(size_t)q * kPeerReduceMax + i];
        p.vals[i] = s;
    }
}

} // namespace hmb200
