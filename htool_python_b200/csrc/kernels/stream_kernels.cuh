// stream_kernels.cuh -- the two streaming phases of the H-matrix product as ONE persistent, warp-specialised kernel
// template (sm_100a):
//
//   PHASE 1:  t_j = V_j x[cols_j]            for every low-rank leaf j   (reference: first gemv of
//             lib/htool/include/htool/hmatrix/lrmat/linalg/add_lrmat_vector_product.hpp:15-16)
//   PHASE 2:  y[bin] = beta y[bin] + alpha (sum_j U_j[bin rows] t_j + sum_d D_d[bin rows] x[cols_d])
//             (second gemv :17 and the dense gemv of matrix/linalg/add_matrix_vector_product.hpp:10-23; the leaf loop
//             and its reduction are hmatrix/linalg/add_hmatrix_vector_product.hpp:90-103 / :139-169)
//
// One CTA per SM.  Warp `NW` is the producer: it claims work units (a target bin in phase 2, a run of packets in
// phase 1) from a global counter and streams their packets through a ring of shared-memory stages with 1-D bulk TMA
// copies (cp.async.bulk, one per packet + one small one per tile for the operand slice), signalling mbarriers.
// Warps 0..NW-1 consume: each tile of a packet is taken by one warp (round robin), lanes own output rows, the
// per-bin cross-warp reduction goes through shared memory, and exactly one CTA writes each output row: no atomics.
// The kernel is HBM-bound by design: 1 FMA per 8 (16) bytes streamed.
#pragma once
#include "../capi/plan.h"
#include "ptx.cuh"
#include "scalar.cuh"

namespace hmb200 {

constexpr int kMaxPeers = 8;

struct StreamParams {
    const uint8_t *stream;
    const StageEntry *stages;
    const ZCopy *zcopies;
    const WorkUnit *units;
    uint32_t n_units;
    uint32_t n_zcopies;
    uint32_t *counter;
    void *xz;               // T*: x | t | partials
    void *y;                // T*: phase-2 output (cluster numbering, or user numbering when out_perm != nullptr)
    const int32_t *out_perm; // optional fused cluster_to_user scatter: y[out_perm[row]] (perm already made local)
    double alpha_re, alpha_im, beta_re, beta_im;
    int stage_bytes, z_bytes, n_stages, consumer_warps;
    int pdl;                // launched as a programmatic dependent of the previous kernel: the packets (which no kernel writes) start
                            // streaming at once, the operand slices (x / t, written by the predecessor) wait for griddepcontrol.wait
    // phase 2 only: the all-gather of the row slabs fused into the epilogue.  Every output row is also stored into the
    // replicated vector of each peer GPU (peer-mapped memory over NVLink; pointers already offset to this rank's slab);
    // when the whole grid is done the last CTA publishes this rank's epoch to every rank's flag word (release, system scope).
    int n_peers;            // number of entries of peer_y (other ranks); 0 = plain product
    int n_flags;            // number of entries of peer_flags (all ranks, own included); 0 = no signalling
    void *peer_y[kMaxPeers];
    unsigned long long *peer_flags[kMaxPeers]; // each points at the word "rank <me> finished epoch e" in that rank's flag block
    unsigned long long *epoch;                 // this rank's product count for the buffer being written
    uint32_t *done;                            // CTA completion counter (self-resetting)
};

constexpr int kBarrierBytes = 256;

__host__ __device__ inline size_t stream_smem_bytes(int phase, int elem_bytes, int stage_bytes, int z_bytes, int n_stages, int consumer_warps) {
    size_t red = phase == 2 ? (size_t)2 * consumer_warps * kMaxBinRows * elem_bytes : 0;
    return kBarrierBytes + red + (size_t)n_stages * (stage_bytes + z_bytes);
}

template <typename T, int PHASE>
__global__ void __launch_bounds__(17 * 32, 1) hmatrix_stream_kernel(const StreamParams p) {
    using S = Scalar<T>;
    extern __shared__ __align__(128) uint8_t smem[];
    const int NW   = p.consumer_warps;
    const int NS   = p.n_stages;
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    const uint32_t bar_base  = ptx::smem_u32(smem);
    const size_t red_bytes   = PHASE == 2 ? (size_t)2 * NW * kMaxBinRows * sizeof(T) : 0;
    uint8_t *stage_base      = smem + kBarrierBytes + red_bytes;
    const uint32_t stage_sz  = (uint32_t)(p.stage_bytes + p.z_bytes);
    auto full_bar  = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (NS + s); };

    if (threadIdx.x == 0) {
        for (int s = 0; s < NS; s++) {
            ptx::mbar_init(full_bar(s), 1);
            ptx::mbar_init(empty_bar(s), NW);
        }
        ptx::fence_mbar_init();
    }
    __syncthreads();
    ptx::grid_launch_dependents(); // the next kernel of the product may be scheduled as soon as an SM frees up (no-op without PDL)

    if (warp == NW) {
        // ============================== producer warp ==============================
        bool dep_pending = p.pdl != 0;
        // All schedule metadata is prefetched so that no global-load latency sits on the per-stage critical path:
        //   * the NEXT work unit is claimed (atomicAdd) while the current one streams;
        //   * the 32-byte StageEntry records of a unit are loaded 32 at a time, one per lane, and handed out by shuffles;
        //   * the ZCopy records live in a 64-entry register window (two coalesced 32-entry loads, the second one in
        //     flight while the first is consumed).
        const uint64_t pol_stream = ptx::policy_evict_first(); // factor bytes are read exactly once
        const uint64_t pol_vec    = ptx::policy_evict_last();  // x / t slices are re-read by many tiles: keep in L2
        const T *xz = (const T *)p.xz;
        const uint2 *ztab = reinterpret_cast<const uint2 *>(p.zcopies);
        uint32_t it = 0;
        uint32_t zw = 0xffffffffu; // first ZCopy index held by z_cur (lane l holds zw + l); z_nxt holds zw + 32 + l
        uint2 z_cur = make_uint2(0, 0), z_nxt = make_uint2(0, 0);
        auto load_z = [&](uint32_t base) { // coalesced 256 B load, zero beyond the table
            const uint32_t i = base + lane;
            return i < p.n_zcopies ? __ldg(ztab + i) : make_uint2(0, 0);
        };
        uint32_t u = 0;
        if (lane == 0) u = atomicAdd(p.counter, 1u);
        u = __shfl_sync(0xffffffffu, u, 0);
        while (u < p.n_units) {
            uint32_t u_next = 0;
            if (lane == 0) u_next = atomicAdd(p.counter, 1u); // claimed early: its latency hides behind this unit
            const uint2 wraw = __ldg(reinterpret_cast<const uint2 *>(p.units + u));
            WorkUnit wu;
            wu.first_stage = wraw.x;
            wu.n_stages    = wraw.y;
            for (uint32_t sb = wu.first_stage; sb < wu.first_stage + wu.n_stages; sb += 32) {
                const uint32_t cnt = min(32u, wu.first_stage + wu.n_stages - sb);
                uint4 me0 = make_uint4(0, 0, 0, 0), me1 = make_uint4(0, 0, 0, 0);
                if ((uint32_t)lane < cnt) {
                    me0 = __ldg(reinterpret_cast<const uint4 *>(p.stages + sb + lane));
                    me1 = __ldg(reinterpret_cast<const uint4 *>(p.stages + sb + lane) + 1);
                }
                for (uint32_t j = 0; j < cnt; j++, it++) {
                    const uint32_t off_lo = __shfl_sync(0xffffffffu, me0.x, j), off_hi = __shfl_sync(0xffffffffu, me0.y, j);
                    const uint32_t bytes = __shfl_sync(0xffffffffu, me0.z, j), z_first = __shfl_sync(0xffffffffu, me0.w, j);
                    const uint32_t n_z = __shfl_sync(0xffffffffu, me1.x, j), z_bytes = __shfl_sync(0xffffffffu, me1.y, j);
                    const uint64_t gmem_off = (uint64_t)off_lo | ((uint64_t)off_hi << 32);
                    // make the window cover z_first (stages are visited in table order inside a unit, so this is a slide
                    // by 32 in the common case and a reload after a jump to another unit)
                    if (n_z) {
                        if (zw == 0xffffffffu || z_first < zw || z_first >= zw + 64) {
                            zw    = z_first & ~31u;
                            z_cur = load_z(zw);
                            z_nxt = load_z(zw + 32);
                        } else if (z_first >= zw + 32) {
                            zw += 32;
                            z_cur = z_nxt;
                            z_nxt = load_z(zw + 32);
                        }
                    }
                    const int slot       = it % NS;
                    const uint32_t par   = (it / NS) & 1u;
                    const uint32_t sbase = ptx::smem_u32(stage_base + (size_t)slot * stage_sz);
                    ptx::mbar_wait(empty_bar(slot), par ^ 1u);
                    if (lane == 0) {
                        ptx::mbar_arrive_expect_tx(full_bar(slot), bytes + z_bytes);
                        ptx::bulk_g2s(sbase, p.stream + gmem_off, bytes, full_bar(slot), pol_stream);
                    }
                    if (dep_pending) { // first packet under way; everything below reads what the previous kernel wrote
                        ptx::grid_dependency_wait();
                        dep_pending = false;
                    }
                    __syncwarp();
                    for (uint32_t b = z_first; b < z_first + n_z; b += 32) {
                        if (b >= zw + 32) { // slide (the chunk [b, b+32) must lie inside [zw, zw+64))
                            zw += 32;
                            z_cur = z_nxt;
                            z_nxt = load_z(zw + 32);
                        }
                        const uint32_t rel = b + lane - zw; // < 64
                        const uint32_t ax = __shfl_sync(0xffffffffu, z_cur.x, rel & 31), ay = __shfl_sync(0xffffffffu, z_cur.y, rel & 31);
                        const uint32_t bx = __shfl_sync(0xffffffffu, z_nxt.x, rel & 31), by = __shfl_sync(0xffffffffu, z_nxt.y, rel & 31);
                        const uint32_t zx = rel < 32 ? ax : bx, zy = rel < 32 ? ay : by;
                        if (b + lane < z_first + n_z) {
                            const uint32_t zb = zy & 0xffffu, zdst = zy >> 16;
                            ptx::bulk_g2s(sbase + p.stage_bytes + zdst, xz + zx, zb, full_bar(slot), pol_vec);
                        }
                    }
                }
            }
            u = __shfl_sync(0xffffffffu, u_next, 0);
        }
        // terminate packet: a header written through the generic proxy, released by a plain arrive
        {
            const int slot     = it % NS;
            const uint32_t par = (it / NS) & 1u;
            ptx::mbar_wait(empty_bar(slot), par ^ 1u);
            if (lane == 0) {
                PacketHeader *h = reinterpret_cast<PacketHeader *>(stage_base + (size_t)slot * stage_sz);
                h->n_pieces = 0;
                h->flags    = (uint16_t)kFlagTerminate;
                ptx::mbar_arrive(full_bar(slot));
            }
        }
        return;
    }
    if (warp > NW) return;

    // ============================== consumer warps ==============================
    T *xz = (T *)p.xz;
    T acc0 = S::zero(), acc1 = S::zero();
    int bin_parity = 0;
    for (uint32_t it = 0;; it++) {
        const int slot     = it % NS;
        const uint32_t par = (it / NS) & 1u;
        ptx::mbar_wait(full_bar(slot), par);
        const uint8_t *pk       = stage_base + (size_t)slot * stage_sz;
        const uint4 hraw        = *reinterpret_cast<const uint4 *>(pk);
        const uint32_t n_pieces = hraw.x, piece_base = hraw.y, row0 = hraw.z;
        const uint32_t bin_rows = hraw.w & 0xffffu, flags = hraw.w >> 16;
        if (flags & kFlagTerminate) break;
        const T *zarea = reinterpret_cast<const T *>(pk + p.stage_bytes);

        int first = (int)((NW - (piece_base % NW) + warp) % NW); // pieces with (piece_base + i) % NW == warp
        for (uint32_t i = first; i < n_pieces; i += NW) {
            const uint4 d       = *reinterpret_cast<const uint4 *>(pk + 16 + 16 * i);
            const T *A          = reinterpret_cast<const T *>(pk + d.x);
            const T *z          = zarea + (d.z & 0xffffu);
            const int nc        = (int)(d.z >> 16);
            const int nr        = (int)(d.w & 0xffu);
            if (PHASE == 2) {
                // tile nr x nc (ld nr) at bin rows [ro, ro+nr): lane owns bin rows lane and lane+32
                const int ro  = (int)((d.w >> 8) & 0xffu);
                const int i0  = lane - ro, i1 = lane + 32 - ro;
                const bool a0 = (unsigned)i0 < (unsigned)nr, a1 = (unsigned)i1 < (unsigned)nr;
                const T *p0   = A + (a0 ? i0 : 0);
                const T *p1   = A + (a1 ? i1 : 0);
                T b0 = S::zero(), b1 = S::zero(), c0 = S::zero(), c1 = S::zero(), e0 = S::zero(), e1 = S::zero();
                int c = 0;
                for (; c + 4 <= nc; c += 4) { // 12 shared loads in flight per lane, 4 independent FMA chains per row
                    const T z0 = z[c], z1 = z[c + 1], z2 = z[c + 2], z3 = z[c + 3];
                    const T v00 = a0 ? p0[0] : S::zero(), v10 = a1 ? p1[0] : S::zero();
                    const T v01 = a0 ? p0[nr] : S::zero(), v11 = a1 ? p1[nr] : S::zero();
                    const T v02 = a0 ? p0[2 * nr] : S::zero(), v12 = a1 ? p1[2 * nr] : S::zero();
                    const T v03 = a0 ? p0[3 * nr] : S::zero(), v13 = a1 ? p1[3 * nr] : S::zero();
                    p0 += 4 * nr;
                    p1 += 4 * nr;
                    S::fma(acc0, v00, z0);
                    S::fma(acc1, v10, z0);
                    S::fma(b0, v01, z1);
                    S::fma(b1, v11, z1);
                    S::fma(c0, v02, z2);
                    S::fma(c1, v12, z2);
                    S::fma(e0, v03, z3);
                    S::fma(e1, v13, z3);
                }
                if (c + 2 <= nc) {
                    const T z0 = z[c], z1 = z[c + 1];
                    const T v00 = a0 ? p0[0] : S::zero(), v10 = a1 ? p1[0] : S::zero();
                    const T v01 = a0 ? p0[nr] : S::zero(), v11 = a1 ? p1[nr] : S::zero();
                    p0 += 2 * nr;
                    p1 += 2 * nr;
                    S::fma(acc0, v00, z0);
                    S::fma(acc1, v10, z0);
                    S::fma(b0, v01, z1);
                    S::fma(b1, v11, z1);
                    c += 2;
                }
                if (c < nc) {
                    const T z0 = z[c];
                    const T v00 = a0 ? p0[0] : S::zero(), v10 = a1 ? p1[0] : S::zero();
                    S::fma(c0, v00, z0);
                    S::fma(c1, v10, z0);
                }
                acc0 = S::add(S::add(acc0, b0), S::add(c0, e0));
                acc1 = S::add(S::add(acc1, b1), S::add(c1, e1));
            } else {
                // tile lr x nq (ld lr): lane = g*lr + l owns row l of the columns k = g, g+G, ...; G = 32 / lr column groups
                // (G and ceil(32768/lr) come precomputed in the descriptor: no integer division in the kernel)
                const int lr = nr, nq = nc;
                const int G  = (int)((d.w >> 8) & 0xffu);
                const int g  = (int)(((uint32_t)lane * (d.w >> 16)) >> 15);
                T s0 = S::zero(), s1 = S::zero(), s2 = S::zero(), s3 = S::zero();
                if (g < G) {
                    const T *pa   = A + lane; // element (l, k=g) sits at g*lr + l = lane
                    const T *pz   = z + g;
                    const int sa  = G * lr;
                    int k = g;
                    for (; k + 3 * G < nq; k += 4 * G) {
                        const T a0 = pa[0], a1 = pa[sa], a2 = pa[2 * sa], a3 = pa[3 * sa];
                        const T z0 = pz[0], z1 = pz[G], z2 = pz[2 * G], z3 = pz[3 * G];
                        pa += 4 * sa;
                        pz += 4 * G;
                        S::fma(s0, a0, z0);
                        S::fma(s1, a1, z1);
                        S::fma(s2, a2, z2);
                        S::fma(s3, a3, z3);
                    }
                    for (; k < nq; k += G) {
                        S::fma(s0, pa[0], pz[0]);
                        pa += sa;
                        pz += G;
                    }
                }
                s0 = S::add(S::add(s0, s1), S::add(s2, s3));
                for (int gg = 1; gg < G; gg++) {
                    const T o = S::shfl(s0, lane < lr ? lane + gg * lr : lane);
                    if (lane < lr) s0 = S::add(s0, o);
                }
                if (lane < lr) xz[d.y + lane] = s0;
            }
        }
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(empty_bar(slot));

        if (PHASE == 2 && (flags & kFlagLastOfBin)) {
            T *red = reinterpret_cast<T *>(smem + kBarrierBytes) + (size_t)bin_parity * NW * kMaxBinRows;
            red[warp * kMaxBinRows + lane]      = acc0;
            red[warp * kMaxBinRows + 32 + lane] = acc1;
            ptx::named_bar_sync(1, NW * 32);
            // every (target, row) pair is taken by one consumer thread: target 0 is this rank's y, targets 1.. are the peers'
            // replicated vectors (fused all-gather).  With 16 consumer warps and 7 peers each thread issues exactly one store, so
            // the remote stores of a bin leave in parallel instead of trailing behind two warps.  All targets get the same bits
            // (same summation order).
            const T alpha = S::make(p.alpha_re, p.alpha_im), beta = S::make(p.beta_re, p.beta_im);
            const uint32_t n_items = (uint32_t)(1 + p.n_peers) * kMaxBinRows;
            for (uint32_t idx = threadIdx.x; idx < n_items; idx += (uint32_t)NW * 32u) {
                const uint32_t r_ = idx % kMaxBinRows, tgt = idx / kMaxBinRows;
                if (r_ >= bin_rows) continue;
                T s = red[r_];
                for (int w = 1; w < NW; w++) s = S::add(s, red[w * kMaxBinRows + r_]);
                T r = S::mul(alpha, s);
                if (tgt == 0) {
                    T *y = (T *)p.y;
                    const size_t o = p.out_perm ? (size_t)p.out_perm[row0 + r_] : (size_t)(row0 + r_);
                    if (!S::is_zero(beta)) r = S::add(r, S::mul(beta, y[o])); // beta == 0 must not propagate NaNs of y (BLAS)
                    y[o] = r;
                } else {
                    reinterpret_cast<T *>(p.peer_y[tgt - 1])[row0 + r_] = r; // beta = 0, cluster numbering (fused path)
                }
            }
            acc0 = S::zero();
            acc1 = S::zero();
            bin_parity ^= 1;
        }
    }
    if (PHASE == 2 && p.n_flags > 0) {
        // every thread that stored rows fences at system scope, the CTA counts itself done, and the last CTA of the grid
        // publishes the new epoch to all ranks (the classic last-block pattern, extended to system scope)
        __threadfence_system();
        ptx::named_bar_sync(1, NW * 32);
        if (threadIdx.x == 0) {
            const uint32_t old = atomicAdd(p.done, 1u);
            if (old == gridDim.x - 1) {
                *p.done = 0;
                const unsigned long long e = *p.epoch + 1ull;
                *p.epoch = e;
                __threadfence_system();
                for (int q = 0; q < p.n_flags; q++) ptx::st_release_sys_u64(p.peer_flags[q], e);
            }
        }
    }
}

// waits until every rank has published at least this rank's own epoch for the buffer (flags: n_ranks words in LOCAL memory
// written by the peers' phase-2 kernels).  Bounded spin: after ~`max_cycles` clocks the error word is set and the kernel exits.
__global__ void peer_wait_kernel(const unsigned long long *flags, int n_ranks, const unsigned long long *epoch, int *err, long long max_cycles) {
    ptx::grid_dependency_wait(); // `epoch` is written by the product kernel in front of this one (no-op without PDL)
    const int q = threadIdx.x;
    if (q >= n_ranks) return;
    const unsigned long long target = *epoch;
    const long long t0 = clock64();
    while (ptx::ld_acquire_sys_u64(flags + q) < target) {
        if (clock64() - t0 > max_cycles) {
            *(volatile int *)err = 1 + q; // mapped host memory: a plain store (no PCIe atomics needed), made visible by the fence
            __threadfence_system();
            break;
        }
        __nanosleep(200);
    }
}

// t[final + l] = sum_q partial[first + q*len + l]   (one warp per segment; deterministic order)
template <typename T>
__global__ void __launch_bounds__(256) reduce_partials_kernel(const ReduceSeg *segs, uint32_t n_segs, T *xz) {
    using S = Scalar<T>;
    ptx::grid_launch_dependents();
    ptx::grid_dependency_wait(); // the partials are written by phase 1 (no-op without PDL)
    const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane   = threadIdx.x & 31;
    if (w >= n_segs) return;
    const uint4 sg = __ldg(reinterpret_cast<const uint4 *>(segs + w));
    const int len = (int)sg.w, G = 32 / len, l = lane % len, g = lane / len;
    // latency-bound (a segment of a wide V tile has > 100 partials): 8 independent loads in flight per lane, fixed order
    T s0 = S::zero(), s1 = S::zero(), s2 = S::zero(), s3 = S::zero();
    if (g < G) {
        const T *pp = xz + sg.y + l;
        const size_t st = (size_t)G * len;
        uint32_t q = g;
        const T *pq = pp + (size_t)q * len;
        for (; q + 7 * G < sg.z; q += 8 * G, pq += 8 * st) {
            const T a0 = pq[0], a1 = pq[st], a2 = pq[2 * st], a3 = pq[3 * st];
            const T a4 = pq[4 * st], a5 = pq[5 * st], a6 = pq[6 * st], a7 = pq[7 * st];
            s0 = S::add(S::add(s0, a0), a4);
            s1 = S::add(S::add(s1, a1), a5);
            s2 = S::add(S::add(s2, a2), a6);
            s3 = S::add(S::add(s3, a3), a7);
        }
        for (; q < sg.z; q += G, pq += st) s0 = S::add(s0, pq[0]);
    }
    s0 = S::add(S::add(s0, s1), S::add(s2, s3));
    for (int gg = 1; gg < G; gg++) {
        const T o = S::shfl(s0, lane < len ? lane + gg * len : lane);
        if (lane < len) s0 = S::add(s0, o);
    }
    if (lane < len) xz[sg.x + lane] = s0;
}

// permutations (bit-exact data movement; clustering/cluster_node.hpp:149-175)
template <typename T>
__global__ void permute_gather_kernel(const int32_t *__restrict__ perm, int off, int n, const T *__restrict__ in, T *__restrict__ out, int mu) {
    const size_t total = (size_t)n * mu;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const size_t c = t / n, i = t - c * n;
        out[t] = in[c * n + (size_t)(perm[off + i] - off)];
    }
}
template <typename T>
__global__ void permute_scatter_kernel(const int32_t *__restrict__ perm, int off, int n, const T *__restrict__ in, T *__restrict__ out, int mu) {
    const size_t total = (size_t)n * mu;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const size_t c = t / n, i = t - c * n;
        out[c * n + (size_t)(perm[off + i] - off)] = in[t];
    }
}

} // namespace hmb200
