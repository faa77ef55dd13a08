// mm_kernels.cuh -- multi-RHS H-matrix product Y = beta Y + alpha op(A) X on the FP64 tensor cores (sm_100a:
// mma.sync.m8n8k4.f64, SASS DMMA), row-major blocks exactly as the reference's *_row_major interfaces
// (hmatrix/linalg/add_hmatrix_matrix_product_row_major.hpp:57-178; the two gemms of
// lrmat/linalg/add_lrmat_matrix_product_row_major.hpp:11-27; the dense gemm of
// matrix/linalg/add_matrix_matrix_product_row_major.hpp:23-46).
//
// Both phases are the SAME persistent kernel: a stream of panel GEMMs  C (rows x mu) += A (rows x K) B (K x mu)  (plan.h):
//   PHASE 1:  A = the V rows of all low-rank leaves that share one source cluster, stacked (<= 128 rows); K = a chunk of the
//             cluster's columns; B = those rows of X; C = the leaves' T rows (row map in the unit's last packet; a cluster wider
//             than a unit leaves partial rows for reduce_rows_kernel)
//   PHASE 2:  A = [U_1 | U_2 | ... | D_1 | ...] of ONE target bin (<= 64 rows) concatenated along K; B = the T rows / X rows of
//             those leaves in the same order; C = the bin's rows of Y.  One CTA owns a bin: no atomics.  The bin's tile is
//             staged in shared memory and stored as full rows to the local Y and -- the all-gather of the block-row product,
//             MPI_Allgatherv of add_distributed_operator_matrix_product_row_major_global_to_global.hpp:75 -- to every peer GPU.
// Warp `kMmConsumerWarps` is the producer: per packet ONE bulk TMA copy of the fragment-major panel chunk (L2 evict_first) and
// one per run of operand rows (evict_last) into a ring of shared-memory stages guarded by full / empty mbarriers.  The eight
// consumer warps tile the panel (WM x WN warps, MT m-tiles x NT n-tiles of 8 x 8 each) and keep C in DMMA accumulator registers
// for the whole unit; A fragments come with one conflict-free 128-bit shared load per lane and k-group, B fragments from
// operand rows at a padded stride (ld = 4 or 12 mod 16 doubles).  The next k-group's fragments are loaded before the DMMAs of
// the current one are issued.  complex<double>: one complex DMMA step = 4 real DMMAs on the (re, im) parts.
#pragma once
#include "../capi/plan.h"
#include "ptx.cuh"
#include "scalar.cuh"
#include "stream_kernels.cuh"

namespace hmb200 {

constexpr int kMmConsumerWarps = 8;
constexpr int kMmThreads       = (kMmConsumerWarps + 1) * 32;
constexpr int kMmYTileBytes    = kMmRows2 * 72 * 8; // staged bin tile: 64 rows x (64 + 8 pad) doubles = 64 x (32 + 4) complex

struct MmParams {
    const uint8_t *stream;
    const StageEntry *stages;
    const ZCopy *zcopies;
    const WorkUnit *units;
    uint32_t n_units;
    uint32_t *counter;
    void *ws;  // T*: workspace rows [x | t | partials | zero rows], row stride ld
    int ld;    // workspace row stride in elements
    int ncols; // right-hand sides of this pass
    void *y;   // phase 2: output block, row-major, stride ldy, already offset to the pass's first column
    long long ldy;
    double alpha_re, alpha_im, beta_re, beta_im;
    int stage_bytes, z_rows, n_stages; // bytes of a stage's packet area, operand rows per stage, ring depth
    // phase 2: the all-gather fused into the bin epilogue (same protocol as StreamParams)
    int n_peers, n_flags;
    void *peer_y[kMaxPeers];                   // peers' replicated blocks, offset to this rank's slab and the pass's first column
    unsigned long long *peer_flags[kMaxPeers];
    unsigned long long *epoch;
    uint32_t *done;
};

__host__ __device__ inline size_t mm_stage_size(int elem_bytes, int stage_bytes, int z_rows, int ld) { return (size_t)stage_bytes + (size_t)z_rows * ld * elem_bytes; }
__host__ __device__ inline size_t mm_smem_bytes(int phase, int elem_bytes, int stage_bytes, int z_rows, int n_stages, int ld) {
    return 256 + (phase == 2 ? kMmYTileBytes : 0) + (size_t)n_stages * mm_stage_size(elem_bytes, stage_bytes, z_rows, ld);
}

__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// accumulators of one warp: MT m-tiles x NT n-tiles (x re / im)
template <typename T, int MT, int NT>
struct MmAcc {
    static constexpr int C = sizeof(T) == 8 ? 1 : 2;
    double v[MT][NT][C][2];
    __device__ __forceinline__ void clear() {
#pragma unroll
        for (int i = 0; i < MT; i++)
#pragma unroll
            for (int j = 0; j < NT; j++)
#pragma unroll
                for (int c = 0; c < C; c++) v[i][j][c][0] = v[i][j][c][1] = 0.0;
    }
};

// B fragments of one k-group: real 2 k4-steps x NT doubles, complex NT double2
template <typename T, int NT>
struct MmBFrag {
    static constexpr int KS = sizeof(T) == 8 ? 2 : 1;
    T v[KS][NT];
};

// acc += panel chunk(s) of this packet x operand rows.  chunks: fragment-major (plan.h), 16 bytes per lane and (kg, mt);
// Z: operand rows of the packet (stride ld).  The warp owns m-tiles wm, wm + WM, ... (nmt of them) and n-tiles c_base/8 ... (nnt).
template <typename T, int MT, int NT>
__device__ __forceinline__ void mm_packet(MmAcc<T, MT, NT> &acc, const double2 *chunks, int mt_panel, int n_kg, int wm, int WM, int nmt,
                                          const T *Z, int ld, int c_base, int nnt, int lane) {
    constexpr int KG = sizeof(T) == 8 ? 8 : 4;
    constexpr int KS = MmBFrag<T, NT>::KS;
    const int gid = lane >> 2, tig = lane & 3;
    const double2 *ap = chunks + (size_t)wm * 32 + lane;
    const int a_mt    = WM * 32, a_kg = mt_panel * 32;
    const T *bp       = Z + (size_t)tig * ld + c_base + gid;
    double2 a[MT];
    MmBFrag<T, NT> b;
    auto load = [&](double2 (&aa)[MT], MmBFrag<T, NT> &bb) {
#pragma unroll
        for (int i = 0; i < MT; i++)
            if (i < nmt) aa[i] = ap[i * a_mt];
#pragma unroll
        for (int ks = 0; ks < KS; ks++)
#pragma unroll
            for (int j = 0; j < NT; j++)
                if (j < nnt) bb.v[ks][j] = bp[(size_t)(4 * ks) * ld + 8 * j];
    };
    load(a, b);
    for (int kg = 0; kg < n_kg; kg++) {
        double2 an[MT];
        MmBFrag<T, NT> bn;
        const bool more = kg + 1 < n_kg;
        if (more) {
            ap += a_kg;
            bp += (size_t)KG * ld;
            load(an, bn);
        }
        if constexpr (sizeof(T) == 8) {
#pragma unroll
            for (int ks = 0; ks < 2; ks++)
#pragma unroll
                for (int i = 0; i < MT; i++)
                    if (i < nmt) {
                        const double av = ks == 0 ? a[i].x : a[i].y;
#pragma unroll
                        for (int j = 0; j < NT; j++)
                            if (j < nnt) dmma884(acc.v[i][j][0][0], acc.v[i][j][0][1], av, b.v[ks][j]);
                    }
        } else {
#pragma unroll
            for (int i = 0; i < MT; i++)
                if (i < nmt) {
#pragma unroll
                    for (int j = 0; j < NT; j++)
                        if (j < nnt) { // (ar + i ai)(br + i bi): re += ar br - ai bi ; im += ar bi + ai br
                            dmma884(acc.v[i][j][0][0], acc.v[i][j][0][1], a[i].x, b.v[0][j].x);
                            dmma884(acc.v[i][j][1][0], acc.v[i][j][1][1], a[i].x, b.v[0][j].y);
                        }
#pragma unroll
                    for (int j = 0; j < NT; j++)
                        if (j < nnt) {
                            dmma884(acc.v[i][j][0][0], acc.v[i][j][0][1], -a[i].y, b.v[0][j].y);
                            dmma884(acc.v[i][j][1][0], acc.v[i][j][1][1], a[i].y, b.v[0][j].x);
                        }
                }
        }
        if (more) {
#pragma unroll
            for (int i = 0; i < MT; i++) a[i] = an[i];
            b = bn;
        }
    }
}

template <typename T>
__device__ __forceinline__ T mm_make(double re, double im);
template <>
__device__ __forceinline__ double mm_make<double>(double re, double) { return re; }
template <>
__device__ __forceinline__ double2 mm_make<double2>(double re, double im) { return make_double2(re, im); }

// WN = 2: "wide" pass (4 x 2 warps, each MT x NT tiles); WN = 1: "narrow" pass (8 x 1 warps).  MT * 8 / WN m-tiles cover the
// panel: 16 (128 rows) in phase 1, 8 (64 rows) in phase 2.
template <typename T, int PHASE, int WN, int MT, int NT>
__global__ void __launch_bounds__(kMmThreads, 1) hmatrix_mm_kernel(const MmParams p) {
    constexpr int NW = kMmConsumerWarps;
    constexpr int WM = NW / WN;
    static_assert(WM * MT * 8 == (PHASE == 1 ? kMmRows1 : kMmRows2), "warp tiling must cover the panel");
    extern __shared__ __align__(128) uint8_t smem[];
    const int NS   = p.n_stages;
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int ld   = p.ld;

    const uint32_t bar_base = ptx::smem_u32(smem);
    uint8_t *ytile_raw      = smem + 256;
    uint8_t *stage_base     = smem + 256 + (PHASE == 2 ? kMmYTileBytes : 0);
    const uint32_t stage_sz = (uint32_t)mm_stage_size((int)sizeof(T), p.stage_bytes, p.z_rows, ld);
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

    if (warp == NW) {
        // ============================== producer warp ==============================
        const uint64_t pol_stream = ptx::policy_evict_first(); // panel chunks are read exactly once
        const uint64_t pol_vec    = ptx::policy_evict_last();  // operand rows are shared by many units: keep them in L2
        const T *ws               = (const T *)p.ws;
        uint32_t it = 0;
        uint32_t u = 0;
        if (lane == 0) u = atomicAdd(p.counter, 1u);
        u = __shfl_sync(0xffffffffu, u, 0);
        while (u < p.n_units) {
            uint32_t u_next = 0;
            if (lane == 0) u_next = atomicAdd(p.counter, 1u); // claimed early: its latency hides behind this unit
            const uint2 wraw = __ldg(reinterpret_cast<const uint2 *>(p.units + u));
            for (uint32_t sb = wraw.x; sb < wraw.x + wraw.y; sb += 32) { // stage records: 32 per coalesced load, handed out by shuffles
                const uint32_t cnt = min(32u, wraw.x + wraw.y - sb);
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
                    uint2 zc = make_uint2(0, 0);
                    if ((uint32_t)lane < n_z) zc = __ldg(reinterpret_cast<const uint2 *>(p.zcopies + z_first) + lane); // in flight during the wait
                    const int slot       = it % NS;
                    const uint32_t par   = (it / NS) & 1u;
                    const uint32_t sbase = ptx::smem_u32(stage_base + (size_t)slot * stage_sz);
                    ptx::mbar_wait(empty_bar(slot), par ^ 1u);
                    if (lane == 0) {
                        ptx::mbar_arrive_expect_tx(full_bar(slot), bytes + z_bytes * (uint32_t)ld);
                        ptx::bulk_g2s(sbase, p.stream + gmem_off, bytes, full_bar(slot), pol_stream);
                    }
                    __syncwarp();
                    for (uint32_t i = lane; i < n_z; i += 32) { // operand rows: ZCopy counts in elements-per-row units, a row is ld elements
                        if (i >= 32) zc = __ldg(reinterpret_cast<const uint2 *>(p.zcopies + z_first) + i);
                        const uint32_t zb = (zc.y & 0xffffu) * (uint32_t)ld, zdst = (zc.y >> 16) * (uint32_t)ld;
                        ptx::bulk_g2s(sbase + p.stage_bytes + zdst, ws + (size_t)zc.x * ld, zb, full_bar(slot), pol_vec);
                    }
                }
            }
            u = __shfl_sync(0xffffffffu, u_next, 0);
        }
        { // terminate packet: a header written through the generic proxy, released by a plain arrive
            const int slot     = it % NS;
            const uint32_t par = (it / NS) & 1u;
            ptx::mbar_wait(empty_bar(slot), par ^ 1u);
            if (lane == 0) {
                *reinterpret_cast<uint4 *>(stage_base + (size_t)slot * stage_sz) = make_uint4(0u, 0u, 0u, kMmFlagTerminate);
                ptx::mbar_arrive(full_bar(slot));
            }
        }
        return;
    }

    // ============================== consumer warps ==============================
    const int gid = lane >> 2, tig = lane & 3;
    const int wm = warp % WM, wn = warp / WM;
    const int c_base = wn * NT * 8;
    const int ncols8 = (p.ncols + 7) & ~7;
    const int nnt    = max(0, min(NT, (ncols8 - c_base) >> 3)); // n-tiles of this warp that hold right-hand sides
    MmAcc<T, MT, NT> acc;
    acc.clear();
    T *ws = (T *)p.ws;
    for (uint32_t it = 0;; it++) {
        const int slot     = it % NS;
        const uint32_t par = (it / NS) & 1u;
        ptx::mbar_wait(full_bar(slot), par);
        const uint8_t *pk   = stage_base + (size_t)slot * stage_sz;
        const uint4 hraw    = *reinterpret_cast<const uint4 *>(pk);
        const int n_kg      = (int)hraw.x;
        const uint32_t out0 = hraw.y;
        const int rows = (int)(hraw.z & 0xffffu), mt_panel = (int)(hraw.z >> 16);
        const uint32_t flags = hraw.w & 0xffffu, map_off16 = hraw.w >> 16;
        if (flags & kMmFlagTerminate) break;
        const int nmt = wm < mt_panel ? min(MT, (mt_panel - wm + WM - 1) / WM) : 0; // m-tiles wm, wm + WM, ... below mt_panel

        if (nmt > 0 && nnt > 0 && n_kg > 0)
            mm_packet<T, MT, NT>(acc, reinterpret_cast<const double2 *>(pk + 16), mt_panel, n_kg, wm, WM, nmt, reinterpret_cast<const T *>(pk + p.stage_bytes), ld, c_base, nnt, lane);

        if (PHASE == 1) {
            if (flags & kMmFlagLast) { // the unit's rows of T (or partial rows): destination row per panel row from the packet's row map
                const uint32_t *map = reinterpret_cast<const uint32_t *>(pk + 16 * (size_t)map_off16);
#pragma unroll
                for (int i = 0; i < MT; i++) {
                    const int r = 8 * (wm + WM * i) + gid;
                    if (i < nmt && r < rows) {
                        T *orow = ws + (size_t)map[r] * ld;
#pragma unroll
                        for (int j = 0; j < NT; j++)
                            if (j < nnt) {
                                const int col = c_base + 8 * j + 2 * tig;
                                if constexpr (sizeof(T) == 8) {
                                    *reinterpret_cast<double2 *>(orow + col) = make_double2(acc.v[i][j][0][0], acc.v[i][j][0][1]);
                                } else {
                                    orow[col]     = make_double2(acc.v[i][j][0][0], acc.v[i][j][1][0]);
                                    orow[col + 1] = make_double2(acc.v[i][j][0][1], acc.v[i][j][1][1]);
                                }
                            }
                    }
                }
                acc.clear();
            }
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(empty_bar(slot));
        } else {
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(empty_bar(slot));
            if (flags & kMmFlagLast) {
                // stage the bin's tile in shared memory, then all consumer threads store full rows: the local Y (with beta) and
                // the same bits into every peer's replicated block (fused all-gather)
                constexpr int ldt = kMmYTileBytes / kMmRows2 / (int)sizeof(T);
                T *yt = reinterpret_cast<T *>(ytile_raw);
#pragma unroll
                for (int i = 0; i < MT; i++) {
                    const int r = 8 * (wm + WM * i) + gid;
                    if (i < nmt) {
#pragma unroll
                        for (int j = 0; j < NT; j++)
                            if (j < nnt) {
                                const int col = c_base + 8 * j + 2 * tig;
                                if constexpr (sizeof(T) == 8) {
                                    *reinterpret_cast<double2 *>(yt + r * ldt + col) = make_double2(acc.v[i][j][0][0], acc.v[i][j][0][1]);
                                } else {
                                    yt[r * ldt + col]     = make_double2(acc.v[i][j][0][0], acc.v[i][j][1][0]);
                                    yt[r * ldt + col + 1] = make_double2(acc.v[i][j][0][1], acc.v[i][j][1][1]);
                                }
                            }
                    }
                }
                ptx::named_bar_sync(1, NW * 32);
                const T alpha = mm_make<T>(p.alpha_re, p.alpha_im), beta = mm_make<T>(p.beta_re, p.beta_im);
                const bool beta_zero = Scalar<T>::is_zero(beta);
                const uint32_t nc = (uint32_t)p.ncols, per_tgt = (uint32_t)rows * nc, total = (uint32_t)(1 + p.n_peers) * per_tgt;
                for (uint32_t idx = threadIdx.x; idx < total; idx += (uint32_t)NW * 32u) {
                    const uint32_t tgt = idx / per_tgt, rem = idx - tgt * per_tgt;
                    const uint32_t r = rem / nc, c = rem - r * nc;
                    T v = Scalar<T>::mul(alpha, yt[r * ldt + c]); // rows above mt_panel * 8 are never read: r < rows
                    const size_t o = (size_t)(out0 + r) * p.ldy + c;
                    if (tgt == 0) {
                        T *y = (T *)p.y;
                        if (!beta_zero) v = Scalar<T>::add(v, Scalar<T>::mul(beta, y[o])); // beta == 0 must not propagate NaNs of y (BLAS)
                        y[o] = v;
                    } else {
                        reinterpret_cast<T *>(p.peer_y[tgt - 1])[o] = v; // beta = 0 on the fused path
                    }
                }
                ptx::named_bar_sync(1, NW * 32); // the tile is rewritten by the next bin
                acc.clear();
            }
        }
    }
    if (PHASE == 2 && p.n_flags > 0) {
        // every thread that stored rows fences at system scope, the CTA counts itself done, and the last CTA of the grid
        // publishes the new epoch to all ranks (as hmatrix_stream_kernel)
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

// final T rows = sum of the partial rows phase 1 wrote for the column chunks of a split unit, in a fixed order; one warp per
// segment (a run of rows of one leaf), lanes over (row, column pairs)
template <typename T>
__global__ void __launch_bounds__(256) reduce_rows_kernel(const ReduceSeg *segs, const uint32_t *part_rows, uint32_t n_segs, T *ws, int ld, int ncols8) {
    const uint32_t wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    const int lane = threadIdx.x & 31;
    constexpr int V = sizeof(T) == 8 ? 2 : 1; // elements per 16-byte access
    const int cpr   = ncols8 / V;              // 16-byte chunks per row
    for (uint32_t sidx = wid; sidx < n_segs; sidx += nw) {
        const uint4 sg = __ldg(reinterpret_cast<const uint4 *>(segs + sidx));
        const int total = (int)sg.w * cpr;
        T *dst_base = ws + (size_t)sg.x * ld;
        for (uint32_t q0 = 0; q0 < sg.z; q0 += 32) { // part rows: 32 at a time in registers, handed out by shuffles
            const uint32_t nq = min(32u, sg.z - q0);
            const uint32_t my_row = (uint32_t)lane < nq ? __ldg(part_rows + sg.y + q0 + lane) : 0u;
            for (int e = lane; e < ((total + 31) & ~31); e += 32) {
                const bool act = e < total;
                const int l = act ? e / cpr : 0, c = act ? (e - l * cpr) * V : 0;
                double2 s = make_double2(0.0, 0.0);
                if (q0 > 0 && act) s = *reinterpret_cast<const double2 *>(dst_base + (size_t)l * ld + c);
                uint32_t q = 0;
                for (; q + 4 <= nq; q += 4) {
                    const uint32_t r0 = __shfl_sync(0xffffffffu, my_row, q), r1 = __shfl_sync(0xffffffffu, my_row, q + 1);
                    const uint32_t r2 = __shfl_sync(0xffffffffu, my_row, q + 2), r3 = __shfl_sync(0xffffffffu, my_row, q + 3);
                    if (act) {
                        const double2 v0 = *reinterpret_cast<const double2 *>(ws + (size_t)(r0 + l) * ld + c);
                        const double2 v1 = *reinterpret_cast<const double2 *>(ws + (size_t)(r1 + l) * ld + c);
                        const double2 v2 = *reinterpret_cast<const double2 *>(ws + (size_t)(r2 + l) * ld + c);
                        const double2 v3 = *reinterpret_cast<const double2 *>(ws + (size_t)(r3 + l) * ld + c);
                        s.x = (((s.x + v0.x) + v1.x) + v2.x) + v3.x;
                        s.y = (((s.y + v0.y) + v1.y) + v2.y) + v3.y;
                    }
                }
                for (; q < nq; q++) {
                    const uint32_t r0 = __shfl_sync(0xffffffffu, my_row, q);
                    if (act) {
                        const double2 v0 = *reinterpret_cast<const double2 *>(ws + (size_t)(r0 + l) * ld + c);
                        s.x += v0.x;
                        s.y += v0.y;
                    }
                }
                if (act) *reinterpret_cast<double2 *>(dst_base + (size_t)l * ld + c) = s;
            }
        }
    }
}

// row-major (n x mu, stride lds) <-> workspace rows (stride ld): X load of one pass; with col_stride > 0 the source is a
// COLUMN-major block (optionally gathered through the cluster permutation: fuses user_to_cluster + the transposes of
// add_hmatrix_matrix_product.hpp:29-31,190-195)
template <typename T>
__global__ void mm_load_rows_kernel(const T *__restrict__ src, long long lds, int n, int ncols, T *__restrict__ ws, int ld, const int32_t *__restrict__ perm, long long col_stride) {
    const size_t total = (size_t)n * ncols;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const size_t i = t / ncols;
        const int j    = (int)(t - i * ncols);
        ws[i * ld + j] = col_stride > 0 ? src[(size_t)(perm ? perm[i] : (int32_t)i) + (size_t)j * col_stride] : src[i * lds + j];
    }
}
// cluster-numbered row-major result rows -> column-major user block (cluster_to_user per column, :202-204)
template <typename T>
__global__ void mm_store_user_kernel(const T *__restrict__ yc, int n, int ncols, T *__restrict__ dst, const int32_t *__restrict__ perm, long long col_stride) {
    const size_t total = (size_t)n * ncols;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const size_t i = t / ncols;
        const int j    = (int)(t - i * ncols);
        dst[(size_t)(perm ? perm[i] : (int32_t)i) + (size_t)j * col_stride] = yc[i * ncols + j];
    }
}

} // namespace hmb200
