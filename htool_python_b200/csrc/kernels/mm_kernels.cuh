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
// operand rows at a padded stride (ld = 4 or 12 mod 16 doubles).  A unit is ONE software-pipelined K loop that runs across its
// packets: the fragments of the next k-group -- also when it lives in the next packet (every packet but a unit's last holds an
// even number of k-groups) -- are loaded before the DMMAs of the current one are issued, so the tensor pipe does not drain at
// packet boundaries.  The hot loops carry no predicates (a runtime predicate in front of mma.sync costs a WARPSYNC + NOP per
// instruction): the number of live m-tiles of a warp is a template parameter chosen once per unit.
// complex<double>: one complex DMMA step = 4 real DMMAs on the (re, im) parts.
#pragma once
#include "../capi/plan.h"
#include "ptx.cuh"
#include "scalar.cuh"
#include "stream_kernels.cuh"

namespace hmb200 {

constexpr int kMmConsumerWarps = 8;
constexpr int kMmThreads       = (kMmConsumerWarps + 1) * 32;
constexpr int kMmYTileRows     = 32;                      // the bin tile is staged in two halves of 32 rows
constexpr int kMmYTileBytes    = kMmYTileRows * 72 * 8;   // 32 rows x (64 + 8 pad) doubles = 32 x (32 + 4) complex

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

// the consumer's view of the stage ring
struct MmRing {
    uint32_t bar_base;
    uint8_t *stage_base;
    uint32_t stage_sz;
    int NS, slot;
    uint32_t par;
    __device__ __forceinline__ const uint8_t *wait() const { // current stage
        ptx::mbar_wait(bar_base + 8u * slot, par);
        return stage_base + (size_t)slot * stage_sz;
    }
    __device__ __forceinline__ const uint8_t *wait_next() const { // the stage after the current one, without leaving the current
        const int ns = slot + 1 == NS ? 0 : slot + 1;
        ptx::mbar_wait(bar_base + 8u * ns, ns == 0 ? par ^ 1u : par);
        return stage_base + (size_t)ns * stage_sz;
    }
    __device__ __forceinline__ void release(int lane) { // all shared loads of the current stage have been issued by this warp
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(bar_base + 8u * (NS + slot));
        if (++slot == NS) {
            slot = 0;
            par ^= 1u;
        }
    }
};

struct MmHeader {
    int n_kg, rows, mt_panel;
    uint32_t out0, flags, map_off16;
    __device__ __forceinline__ explicit MmHeader(const uint8_t *pk) {
        const uint4 h = *reinterpret_cast<const uint4 *>(pk);
        n_kg = (int)h.x; out0 = h.y; rows = (int)(h.z & 0xffffu); mt_panel = (int)(h.z >> 16); flags = h.w & 0xffffu; map_off16 = h.w >> 16;
    }
};

// One unit = one K loop over all its packets.  The warp owns m-tiles wm, wm + WM, ... (NMT live ones; with PRED the live counts
// nmt <= NMT, nnt <= NT are runtime values and everything is predicated: only column tails and -- in the square tilings --
// panels that end inside the warp's rows take that path).  TALL: a warp with many m-tiles and few n-tiles (wide pass of phase 1).
// Returns with the unit's LAST packet still held (the caller reads the row map from it, then releases it).
//   square tiling: A chunks ping-pong between two register sets; real: the B fragments of a k4-step are reloaded for the next
//                  k-group right after their DMMAs were issued; complex: B ping-pongs too
//   tall tiling  : B ping-pongs; an A chunk is reloaded for the next k-group right after the last DMMA that reads it; two sweeps
//                  over the m-tiles per k-group keep DMMAs on one accumulator 2 x NMT x NT ... instructions apart
template <typename T, int MT, int NT, int NMT, bool PRED, bool TALL>
__device__ __forceinline__ const uint8_t *mm_unit(MmAcc<T, MT, NT> &acc, MmRing &ring, const uint8_t *pk, MmHeader &h, int stage_bytes, int ld, int wm, int WM,
                                                  int c_base, int nnt, int nmt, int lane) {
    constexpr int KG = sizeof(T) == 8 ? 8 : 4;
    constexpr int KS = MmBFrag<T, NT>::KS;
    const int gid = lane >> 2, tig = lane & 3;
    const int a_mt = WM * 32, a_kg = h.mt_panel * 32;
    const size_t b_kg = (size_t)KG * ld, b_k4 = (size_t)4 * ld;
    const int a_off = wm * 32 + lane;
    const size_t b_off = (size_t)tig * ld + c_base + gid;
    const double2 *ap = reinterpret_cast<const double2 *>(pk + 16) + a_off;
    const T *bp       = reinterpret_cast<const T *>(pk + stage_bytes) + b_off;
    int n_kg = h.n_kg;
    uint32_t flags = h.flags;
    if (n_kg == 0) return pk; // a bin without coefficients: one empty packet

    auto load_a = [&](double2 (&a)[NMT], const double2 *q) {
#pragma unroll
        for (int i = 0; i < NMT; i++)
            if (!PRED || i < nmt) a[i] = q[i * a_mt];
    };
    auto load_b = [&](MmBFrag<T, NT> &b, const T *q) {
#pragma unroll
        for (int ks = 0; ks < KS; ks++)
#pragma unroll
            for (int j = 0; j < NT; j++)
                if (!PRED || j < nnt) b.v[ks][j] = q[(size_t)(4 * ks) * ld + 8 * j];
    };
    auto load_bh = [&](T (&b)[NT], const T *q) { // one k4-step (real)
#pragma unroll
        for (int j = 0; j < NT; j++)
            if (!PRED || j < nnt) b[j] = q[8 * j];
    };
    // the next packet of this unit: wait for it, leave the current one, re-base the fragment pointers
    auto next_packet = [&]() {
        const uint8_t *npk = ring.wait_next();
        ring.release(lane);
        pk = npk;
        const uint4 hr = *reinterpret_cast<const uint4 *>(pk);
        n_kg  = (int)hr.x;
        flags = hr.w & 0xffffu;
        ap    = reinterpret_cast<const double2 *>(pk + 16) + a_off;
        bp    = reinterpret_cast<const T *>(pk + stage_bytes) + b_off;
    };

    if constexpr (TALL) {
        double2 a[NMT];
        MmBFrag<T, NT> b0, b1;
        auto group = [&](const MmBFrag<T, NT> &b, bool reload, const double2 *anext) {
#pragma unroll
            for (int i = 0; i < NMT; i++)
#pragma unroll
                for (int j = 0; j < NT; j++) {
                    if (PRED && j >= nnt) continue;
                    if constexpr (sizeof(T) == 8) {
                        dmma884(acc.v[i][j][0][0], acc.v[i][j][0][1], a[i].x, b.v[0][j]);
                    } else {
                        dmma884(acc.v[i][j][0][0], acc.v[i][j][0][1], a[i].x, b.v[0][j].x);
                        dmma884(acc.v[i][j][1][0], acc.v[i][j][1][1], a[i].x, b.v[0][j].y);
                    }
                }
#pragma unroll
            for (int i = 0; i < NMT; i++) {
#pragma unroll
                for (int j = 0; j < NT; j++) {
                    if (PRED && j >= nnt) continue;
                    if constexpr (sizeof(T) == 8) {
                        dmma884(acc.v[i][j][0][0], acc.v[i][j][0][1], a[i].y, b.v[KS - 1][j]);
                    } else {
                        dmma884(acc.v[i][j][0][0], acc.v[i][j][0][1], -a[i].y, b.v[0][j].y);
                        dmma884(acc.v[i][j][1][0], acc.v[i][j][1][1], a[i].y, b.v[0][j].x);
                    }
                }
                if (reload) a[i] = anext[i * a_mt]; // the chunk of the next k-group, as soon as this one has been read
            }
        };
        load_a(a, ap);
        load_b(b0, bp);
        for (;;) {
            for (int kg = 0; kg < n_kg; kg += 2) {
                const bool has1 = kg + 1 < n_kg;
                if (has1) load_b(b1, bp + b_kg);
                group(b0, has1, ap + a_kg);
                if (has1) {
                    if (kg + 2 < n_kg) {
                        ap += 2 * a_kg;
                        bp += 2 * b_kg;
                        load_b(b0, bp);
                        group(b1, true, ap);
                    } else if (!(flags & kMmFlagLast)) { // the next k-group is the first of the next packet
                        next_packet();
                        load_b(b0, bp);
                        group(b1, true, ap);
                        kg = -2; // restart the k-group loop on the new packet
                    } else {
                        group(b1, false, ap);
                    }
                }
            }
            if (flags & kMmFlagLast) return pk;
            // (only reached when a non-last packet held an odd number of k-groups, which the planner never emits)
            next_packet();
            load_a(a, ap);
            load_b(b0, bp);
        }
    } else if constexpr (sizeof(T) == 8) {
        double2 a0[NMT], a1[NMT];
        double b0[NT], b1[NT];
        auto half = [&](const double2 (&a)[NMT], bool hi, const double (&b)[NT]) {
#pragma unroll
            for (int i = 0; i < NMT; i++) {
                if (PRED && i >= nmt) continue;
                const double av = hi ? a[i].y : a[i].x;
#pragma unroll
                for (int j = 0; j < NT; j++)
                    if (!PRED || j < nnt) dmma884(acc.v[i][j][0][0], acc.v[i][j][0][1], av, b[j]);
            }
        };
        load_a(a0, ap);
        load_bh(b0, bp);
        load_bh(b1, bp + b_k4);
        for (;;) {
            for (int kg = 0; kg < n_kg; kg += 2) {
                const bool has1 = kg + 1 < n_kg;
                if (has1) load_a(a1, ap + a_kg);
                half(a0, false, b0);
                if (has1) load_bh(b0, bp + b_kg);
                half(a0, true, b1);
                if (has1) {
                    load_bh(b1, bp + b_kg + b_k4);
                    const bool more = kg + 2 < n_kg, cross = !more && !(flags & kMmFlagLast);
                    if (more) {
                        ap += 2 * a_kg;
                        bp += 2 * b_kg;
                    } else if (cross) {
                        next_packet();
                        kg = -2;
                    }
                    if (more || cross) load_a(a0, ap);
                    half(a1, false, b0);
                    if (more || cross) load_bh(b0, bp);
                    half(a1, true, b1);
                    if (more || cross) load_bh(b1, bp + b_k4);
                }
            }
            if (flags & kMmFlagLast) return pk;
            next_packet();
            load_a(a0, ap);
            load_bh(b0, bp);
            load_bh(b1, bp + b_k4);
        }
    } else {
        double2 a0[NMT], a1[NMT];
        MmBFrag<T, NT> b0, b1;
        auto mma = [&](const double2 (&a)[NMT], const MmBFrag<T, NT> &b) {
#pragma unroll
            for (int i = 0; i < NMT; i++) {
                if (PRED && i >= nmt) continue;
#pragma unroll
                for (int j = 0; j < NT; j++)
                    if (!PRED || j < nnt) { // (ar + i ai)(br + i bi): re += ar br - ai bi ; im += ar bi + ai br
                        dmma884(acc.v[i][j][0][0], acc.v[i][j][0][1], a[i].x, b.v[0][j].x);
                        dmma884(acc.v[i][j][1][0], acc.v[i][j][1][1], a[i].x, b.v[0][j].y);
                    }
#pragma unroll
                for (int j = 0; j < NT; j++)
                    if (!PRED || j < nnt) {
                        dmma884(acc.v[i][j][0][0], acc.v[i][j][0][1], -a[i].y, b.v[0][j].y);
                        dmma884(acc.v[i][j][1][0], acc.v[i][j][1][1], a[i].y, b.v[0][j].x);
                    }
            }
        };
        load_a(a0, ap);
        load_b(b0, bp);
        for (;;) {
            for (int kg = 0; kg < n_kg; kg += 2) {
                const bool has1 = kg + 1 < n_kg;
                if (has1) {
                    load_a(a1, ap + a_kg);
                    load_b(b1, bp + b_kg);
                }
                mma(a0, b0);
                if (has1) {
                    const bool more = kg + 2 < n_kg, cross = !more && !(flags & kMmFlagLast);
                    if (more) {
                        ap += 2 * a_kg;
                        bp += 2 * b_kg;
                    } else if (cross) {
                        next_packet();
                        kg = -2;
                    }
                    if (more || cross) {
                        load_a(a0, ap);
                        load_b(b0, bp);
                    }
                    mma(a1, b1);
                }
            }
            if (flags & kMmFlagLast) return pk;
            next_packet();
            load_a(a0, ap);
            load_b(b0, bp);
        }
    }
}

// column tail of a pass inside a tall warp tile (fewer live n-tiles than NT): plain predicated loop over the unit's packets, no
// register double buffering (few registers; only the warps that straddle the last columns of a 33..63-column pass take it)
template <typename T, int MT, int NT>
__device__ __forceinline__ const uint8_t *mm_unit_plain(MmAcc<T, MT, NT> &acc, MmRing &ring, const uint8_t *pk, MmHeader &h, int stage_bytes, int ld, int wm, int WM,
                                                        int c_base, int nnt, int nmt, int lane) {
    constexpr int KG = sizeof(T) == 8 ? 8 : 4;
    constexpr int KS = MmBFrag<T, NT>::KS;
    const int gid = lane >> 2, tig = lane & 3;
    int n_kg = h.n_kg;
    uint32_t flags = h.flags;
    for (;;) {
        const double2 *ap = reinterpret_cast<const double2 *>(pk + 16) + wm * 32 + lane;
        const T *bp       = reinterpret_cast<const T *>(pk + stage_bytes) + (size_t)tig * ld + c_base + gid;
        for (int kg = 0; kg < n_kg; kg++, ap += h.mt_panel * 32, bp += (size_t)KG * ld) {
            MmBFrag<T, NT> b;
#pragma unroll
            for (int ks = 0; ks < KS; ks++)
#pragma unroll
                for (int j = 0; j < NT; j++)
                    if (j < nnt) b.v[ks][j] = bp[(size_t)(4 * ks) * ld + 8 * j];
#pragma unroll
            for (int i = 0; i < MT; i++) {
                if (i >= nmt) continue;
                const double2 a = ap[i * WM * 32];
#pragma unroll
                for (int j = 0; j < NT; j++) {
                    if (j >= nnt) continue;
                    if constexpr (sizeof(T) == 8) {
                        dmma884(acc.v[i][j][0][0], acc.v[i][j][0][1], a.x, b.v[0][j]);
                        dmma884(acc.v[i][j][0][0], acc.v[i][j][0][1], a.y, b.v[KS - 1][j]);
                    } else {
                        dmma884(acc.v[i][j][0][0], acc.v[i][j][0][1], a.x, b.v[0][j].x);
                        dmma884(acc.v[i][j][1][0], acc.v[i][j][1][1], a.x, b.v[0][j].y);
                        dmma884(acc.v[i][j][0][0], acc.v[i][j][0][1], -a.y, b.v[0][j].y);
                        dmma884(acc.v[i][j][1][0], acc.v[i][j][1][1], a.y, b.v[0][j].x);
                    }
                }
            }
        }
        if (flags & kMmFlagLast) return pk;
        ring.release(lane);
        pk = ring.wait();
        const uint4 hr = *reinterpret_cast<const uint4 *>(pk);
        n_kg  = (int)hr.x;
        flags = hr.w & 0xffffu;
    }
}

// a warp without live tiles in this unit (panel shorter than its rows, or a column tail) only walks the ring
__device__ __forceinline__ const uint8_t *mm_skip_unit(MmRing &ring, const uint8_t *pk, uint32_t flags, int lane) {
    while (!(flags & kMmFlagLast)) {
        ring.release(lane);
        pk    = ring.wait();
        flags = reinterpret_cast<const uint4 *>(pk)->w & 0xffffu;
    }
    return pk;
}

template <typename T, int MT, int NT, int NMT, bool PRED>
__device__ __forceinline__ const uint8_t *mm_tall_switch(MmAcc<T, MT, NT> &acc, MmRing &ring, const uint8_t *pk, MmHeader &h, int stage_bytes, int ld, int wm, int WM,
                                                         int c_base, int nnt, int nmt, int lane) {
    if (nmt == NMT) return mm_unit<T, MT, NT, NMT, PRED, true>(acc, ring, pk, h, stage_bytes, ld, wm, WM, c_base, nnt, NMT, lane);
    if constexpr (NMT > 1) return mm_tall_switch<T, MT, NT, NMT - 1, PRED>(acc, ring, pk, h, stage_bytes, ld, wm, WM, c_base, nnt, nmt, lane);
    else return pk;
}

template <typename T, int MT, int NT>
__device__ __forceinline__ const uint8_t *mm_unit_dispatch(MmAcc<T, MT, NT> &acc, MmRing &ring, const uint8_t *pk, MmHeader &h, int stage_bytes, int ld, int wm, int WM,
                                                           int c_base, int nnt, int nmt, int lane) {
    if (nmt <= 0 || nnt <= 0) return mm_skip_unit(ring, pk, h.flags, lane);
    if constexpr (MT >= 4) { // tall tiling: live m-tile count at compile time; a column tail predicates the n-tiles only
        if (nnt == NT) return mm_tall_switch<T, MT, NT, MT, false>(acc, ring, pk, h, stage_bytes, ld, wm, WM, c_base, nnt, nmt, lane);
        else return mm_unit_plain<T, MT, NT>(acc, ring, pk, h, stage_bytes, ld, wm, WM, c_base, nnt, nmt, lane);
    } else {
        if (nnt == NT && nmt == MT) return mm_unit<T, MT, NT, MT, false, false>(acc, ring, pk, h, stage_bytes, ld, wm, WM, c_base, nnt, nmt, lane);
        if constexpr (MT == 2) {
            if (nnt == NT && nmt == 1) return mm_unit<T, MT, NT, 1, false, false>(acc, ring, pk, h, stage_bytes, ld, wm, WM, c_base, nnt, nmt, lane);
        }
        return mm_unit<T, MT, NT, MT, true, false>(acc, ring, pk, h, stage_bytes, ld, wm, WM, c_base, nnt, nmt, lane);
    }
}

template <typename T>
__device__ __forceinline__ T mm_make(double re, double im);
template <>
__device__ __forceinline__ double mm_make<double>(double re, double) { return re; }
template <>
__device__ __forceinline__ double2 mm_make<double2>(double re, double im) { return make_double2(re, im); }

// WM x WN = 8 consumer warps, each MT x NT tiles; WM * MT m-tiles cover the panel: 16 (128 rows) in phase 1, 8 (64 rows) in phase 2.
// Wide passes: phase 2 runs 4 x 2 warps of 2 x 4 tiles, phase 1 runs 2 x 4 warps of 8 x 2 tiles (panels there have any number
// of m-tiles: two warp rows with interleaved m-tiles balance to within one tile).  Narrow passes (<= 32 columns): 8 x 1 warps.
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
        int slot = 0;
        uint32_t par = 0;
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
                for (uint32_t j = 0; j < cnt; j++) {
                    const uint32_t off_lo = __shfl_sync(0xffffffffu, me0.x, j), off_hi = __shfl_sync(0xffffffffu, me0.y, j);
                    const uint32_t bytes = __shfl_sync(0xffffffffu, me0.z, j), z_first = __shfl_sync(0xffffffffu, me0.w, j);
                    const uint32_t n_z = __shfl_sync(0xffffffffu, me1.x, j), z_bytes = __shfl_sync(0xffffffffu, me1.y, j);
                    const uint64_t gmem_off = (uint64_t)off_lo | ((uint64_t)off_hi << 32);
                    uint2 zc = make_uint2(0, 0);
                    if ((uint32_t)lane < n_z) zc = __ldg(reinterpret_cast<const uint2 *>(p.zcopies + z_first) + lane); // in flight during the wait
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
                    if (++slot == NS) {
                        slot = 0;
                        par ^= 1u;
                    }
                }
            }
            u = __shfl_sync(0xffffffffu, u_next, 0);
        }
        { // terminate packet: a header written through the generic proxy, released by a plain arrive
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
    // warp -> (row, column) of the warp grid.  Warps w and w + 4 share a scheduler (and its DMMA pipe); with two warp rows
    // (tall tiling: a panel of an odd number of m-tiles gives row 0 one tile more than row 1) they are put on DIFFERENT rows, so
    // that every pipe sees the same number of DMMAs per k-group
    const int wm = WM == 2 ? ((warp + (warp >> 2)) & 1) : warp % WM;
    const int wn = WM == 2 ? (warp >> 1) : warp / WM;
    const int c_base = wn * NT * 8;
    const int ncols8 = (p.ncols + 7) & ~7;
    const int nnt    = max(0, min(NT, (ncols8 - c_base) >> 3)); // n-tiles of this warp that hold right-hand sides
    MmAcc<T, MT, NT> acc;
    acc.clear();
    MmRing ring{bar_base, stage_base, stage_sz, NS, 0, 0u};
    T *ws = (T *)p.ws;
    for (;;) {
        const uint8_t *pk = ring.wait();
        MmHeader h(pk);
        if (h.flags & kMmFlagTerminate) break;
        const int nmt = wm < h.mt_panel ? min(MT, (h.mt_panel - wm + WM - 1) / WM) : 0; // m-tiles wm, wm + WM, ... below mt_panel
        const bool live = nmt > 0 && nnt > 0;
        const uint8_t *lpk = mm_unit_dispatch<T, MT, NT>(acc, ring, pk, h, p.stage_bytes, ld, wm, WM, c_base, nnt, nmt, lane);

        if (PHASE == 1) {
            if (live) { // the unit's rows of T (or partial rows): destination row per panel row from the last packet's row map
                const uint32_t *map = reinterpret_cast<const uint32_t *>(lpk + 16 * (size_t)(reinterpret_cast<const uint4 *>(lpk)->w >> 16));
#pragma unroll
                for (int i = 0; i < MT; i++) {
                    const int r = 8 * (wm + WM * i) + gid;
                    if (i < nmt && r < h.rows) {
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
            ring.release(lane);
        } else {
            ring.release(lane);
            // stage the bin's tile in shared memory (two halves of 32 rows), then all consumer threads store full rows: the local
            // Y (with beta) and the same bits into every peer's replicated block (fused all-gather)
            constexpr int ldt = kMmYTileBytes / kMmYTileRows / (int)sizeof(T);
            T *yt = reinterpret_cast<T *>(ytile_raw);
            const T alpha = mm_make<T>(p.alpha_re, p.alpha_im), beta = mm_make<T>(p.beta_re, p.beta_im);
            const bool beta_zero = Scalar<T>::is_zero(beta);
            for (int hf = 0; hf * kMmYTileRows < h.rows; hf++) {
#pragma unroll
                for (int i = 0; i < MT; i++) {
                    const int mt = wm + WM * i;
                    if (live && i < nmt && (mt >> 2) == hf) {
                        const int r = 8 * (mt & 3) + gid;
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
                // thread -> (column, every 4th row): no divisions; a value is scaled once and stored to the local Y and to every peer
                const int rows_h = min(kMmYTileRows, h.rows - hf * kMmYTileRows);
                const int cs_log = p.ncols <= 8 ? 3 : (p.ncols <= 16 ? 4 : (p.ncols <= 32 ? 5 : 6)); // column slots: 8 / 16 / 32 / 64
                const int c = (int)threadIdx.x & ((1 << cs_log) - 1), row_step = (NW * 32) >> cs_log;
                if (c < p.ncols) {
                    for (int r = (int)threadIdx.x >> cs_log; r < rows_h; r += row_step) {
                        T v = Scalar<T>::mul(alpha, yt[r * ldt + c]);
                        const size_t o = (size_t)(h.out0 + hf * kMmYTileRows + r) * p.ldy + c;
                        T *y = (T *)p.y;
                        if (!beta_zero) v = Scalar<T>::add(v, Scalar<T>::mul(beta, y[o])); // beta == 0 must not propagate NaNs of y (BLAS)
                        y[o] = v;
                        for (int q = 0; q < p.n_peers; q++) reinterpret_cast<T *>(p.peer_y[q])[o] = v; // beta = 0 on the fused path
                    }
                }
                ptx::named_bar_sync(1, NW * 32); // the tile is rewritten by the next half / the next bin
            }
            acc.clear();
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
// the same for a row-major source without permutation (the operator interface's own layout): one warp per row, 16-byte chunks,
// no index division -- this copy is pure overhead of the multi-RHS product (the B fragments need the padded row stride) and does
// not shrink with the number of GPUs, so it has to run at memory speed
template <typename T>
__global__ void __launch_bounds__(256) mm_load_rows_fast_kernel(const T *__restrict__ src, long long lds, int n, int ncols, T *__restrict__ ws, int ld, int vec_ok,
                                                                const int32_t *__restrict__ row_perm) {
    constexpr int V = 16 / (int)sizeof(T); // elements per 16-byte chunk
    const int cpr   = (ncols + V - 1) / V;
    for (size_t row = (size_t)blockIdx.x * blockDim.y + threadIdx.y; row < (size_t)n; row += (size_t)gridDim.x * blockDim.y) {
        const T *s = src + (row_perm ? (size_t)row_perm[row] : row) * lds; // optional row gather (user -> cluster numbering)
        T *d       = ws + row * ld;
        for (int c = threadIdx.x; c < cpr; c += blockDim.x) {
            const int col = c * V;
            if (vec_ok && col + V <= ncols) {
                *reinterpret_cast<double2 *>(d + col) = *reinterpret_cast<const double2 *>(s + col);
            } else {
#pragma unroll
                for (int e = 0; e < V; e++)
                    if (col + e < ncols) d[col + e] = s[col + e];
            }
        }
    }
}
// rows [r0, r0 + cnt) of the USER-numbered result as a column-major chunk (cnt x ncols, leading dimension cnt), gathered from the
// cluster-numbered row-major block through the inverse target permutation: the rank's share of the replicated result of the
// distributed matrix product, ready for one strided D2H copy
template <typename T>
__global__ void mm_gather_user_chunk_kernel(const T *__restrict__ yc, int ncols, const int32_t *__restrict__ inv_perm, long long r0, int cnt, T *__restrict__ dst) {
    const size_t total = (size_t)cnt * ncols;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const size_t i = t / ncols;
        const int j    = (int)(t - i * ncols);
        dst[i + (size_t)j * cnt] = yc[(size_t)inv_perm[r0 + i] * ncols + j];
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
