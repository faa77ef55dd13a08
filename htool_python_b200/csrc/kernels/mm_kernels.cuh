// mm_kernels.cuh -- multi-RHS H-matrix product Y += A X on the FP64 tensor cores (sm_100a: mma.sync.m8n8k4.f64, SASS DMMA).
//
// Same two phases and the same packed streams as the single-RHS kernels (stream_kernels.cuh), but every leaf product is
// now a real contraction (reference: the two gemms of hmatrix/lrmat/linalg/add_lrmat_matrix_product_row_major.hpp:11-27
// and the dense gemm of matrix/linalg/add_matrix_matrix_product_row_major.hpp:23-46, driven by the leaf loop of
// hmatrix/linalg/add_hmatrix_matrix_product_row_major.hpp:57-178):
//
//   PHASE 1:  T_j (r x mu) = V_j X[cols_j]      V tiles grouped by SOURCE bin; the bin's rows of X are resident in shared
//             memory (double buffered) while the tiles stream; outputs go to the workspace (final T rows, or partial
//             rows summed by reduce_rows_kernel when a leaf spans several source bins)
//   PHASE 2:  Y[bin] = beta Y[bin] + alpha (sum_j U_j T_j + sum_d D_d X[cols_d])   the packets of the single-RHS stream2;
//             the operand ROWS (T_j / X rows, mu wide) are fetched per packet by bulk copies; each consumer warp owns a
//             16 x (4 n-tiles) block of the bin's output in DMMA accumulator registers for the whole bin: no atomics,
//             no cross-warp reduction.
//
// Row-major multi-RHS blocks (row i, rhs j at i*ld + j) exactly as the reference's *_row_major interfaces, so the
// transposes of add_hmatrix_matrix_product.hpp:29-31,48 never happen.  The workspace keeps rows at a padded stride
// (ld = mu rounded up to 8, plus 4 or 12) so that B-fragment loads are bank-conflict free.
// complex<double>: one complex DMMA step = 4 real DMMAs on the (re, im) parts of the interleaved fragments.
#pragma once
#include "../capi/plan.h"
#include "ptx.cuh"
#include "scalar.cuh"

namespace hmb200 {

constexpr int kMmConsumerWarps = 8; // 4 row groups (16 rows) x 2 column halves (4 / 2 n-tiles of 8)
constexpr int kMmSlack         = 256;

struct MmParams {
    const uint8_t *stream;
    const StageEntry *stages;
    const ZCopy *zcopies;
    const WorkUnit *units;
    uint32_t n_units;
    uint32_t n_zcopies;
    uint32_t *counter;
    void *ws;   // T*: workspace rows [x | t | partials], row stride ld
    int ld;     // workspace row stride in elements
    int ncols;  // right-hand sides of this pass (<= 64 real, <= 32 complex)
    void *y;    // phase 2: output block, row-major, stride ldy, already offset to the pass's first column
    long long ldy;
    double alpha_re, alpha_im, beta_re, beta_im;
    int stage_bytes, z_rows, n_stages, x_rows; // z_rows: operand rows per stage (phase 2); x_rows: resident rows (phase 1)
};

template <typename T>
__host__ __device__ constexpr int mm_ntiles_per_warp() { return sizeof(T) == 8 ? 4 : 2; } // columns per warp = 32 / 16
template <typename T>
__host__ __device__ constexpr int mm_pass_cols() { return sizeof(T) == 8 ? 64 : 32; }

__host__ __device__ inline size_t mm_smem_bytes(int phase, int elem_bytes, int stage_bytes, int z_rows, int n_stages, int x_rows, int ld) {
    size_t z = phase == 2 ? (size_t)z_rows * ld * elem_bytes : 0;
    size_t x = phase == 1 ? (size_t)2 * x_rows * ld * elem_bytes : 0;
    return 256 + x + (size_t)n_stages * (stage_bytes + z) + kMmSlack;
}

__device__ __forceinline__ void dmma884v(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// accumulators of one warp: MT m-tiles x NT n-tiles (x re/im)
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

// acc(16 rows x NT*8 cols) += A[rows r_base.., k] * B[k, cols c_base..] over k in [0, nk)
//   A: column-major tile (ld = lda elements), row index = r_base + 8*mt + gid, valid rows [0, nr)
//   B: row-major operand rows (ld = ldb), row k, column c_base + 8*nt + gid
template <typename T, int NT>
__device__ __forceinline__ void mm_tile(MmAcc<T, 2, NT> &acc, const T *A, int lda, int nr, int r_base, const T *B, int ldb, int c_base, int nk, int lane) {
    using S = Scalar<T>;
    const int gid = lane >> 2, tig = lane & 3;
    const int r0 = r_base + gid, r1 = r_base + 8 + gid;
    const bool v0 = (unsigned)r0 < (unsigned)nr, v1 = (unsigned)r1 < (unsigned)nr;
    const T *a0p = A + (v0 ? r0 : 0) + (size_t)tig * lda;
    const T *a1p = A + (v1 ? r1 : 0) + (size_t)tig * lda;
    const T *bp  = B + (size_t)tig * ldb + c_base + gid;
    // fragments of the current k-step; the next step's are loaded before the DMMAs of this one are issued
    // (the K tail is zero-filled: the bytes next to a tile are other tiles / headers)
    T a0, a1, b[NT];
    {
        const bool kv = tig < nk;
        a0 = (kv && v0) ? a0p[0] : S::zero();
        a1 = (kv && v1) ? a1p[0] : S::zero();
#pragma unroll
        for (int j = 0; j < NT; j++) b[j] = kv ? bp[8 * j] : S::zero();
    }
    for (int k0 = 0; k0 < nk; k0 += 4) {
        a0p += 4 * (size_t)lda;
        a1p += 4 * (size_t)lda;
        bp += 4 * (size_t)ldb;
        const bool kv = k0 + 4 + tig < nk;
        const T na0 = (kv && v0) ? a0p[0] : S::zero();
        const T na1 = (kv && v1) ? a1p[0] : S::zero();
        T nb[NT];
#pragma unroll
        for (int j = 0; j < NT; j++) nb[j] = kv ? bp[8 * j] : S::zero();
        if constexpr (sizeof(T) == 8) {
#pragma unroll
            for (int j = 0; j < NT; j++) {
                dmma884(acc.v[0][j][0][0], acc.v[0][j][0][1], a0, b[j]);
                dmma884(acc.v[1][j][0][0], acc.v[1][j][0][1], a1, b[j]);
            }
        } else {
#pragma unroll
            for (int j = 0; j < NT; j++) {
                // (ar + i ai)(br + i bi): re += ar br - ai bi ; im += ar bi + ai br
                dmma884(acc.v[0][j][0][0], acc.v[0][j][0][1], a0.x, b[j].x);
                dmma884(acc.v[0][j][1][0], acc.v[0][j][1][1], a0.x, b[j].y);
                dmma884(acc.v[1][j][0][0], acc.v[1][j][0][1], a1.x, b[j].x);
                dmma884(acc.v[1][j][1][0], acc.v[1][j][1][1], a1.x, b[j].y);
                dmma884(acc.v[0][j][0][0], acc.v[0][j][0][1], -a0.y, b[j].y);
                dmma884(acc.v[0][j][1][0], acc.v[0][j][1][1], a0.y, b[j].x);
                dmma884(acc.v[1][j][0][0], acc.v[1][j][0][1], -a1.y, b[j].y);
                dmma884(acc.v[1][j][1][0], acc.v[1][j][1][1], a1.y, b[j].x);
            }
        }
        a0 = na0;
        a1 = na1;
#pragma unroll
        for (int j = 0; j < NT; j++) b[j] = nb[j];
    }
}

// the same contraction without the register double-buffering (fewer registers, no prologue): better for the short-K
// tiles of phase 2 (K = rank)
template <typename T, int NT>
__device__ __forceinline__ void mm_tile_plain(MmAcc<T, 2, NT> &acc, const T *A, int lda, int nr, int r_base, const T *B, int ldb, int c_base, int nk, int lane) {
    using S = Scalar<T>;
    const int gid = lane >> 2, tig = lane & 3;
    const int r0 = r_base + gid, r1 = r_base + 8 + gid;
    const bool v0 = (unsigned)r0 < (unsigned)nr, v1 = (unsigned)r1 < (unsigned)nr;
    const T *a0p = A + (v0 ? r0 : 0) + (size_t)tig * lda;
    const T *a1p = A + (v1 ? r1 : 0) + (size_t)tig * lda;
    const T *bp  = B + (size_t)tig * ldb + c_base + gid;
    for (int k0 = 0; k0 < nk; k0 += 4) {
        const bool kv = k0 + tig < nk;
        const T a0 = (kv && v0) ? a0p[0] : S::zero(), a1 = (kv && v1) ? a1p[0] : S::zero();
        T b[NT];
#pragma unroll
        for (int j = 0; j < NT; j++) b[j] = kv ? bp[8 * j] : S::zero();
        if constexpr (sizeof(T) == 8) {
#pragma unroll
            for (int j = 0; j < NT; j++) {
                dmma884v(acc.v[0][j][0][0], acc.v[0][j][0][1], a0, b[j]);
                dmma884v(acc.v[1][j][0][0], acc.v[1][j][0][1], a1, b[j]);
            }
        } else {
#pragma unroll
            for (int j = 0; j < NT; j++) {
                dmma884v(acc.v[0][j][0][0], acc.v[0][j][0][1], a0.x, b[j].x);
                dmma884v(acc.v[0][j][1][0], acc.v[0][j][1][1], a0.x, b[j].y);
                dmma884v(acc.v[1][j][0][0], acc.v[1][j][0][1], a1.x, b[j].x);
                dmma884v(acc.v[1][j][1][0], acc.v[1][j][1][1], a1.x, b[j].y);
                dmma884v(acc.v[0][j][0][0], acc.v[0][j][0][1], -a0.y, b[j].y);
                dmma884v(acc.v[0][j][1][0], acc.v[0][j][1][1], a0.y, b[j].x);
                dmma884v(acc.v[1][j][0][0], acc.v[1][j][0][1], -a1.y, b[j].y);
                dmma884v(acc.v[1][j][1][0], acc.v[1][j][1][1], a1.y, b[j].x);
            }
        }
        a0p += 4 * (size_t)lda;
        a1p += 4 * (size_t)lda;
        bp += 4 * (size_t)ldb;
    }
}

template <typename T>
__device__ __forceinline__ T mm_make(double re, double im);
template <>
__device__ __forceinline__ double mm_make<double>(double re, double) { return re; }
template <>
__device__ __forceinline__ double2 mm_make<double2>(double re, double im) { return make_double2(re, im); }

template <typename T, int PHASE, bool PIPE>
__global__ void __launch_bounds__((kMmConsumerWarps + 1) * 32, 1) hmatrix_mm_kernel(const MmParams p) {
    constexpr int NW = kMmConsumerWarps;
    constexpr int NT = mm_ntiles_per_warp<T>();
    extern __shared__ __align__(128) uint8_t smem[];
    const int NS   = p.n_stages;
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int ld   = p.ld;

    const uint32_t bar_base = ptx::smem_u32(smem);
    const size_t x_bytes    = PHASE == 1 ? (size_t)p.x_rows * ld * sizeof(T) : 0; // one resident X buffer
    const size_t z_bytes_st = PHASE == 2 ? (size_t)p.z_rows * ld * sizeof(T) : 0; // operand rows of one stage
    uint8_t *x_base         = smem + 256;
    uint8_t *stage_base     = x_base + 2 * x_bytes;
    const uint32_t stage_sz = (uint32_t)(p.stage_bytes + z_bytes_st);
    auto full_bar   = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar  = [&](int s) { return bar_base + 8u * (NS + s); };
    auto xfull_bar  = [&](int b) { return bar_base + 8u * (2 * NS + b); };
    auto xempty_bar = [&](int b) { return bar_base + 8u * (2 * NS + 2 + b); };

    if (threadIdx.x == 0) {
        for (int s = 0; s < NS; s++) {
            ptx::mbar_init(full_bar(s), 1);
            ptx::mbar_init(empty_bar(s), NW);
        }
        for (int b = 0; b < 2; b++) {
            ptx::mbar_init(xfull_bar(b), 1);
            ptx::mbar_init(xempty_bar(b), NW);
        }
        ptx::fence_mbar_init();
    }
    __syncthreads();

    if (warp == NW) {
        // ============================== producer warp ==============================
        const uint64_t pol_stream = ptx::policy_evict_first();
        const uint64_t pol_vec    = ptx::policy_evict_last();
        const T *ws               = (const T *)p.ws;
        uint32_t it = 0, ul = 0;
        uint32_t u = 0;
        if (lane == 0) u = atomicAdd(p.counter, 1u);
        u = __shfl_sync(0xffffffffu, u, 0);
        while (u < p.n_units) {
            uint32_t u_next = 0;
            if (lane == 0) u_next = atomicAdd(p.counter, 1u);
            const uint2 wraw = __ldg(reinterpret_cast<const uint2 *>(p.units + u));
            for (uint32_t s = wraw.x; s < wraw.x + wraw.y; s++, it++) {
                const uint4 e0 = __ldg(reinterpret_cast<const uint4 *>(p.stages + s));
                const uint4 e1 = __ldg(reinterpret_cast<const uint4 *>(p.stages + s) + 1);
                const uint64_t gmem_off = (uint64_t)e0.x | ((uint64_t)e0.y << 32);
                const uint32_t bytes = e0.z, z_first = e0.w, n_z = e1.x, z_bytes = e1.y;
                if (PHASE == 1 && s == wraw.x) { // the unit's resident slice of X: rows [e1.z, e1.z + e1.w)
                    const int buf = ul & 1;
                    ptx::mbar_wait(xempty_bar(buf), ((ul >> 1) & 1u) ^ 1u);
                    if (lane == 0) {
                        const uint32_t xb = e1.w * (uint32_t)ld * (uint32_t)sizeof(T);
                        ptx::mbar_arrive_expect_tx(xfull_bar(buf), xb);
                        ptx::bulk_g2s(ptx::smem_u32(x_base + (size_t)buf * x_bytes), ws + (size_t)e1.z * ld, xb, xfull_bar(buf), pol_vec);
                    }
                    ul++;
                }
                const int slot       = it % NS;
                const uint32_t par   = (it / NS) & 1u;
                const uint32_t sbase = ptx::smem_u32(stage_base + (size_t)slot * stage_sz);
                ptx::mbar_wait(empty_bar(slot), par ^ 1u);
                if (lane == 0) {
                    ptx::mbar_arrive_expect_tx(full_bar(slot), bytes + (PHASE == 2 ? z_bytes * (uint32_t)ld : 0u));
                    ptx::bulk_g2s(sbase, p.stream + gmem_off, bytes, full_bar(slot), pol_stream);
                }
                __syncwarp();
                if (PHASE == 2) {
                    for (uint32_t i = lane; i < n_z; i += 32) { // operand rows: each matvec operand element is a row of ld here
                        const uint2 zc    = __ldg(reinterpret_cast<const uint2 *>(p.zcopies + z_first) + i);
                        const uint32_t zb = (zc.y & 0xffffu) * (uint32_t)ld, zdst = (zc.y >> 16) * (uint32_t)ld;
                        ptx::bulk_g2s(sbase + p.stage_bytes + zdst, ws + (size_t)zc.x * ld, zb, full_bar(slot), pol_vec);
                    }
                }
            }
            u = __shfl_sync(0xffffffffu, u_next, 0);
        }
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

    // ============================== consumer warps ==============================
    const int gid = lane >> 2, tig = lane & 3;
    const int mg = warp & 3, nh = warp >> 2;
    const int c_base  = nh * NT * 8;
    const int ncols8  = (p.ncols + 7) & ~7;
    MmAcc<T, 2, NT> acc;
    acc.clear();
    T *ws = (T *)p.ws;
    uint32_t ul = 0;
    const T *xres = nullptr;
    for (uint32_t it = 0;; it++) {
        const int slot     = it % NS;
        const uint32_t par = (it / NS) & 1u;
        ptx::mbar_wait(full_bar(slot), par);
        const uint8_t *pk       = stage_base + (size_t)slot * stage_sz;
        const uint4 hraw        = *reinterpret_cast<const uint4 *>(pk);
        const uint32_t n_pieces = hraw.x, row0 = hraw.z;
        const uint32_t bin_rows = hraw.w & 0xffffu, flags = hraw.w >> 16;
        if (flags & kFlagTerminate) break;

        if (PHASE == 2) {
            const T *zarea = reinterpret_cast<const T *>(pk + p.stage_bytes);
            if (16 * mg < (int)bin_rows && c_base < ncols8) {
                for (uint32_t i = 0; i < n_pieces; i++) {
                    const uint4 d = *reinterpret_cast<const uint4 *>(pk + 16 + 16 * i);
                    const T *A    = reinterpret_cast<const T *>(pk + d.x);
                    const int nc = (int)(d.z >> 16), nr = (int)(d.w & 0xffu), ro = (int)((d.w >> 8) & 0xffu);
                    if (PIPE) mm_tile<T, NT>(acc, A, nr, nr, 16 * mg - ro, zarea + (size_t)(d.z & 0xffffu) * ld, ld, c_base, nc, lane);
                    else mm_tile_plain<T, NT>(acc, A, nr, nr, 16 * mg - ro, zarea + (size_t)(d.z & 0xffffu) * ld, ld, c_base, nc, lane);
                }
            }
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(empty_bar(slot));
            if (flags & kFlagLastOfBin) {
                const T alpha = mm_make<T>(p.alpha_re, p.alpha_im), beta = mm_make<T>(p.beta_re, p.beta_im);
                const bool beta_zero = Scalar<T>::is_zero(beta);
                T *y = (T *)p.y;
#pragma unroll
                for (int mt = 0; mt < 2; mt++) {
                    const int br = 16 * mg + 8 * mt + gid;
                    if (br < (int)bin_rows) {
                        T *yr = y + (size_t)(row0 + br) * p.ldy;
#pragma unroll
                        for (int j = 0; j < NT; j++)
#pragma unroll
                            for (int e = 0; e < 2; e++) {
                                const int col = c_base + 8 * j + 2 * tig + e;
                                if (col < p.ncols) {
                                    T s;
                                    if constexpr (sizeof(T) == 8) s = acc.v[mt][j][0][e];
                                    else s = make_double2(acc.v[mt][j][0][e], acc.v[mt][j][1][e]);
                                    T r = Scalar<T>::mul(alpha, s);
                                    if (!beta_zero) r = Scalar<T>::add(r, Scalar<T>::mul(beta, yr[col]));
                                    yr[col] = r;
                                }
                            }
                    }
                }
                acc.clear();
            }
        } else {
            if (flags & kFlagFirstOfBin) {
                const int buf = ul & 1;
                ptx::mbar_wait(xfull_bar(buf), (ul >> 1) & 1u);
                xres = reinterpret_cast<const T *>(x_base + (size_t)buf * x_bytes);
            }
            // work items = (tile, 16-row pair, column quarter): equal-sized, dealt round-robin to the consumer warps
            constexpr int NQ = NT / 2; // n-tiles per item (2 real / 1 complex): a quarter of the pass's columns
            uint32_t item = 0;
            for (uint32_t i = 0; i < n_pieces; i++) {
                const uint4 d = *reinterpret_cast<const uint4 *>(pk + 16 + 16 * i);
                const int nq = (int)(d.z >> 16), lr = (int)(d.w & 0xffu);
                const int n_mp = (lr + 15) >> 4;
                for (int q = 0; q < 4 * n_mp; q++, item++) {
                    if ((int)(item % NW) != warp) continue;
                    const int mp = q >> 2, cb = (q & 3) * NQ * 8;
                    if (cb >= ncols8) continue;
                    const T *A = reinterpret_cast<const T *>(pk + d.x);
                    MmAcc<T, 2, NQ> a1;
                    a1.clear();
                    if (PIPE) mm_tile<T, NQ>(a1, A, lr, lr, 16 * mp, xres + (size_t)(d.z & 0xffffu) * ld, ld, cb, nq, lane);
                    else mm_tile_plain<T, NQ>(a1, A, lr, lr, 16 * mp, xres + (size_t)(d.z & 0xffffu) * ld, ld, cb, nq, lane);
#pragma unroll
                    for (int mt = 0; mt < 2; mt++) {
                        const int r = 16 * mp + 8 * mt + gid;
                        if (r < lr) {
                            T *orow = ws + (size_t)(d.y + r) * ld;
#pragma unroll
                            for (int j = 0; j < NQ; j++) {
                                const int col = cb + 8 * j + 2 * tig;
                                if (col < ncols8) {
                                    if constexpr (sizeof(T) == 8) {
                                        *reinterpret_cast<double2 *>(orow + col) = make_double2(a1.v[mt][j][0][0], a1.v[mt][j][0][1]);
                                    } else {
                                        orow[col]     = make_double2(a1.v[mt][j][0][0], a1.v[mt][j][1][0]);
                                        orow[col + 1] = make_double2(a1.v[mt][j][0][1], a1.v[mt][j][1][1]);
                                    }
                                }
                            }
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) {
                ptx::mbar_arrive(empty_bar(slot));
                if (flags & kFlagLastOfBin) ptx::mbar_arrive(xempty_bar(ul & 1));
            }
            if (flags & kFlagLastOfBin) ul++;
        }
    }
}

// final T rows = sum of the output rows phase 1 wrote for the parts (source bins) of each leaf, in a fixed order
// (also the plain copy into the t slot for a leaf with one part); one warp per segment, lanes over (row, column pairs)
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

// row-major (n x mu, stride lds) <-> workspace rows (stride ld): X load of one pass, with optional cluster permutation of
// a COLUMN-major user block (fuses user_to_cluster + the transposes of add_hmatrix_matrix_product.hpp:29-31,190-195)
template <typename T>
__global__ void mm_load_rows_kernel(const T *__restrict__ src, long long lds, int n, int ncols, T *__restrict__ ws, int ld, const int32_t *__restrict__ perm, long long col_stride) {
    const size_t total = (size_t)n * ncols;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const size_t i = t / ncols;
        const int j    = (int)(t - i * ncols);
        ws[i * ld + j] = perm ? src[(size_t)perm[i] + (size_t)j * col_stride] : src[i * lds + j];
    }
}
// cluster-numbered row-major result rows -> column-major user block (cluster_to_user per column, :202-204)
template <typename T>
__global__ void mm_store_user_kernel(const T *__restrict__ yc, int n, int ncols, T *__restrict__ dst, const int32_t *__restrict__ perm, long long col_stride) {
    const size_t total = (size_t)n * ncols;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const size_t i = t / ncols;
        const int j    = (int)(t - i * ncols);
        dst[(size_t)perm[i] + (size_t)j * col_stride] = yc[i * ncols + j];
    }
}

} // namespace hmb200
