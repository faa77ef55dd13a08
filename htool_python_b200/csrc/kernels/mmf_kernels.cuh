// mmf_kernels.cuh -- EXPERIMENTAL (opt-in: Options.mm_variant bit 2; NOT yet run on a GPU): the multi-RHS product with the
// data flow of mm_kernels.cuh (same packed streams, tables, workspace rows, producer warp) but a plain-DFMA consumer.
//
// Why (profiles/r01_block_gmres_profile.log): a pass of the DMMA kernels costs ~10 ms at N=1M whatever its
// width, because each of the 8 consumer warps walks EVERY tile of its bin serially (one short dependent LDS->DMMA chain per
// tile).  On B200 the vector FP64 rate is 90 % of the DMMA rate (33 vs 37 TFLOP/s, profiles/r01_dmma_rate.log), so the mapping
// of the single-RHS stream kernel can be kept instead: the tiles of a packet are dealt round-robin to the consumer warps, lanes
// own output rows, every lane keeps NCOL accumulators per row (NCOL independent FMA chains), and the warps' partial sums of a
// bin are added through shared memory in a fixed order.  Per coefficient: 1 shared load + NCOL/2 broadcast 128-bit operand
// loads + NCOL DFMA; a pass is bounded by max(bytes / HBM, 2 W NCOL / DFMA rate).
//
//   PHASE 1:  T_j (r x nc) = V_j X[cols_j]     lane = g*rows + l owns row l for the columns k = g, g+G, ... of the tile
//                                              (G = 32 / rows groups), groups added by shuffles      [stream_kernels.cuh phase 1]
//   PHASE 2:  Y[bin] = beta Y[bin] + alpha (sum_j U_j T_j + sum_d D_d X[cols_d])   lane owns bin rows lane and lane + 32
//                                                                                                   [stream_kernels.cuh phase 2]
// Reference loops: add_hmatrix_matrix_product_row_major.hpp:57-178, add_lrmat_matrix_product_row_major.hpp:11-27.
#pragma once
#include "mm_kernels.cuh"

namespace hmb200 {

template <typename T>
__host__ __device__ constexpr int mmf_chunk_cols() { return sizeof(T) == 8 ? 8 : 4; } // columns per cross-warp reduction round
template <typename T>
__host__ __device__ constexpr int mmf_pass_cols() { return sizeof(T) == 8 ? 32 : 16; } // widest pass (accumulators in registers)

// shared memory of the cross-warp reduction: [warp][bin row][chunk column], row stride chunk+1 (bank-conflict free)
__host__ __device__ inline size_t mmf_red_bytes(int elem_bytes) {
    const int cc = elem_bytes == 8 ? 8 : 4;
    return (size_t)kMmConsumerWarps * kMaxBinRows * (cc + 1) * elem_bytes;
}
__host__ __device__ inline size_t mmf_smem_bytes(int phase, int elem_bytes, int stage_bytes, int z_rows, int n_stages, int x_rows, int ld) {
    return mm_smem_bytes(phase, elem_bytes, stage_bytes, z_rows, n_stages, x_rows, ld) + (phase == 2 ? mmf_red_bytes(elem_bytes) : 0);
}

// acc[j] += a * row[j], j < NCOL: the operand row is read with 16-byte shared loads (all lanes of a group read the same row)
template <typename T, int NCOL>
__device__ __forceinline__ void mmf_axpy_row(T (&acc)[NCOL], const T a, const T *row) {
    if constexpr (sizeof(T) == 8) {
#pragma unroll
        for (int j = 0; j < NCOL; j += 2) {
            const double2 z = *reinterpret_cast<const double2 *>(row + j);
            acc[j]     = ::fma(a, z.x, acc[j]);
            acc[j + 1] = ::fma(a, z.y, acc[j + 1]);
        }
    } else {
#pragma unroll
        for (int j = 0; j < NCOL; j++) Scalar<T>::fma(acc[j], a, row[j]);
    }
}

template <typename T, int PHASE, int NCOL>
__global__ void __launch_bounds__((kMmConsumerWarps + 1) * 32, 1) hmatrix_mmf_kernel(const MmParams p) {
    using S = Scalar<T>;
    constexpr int NW = kMmConsumerWarps;
    constexpr int CC = mmf_chunk_cols<T>();
    extern __shared__ __align__(128) uint8_t smem[];
    const int NS   = p.n_stages;
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int ld   = p.ld;

    // ---- layout and barriers: identical to hmatrix_mm_kernel (the reduction buffer is appended after the stages) ----
    const uint32_t bar_base = ptx::smem_u32(smem);
    const size_t x_bytes    = PHASE == 1 ? (size_t)p.x_rows * ld * sizeof(T) : 0;
    const size_t z_bytes_st = PHASE == 2 ? (size_t)p.z_rows * ld * sizeof(T) : 0;
    uint8_t *x_base         = smem + 256;
    uint8_t *stage_base     = x_base + 2 * x_bytes;
    const uint32_t stage_sz = (uint32_t)(p.stage_bytes + z_bytes_st);
    T *red                  = reinterpret_cast<T *>(stage_base + (size_t)NS * stage_sz); // phase 2 only
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
        // ============================== producer warp (as hmatrix_mm_kernel) ==============================
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
                    for (uint32_t i = lane; i < n_z; i += 32) {
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
    const int ncols8 = (p.ncols + 7) & ~7;
    T *ws = (T *)p.ws;
    uint32_t ul = 0;
    const T *xres = nullptr;
    T acc0[NCOL], acc1[NCOL]; // phase 2: bin rows lane and lane + 32
#pragma unroll
    for (int j = 0; j < NCOL; j++) acc0[j] = acc1[j] = S::zero();

    for (uint32_t it = 0;; it++) {
        const int slot     = it % NS;
        const uint32_t par = (it / NS) & 1u;
        ptx::mbar_wait(full_bar(slot), par);
        const uint8_t *pk       = stage_base + (size_t)slot * stage_sz;
        const uint4 hraw        = *reinterpret_cast<const uint4 *>(pk);
        const uint32_t n_pieces = hraw.x, piece_base = hraw.y, row0 = hraw.z;
        const uint32_t bin_rows = hraw.w & 0xffffu, flags = hraw.w >> 16;
        if (flags & kFlagTerminate) break;
        const int first = (int)((NW - (piece_base % NW) + warp) % NW); // pieces with (piece_base + i) % NW == warp

        if (PHASE == 2) {
            const T *zarea = reinterpret_cast<const T *>(pk + p.stage_bytes);
            for (uint32_t i = first; i < n_pieces; i += NW) {
                const uint4 d = *reinterpret_cast<const uint4 *>(pk + 16 + 16 * i);
                const T *A    = reinterpret_cast<const T *>(pk + d.x);
                const int nc = (int)(d.z >> 16), nr = (int)(d.w & 0xffu), ro = (int)((d.w >> 8) & 0xffu);
                // tile nr x nc (ld nr) at bin rows [ro, ro + nr)
                const int i0  = lane - ro, i1 = lane + 32 - ro;
                const bool a0 = (unsigned)i0 < (unsigned)nr, a1 = (unsigned)i1 < (unsigned)nr;
                const T *p0   = A + (a0 ? i0 : 0);
                const T *p1   = A + (a1 ? i1 : 0);
                const T *zr   = zarea + (size_t)(d.z & 0xffffu) * ld;
                for (int c = 0; c < nc; c++, p0 += nr, p1 += nr, zr += ld) {
                    const T v0 = a0 ? p0[0] : S::zero(), v1 = a1 ? p1[0] : S::zero();
                    mmf_axpy_row<T, NCOL>(acc0, v0, zr);
                    mmf_axpy_row<T, NCOL>(acc1, v1, zr);
                }
            }
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(empty_bar(slot));
            if (flags & kFlagLastOfBin) {
                const T alpha = mm_make<T>(p.alpha_re, p.alpha_im), beta = mm_make<T>(p.beta_re, p.beta_im);
                const bool beta_zero = S::is_zero(beta);
                T *y = (T *)p.y;
#pragma unroll
                for (int cb = 0; cb < NCOL; cb += CC) { // one reduction round per chunk of CC columns
                    if (cb < ncols8) {
#pragma unroll
                        for (int j = 0; j < CC; j++) {
                            red[((size_t)warp * kMaxBinRows + lane) * (CC + 1) + j]      = acc0[cb + j];
                            red[((size_t)warp * kMaxBinRows + 32 + lane) * (CC + 1) + j] = acc1[cb + j];
                        }
                    }
                    ptx::named_bar_sync(1, NW * 32);
                    if (cb < ncols8) {
                        for (uint32_t idx = threadIdx.x; idx < (uint32_t)(kMaxBinRows * CC); idx += (uint32_t)NW * 32u) {
                            const uint32_t r_ = idx / CC, j = idx % CC;
                            const int col = cb + (int)j;
                            if (r_ >= bin_rows || col >= p.ncols) continue;
                            T s = red[(size_t)r_ * (CC + 1) + j];
                            for (int w = 1; w < NW; w++) s = S::add(s, red[((size_t)w * kMaxBinRows + r_) * (CC + 1) + j]);
                            T r = S::mul(alpha, s);
                            T *yr = y + (size_t)(row0 + r_) * p.ldy + col;
                            if (!beta_zero) r = S::add(r, S::mul(beta, *yr));
                            *yr = r;
                        }
                    }
                    ptx::named_bar_sync(1, NW * 32); // the buffer is reused by the next round / the next bin
                }
#pragma unroll
                for (int j = 0; j < NCOL; j++) acc0[j] = acc1[j] = S::zero();
            }
        } else {
            if (flags & kFlagFirstOfBin) {
                const int buf = ul & 1;
                ptx::mbar_wait(xfull_bar(buf), (ul >> 1) & 1u);
                xres = reinterpret_cast<const T *>(x_base + (size_t)buf * x_bytes);
            }
            for (uint32_t i = first; i < n_pieces; i += NW) {
                const uint4 d = *reinterpret_cast<const uint4 *>(pk + 16 + 16 * i);
                const T *A    = reinterpret_cast<const T *>(pk + d.x);
                const int nq = (int)(d.z >> 16), lr = (int)(d.w & 0xffu);
                const T *xr0 = xres + (size_t)(d.z & 0xffffu) * ld;
                for (int rb = 0; rb < lr; rb += 32) { // 32 output rows at a time (lr <= 32 for every low-rank tile in practice)
                    const int rows = min(32, lr - rb);
                    const int G    = 32 / rows;            // groups of lanes sharing the contraction index
                    const int g    = lane / rows, l = lane - g * rows;
                    T s[NCOL];
#pragma unroll
                    for (int j = 0; j < NCOL; j++) s[j] = S::zero();
                    if (g < G) {
                        const T *pa = A + rb + l + (size_t)g * lr; // element (rb + l, k) sits at k*lr + rb + l
                        const T *pz = xr0 + (size_t)g * ld;
                        for (int k = g; k < nq; k += G, pa += (size_t)G * lr, pz += (size_t)G * ld) mmf_axpy_row<T, NCOL>(s, pa[0], pz);
                    }
                    for (int gg = 1; gg < G; gg++) {
#pragma unroll
                        for (int j = 0; j < NCOL; j++) {
                            const T o = S::shfl(s[j], lane < rows ? lane + gg * rows : lane);
                            if (lane < rows) s[j] = S::add(s[j], o);
                        }
                    }
                    if (lane < rows) {
                        T *orow = ws + (size_t)(d.y + rb + lane) * ld;
#pragma unroll
                        for (int j = 0; j < NCOL; j++)
                            if (j < ncols8) orow[j] = s[j];
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

} // namespace hmb200
