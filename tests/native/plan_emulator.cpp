// tests/native/plan_emulator.cpp -- TEST-ONLY host interpreter of the packed packet streams.
//
// Runs the planner + packer of the product library (htool_python_b200/csrc/capi/planner.cpp) and then walks the packed
// bytes exactly as the CUDA kernel does (packet header -> piece descriptors -> tiles, operand slices fetched through
// the ZCopy table into a stage-local area, per-piece partials, reduce segments, per-bin accumulation), on the CPU.
// It lets the CPU test-suite (-m "not gpu") validate the schedule and the packed layout against the oracle without a
// GPU.  It is NOT part of libhmb200.so and is never used by the product path.
#include "planner.h"

#include <limits>
#include <complex>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

using namespace hmb200;

template <typename T>
static void emulate(const hmb200_flat &flat, char op, const Options &opt, T alpha, const T *in, T beta, T *out, int64_t *stats) {
    const int es = sizeof(T);
    Plan plan;
    build_plan(flat, es, op, opt, plan);
    std::vector<T> xz(plan.xz_elems + 4, T(0));
    std::copy(in, in + plan.n_in, xz.begin());
    std::vector<uint8_t> zarea(opt.stage_z_bytes);
    for (int phase = 1; phase <= 2; phase++) {
        const PhasePlan &ph = phase == 1 ? plan.phase1 : plan.phase2;
        std::vector<uint8_t> stream(ph.stream_bytes + 256, 0xAB);
        if (!ph.stages.empty()) pack_stages(flat, plan, phase, 0, ph.stages.size(), stream.data());
        if (phase == 2) { // between the phases: the partial reduction kernel
            for (const ReduceSeg &rs : plan.reduce)
                for (uint32_t l = 0; l < rs.len; l++) {
                    T s = 0;
                    for (uint32_t q = 0; q < rs.n_parts; q++) s += xz[rs.first_partial + (size_t)q * rs.len + l];
                    xz[rs.final_off + l] = s;
                }
        }
        std::vector<char> row_written(plan.n_out, 0);
        for (const WorkUnit &wu : ph.units) {
            T acc[kMaxBinRows];
            for (auto &a : acc) a = T(0);
            for (uint32_t s = wu.first_stage; s < wu.first_stage + wu.n_stages; s++) {
                const StageEntry &se = ph.stages[s];
                if (se.bytes > (uint32_t)opt.stage_bytes || (se.bytes & 15) || (se.gmem_off & 15)) throw std::runtime_error("bad packet size/alignment");
                if (se.z_bytes > (uint32_t)opt.stage_z_bytes) throw std::runtime_error("z-area overflow");
                const uint8_t *pk = stream.data() + se.gmem_off;
                std::memset(zarea.data(), 0xCD, zarea.size());
                uint32_t zsum = 0;
                for (uint32_t i = 0; i < se.n_z; i++) {
                    const ZCopy &zc = ph.zcopies[se.z_first + i];
                    if ((zc.bytes & 15) || (zc.dst_off & 15) || (((size_t)zc.src_elem * es) & 15)) throw std::runtime_error("misaligned operand copy");
                    if ((size_t)zc.src_elem * es + zc.bytes > xz.size() * es) throw std::runtime_error("operand copy out of bounds");
                    if ((uint32_t)zc.dst_off + zc.bytes > (uint32_t)opt.stage_z_bytes) throw std::runtime_error("operand copy overflows the z-area");
                    std::memcpy(zarea.data() + zc.dst_off, (const uint8_t *)xz.data() + (size_t)zc.src_elem * es, zc.bytes);
                    zsum += zc.bytes;
                }
                if (zsum != se.z_bytes) throw std::runtime_error("z_bytes mismatch");
                PacketHeader h;
                std::memcpy(&h, pk, 16);
                const T *z_base = (const T *)zarea.data();
                for (uint32_t i = 0; i < h.n_pieces; i++) {
                    PieceDesc d;
                    std::memcpy(&d, pk + 16 + 16 * (size_t)i, 16);
                    if (d.data_off & 15) throw std::runtime_error("misaligned tile");
                    const T *A = (const T *)(pk + d.data_off);
                    const T *z = z_base + d.z_off;
                    if (d.data_off + (size_t)d.nr * d.nc * es > se.bytes) throw std::runtime_error("tile exceeds packet");
                    if (phase == 1) {
                        if (d.nr > kMaxPhase1Rows) throw std::runtime_error("phase-1 tile too tall");
                        for (int l = 0; l < d.nr; l++) {
                            T s = 0;
                            for (int k = 0; k < d.nc; k++) s += A[(size_t)k * d.nr + l] * z[k];
                            xz[d.out_off + l] = s;
                        }
                    } else {
                        if (d.ro + d.nr > h.bin_rows) throw std::runtime_error("tile exceeds bin");
                        for (int c = 0; c < d.nc; c++)
                            for (int i2 = 0; i2 < d.nr; i2++) acc[d.ro + i2] += A[(size_t)c * d.nr + i2] * z[c];
                    }
                }
                if (phase == 2 && (h.flags & kFlagLastOfBin)) {
                    for (int r = 0; r < h.bin_rows; r++) {
                        T v = alpha * acc[r];
                        if (beta != T(0)) v += beta * out[h.row0 + r];
                        out[h.row0 + r] = v;
                        if (row_written[h.row0 + r]) throw std::runtime_error("output row written twice");
                        row_written[h.row0 + r] = 1;
                    }
                    for (auto &a : acc) a = T(0);
                }
            }
        }
        if (phase == 2)
            for (int r = 0; r < plan.n_out; r++)
                if (!row_written[r]) throw std::runtime_error("output row " + std::to_string(r) + " never written");
    }
    if (stats) {
        stats[0] = (int64_t)plan.phase1.stream_bytes; stats[1] = (int64_t)plan.phase2.stream_bytes;
        stats[2] = (int64_t)plan.pieces1.size(); stats[3] = (int64_t)plan.pieces2.size();
        stats[4] = (int64_t)plan.phase1.stages.size(); stats[5] = (int64_t)plan.phase2.stages.size();
        stats[6] = (int64_t)plan.reduce.size(); stats[7] = (int64_t)plan.bin_start.size() - 1;
        stats[8] = (int64_t)plan.xz_elems; stats[9] = (int64_t)plan.stored_coefficients;
        stats[10] = (int64_t)plan.phase1.zcopies.size(); stats[11] = (int64_t)plan.phase2.zcopies.size();
    }
}


// multi-RHS path: phase-1m packets (source-bin grouped V tiles, resident X rows), reduce_mm, then the phase-2 packets with
// every operand element widened to a row of mu -- the data flow of kernels/mm_kernels.cuh, on the CPU.
template <typename T>
static void emulate_mm(const hmb200_flat &flat, char op, const Options &opt, T alpha, const T *in, T beta, T *out, int mu, int64_t *stats) {
    const int es = sizeof(T);
    Plan plan;
    build_plan(flat, es, op, opt, plan);
    if (!opt.build_matmat) throw std::runtime_error("plan has no multi-RHS stream");
    std::vector<T> ws((plan.mm_rows + 4) * (size_t)mu, T(0));
    std::copy(in, in + (size_t)plan.n_in * mu, ws.begin());
    {   // phase 1m
        const PhasePlan &ph = plan.phase1m;
        std::vector<uint8_t> stream(ph.stream_bytes + 256, 0xAB);
        if (!ph.stages.empty()) pack_stages(flat, plan, 3, 0, ph.stages.size(), stream.data());
        for (const WorkUnit &wu : ph.units)
            for (uint32_t s = wu.first_stage; s < wu.first_stage + wu.n_stages; s++) {
                const StageEntry &se = ph.stages[s];
                if (se.bytes > (uint32_t)opt.mm_stage_bytes || (se.bytes & 15) || (se.gmem_off & 15)) throw std::runtime_error("bad mm packet size/alignment");
                if (se.n_z != 0) throw std::runtime_error("mm phase-1 packet with operand copies");
                const uint32_t x0 = se.pad[0], xr = se.pad[1];
                if (xr > (uint32_t)opt.mm_bin_cols || x0 + xr > (uint32_t)plan.n_in) throw std::runtime_error("bad resident X slice");
                const uint8_t *pk = stream.data() + se.gmem_off;
                PacketHeader h;
                std::memcpy(&h, pk, 16);
                if (h.row0 != x0 || h.bin_rows != xr) throw std::runtime_error("header / stage entry mismatch");
                for (uint32_t i = 0; i < h.n_pieces; i++) {
                    PieceDesc d;
                    std::memcpy(&d, pk + 16 + 16 * (size_t)i, 16);
                    const T *A = (const T *)(pk + d.data_off);
                    if (d.data_off + (size_t)d.nr * d.nc * es > se.bytes) throw std::runtime_error("mm tile exceeds packet");
                    if ((uint32_t)d.z_off + d.nc > xr) throw std::runtime_error("mm tile reads outside the resident X slice");
                    if ((uint64_t)d.out_off + d.nr > plan.mm_rows) throw std::runtime_error("mm output row out of range");
                    for (int l = 0; l < d.nr; l++)
                        for (int j = 0; j < mu; j++) {
                            T sacc = 0;
                            for (int k = 0; k < d.nc; k++) sacc += A[(size_t)k * d.nr + l] * ws[(size_t)(x0 + d.z_off + k) * mu + j];
                            ws[(size_t)(d.out_off + l) * mu + j] = sacc;
                        }
                }
            }
    }
    for (const ReduceSeg &rs : plan.reduce_mm)
        for (uint32_t l = 0; l < rs.len; l++)
            for (int j = 0; j < mu; j++) {
                T sacc = 0;
                for (uint32_t q = 0; q < rs.n_parts; q++) sacc += ws[(size_t)(plan.reduce_mm_rows[rs.first_partial + q] + l) * mu + j];
                ws[(size_t)(rs.final_off + l) * mu + j] = sacc;
            }
    {   // phase 2 on the single-RHS stream2, operands as rows
        const PhasePlan &ph = plan.phase2;
        std::vector<uint8_t> stream(ph.stream_bytes + 256, 0xAB);
        if (!ph.stages.empty()) pack_stages(flat, plan, 2, 0, ph.stages.size(), stream.data());
        std::vector<T> zrows((size_t)opt.stage_z_elems * mu);
        std::vector<T> acc((size_t)kMaxBinRows * mu);
        for (const WorkUnit &wu : ph.units) {
            std::fill(acc.begin(), acc.end(), T(0));
            for (uint32_t s = wu.first_stage; s < wu.first_stage + wu.n_stages; s++) {
                const StageEntry &se = ph.stages[s];
                if (se.z_bytes > (uint32_t)opt.stage_z_elems * es) throw std::runtime_error("phase-2 packet references more operand rows than stage_z_elems");
                for (uint32_t i = 0; i < se.n_z; i++) {
                    const ZCopy &zc = ph.zcopies[se.z_first + i];
                    if ((uint64_t)zc.src_elem + zc.bytes / es > plan.mm_rows + 2) throw std::runtime_error("operand rows out of range");
                    std::memcpy(zrows.data() + (size_t)(zc.dst_off / es) * mu, ws.data() + (size_t)zc.src_elem * mu, (size_t)(zc.bytes / es) * mu * es);
                }
                const uint8_t *pk = stream.data() + se.gmem_off;
                PacketHeader h;
                std::memcpy(&h, pk, 16);
                for (uint32_t i = 0; i < h.n_pieces; i++) {
                    PieceDesc d;
                    std::memcpy(&d, pk + 16 + 16 * (size_t)i, 16);
                    const T *A = (const T *)(pk + d.data_off);
                    for (int c = 0; c < d.nc; c++)
                        for (int i2 = 0; i2 < d.nr; i2++)
                            for (int j = 0; j < mu; j++) acc[(size_t)(d.ro + i2) * mu + j] += A[(size_t)c * d.nr + i2] * zrows[(size_t)(d.z_off + c) * mu + j];
                }
                if (h.flags & kFlagLastOfBin) {
                    for (int r = 0; r < h.bin_rows; r++)
                        for (int j = 0; j < mu; j++) {
                            T v = alpha * acc[(size_t)r * mu + j];
                            if (beta != T(0)) v += beta * out[(size_t)(h.row0 + r) * mu + j];
                            out[(size_t)(h.row0 + r) * mu + j] = v;
                        }
                    std::fill(acc.begin(), acc.end(), T(0));
                }
            }
        }
    }
    if (stats) {
        stats[0] = (int64_t)plan.phase1m.stream_bytes; stats[1] = (int64_t)plan.phase2.stream_bytes;
        stats[2] = (int64_t)plan.mm_tiles.size(); stats[3] = (int64_t)plan.pieces2.size();
        stats[4] = (int64_t)plan.phase1m.stages.size(); stats[5] = (int64_t)plan.phase2.stages.size();
        stats[6] = (int64_t)plan.reduce_mm.size(); stats[7] = (int64_t)plan.src_bin_start.size() - 1;
        stats[8] = (int64_t)plan.mm_rows; stats[9] = (int64_t)plan.stored_coefficients;
    }
}

// LANE-ACCURATE emulation of the experimental DFMA consumer (kernels/mmf_kernels.cuh): same packets and tables as emulate_mm,
// but the work is split exactly like the kernel does it -- tiles dealt to 8 warps by (piece_base + i) % 8, phase 1 lanes
// = (column group g, row l) with shuffle-style group sums, phase 2 lanes own bin rows lane / lane + 32 with NCOL accumulators,
// cross-warp sums in chunks through a [warp][row][chunk + 1] buffer, workspace rows at the padded stride ld with NaN in the
// pad columns -- so that the index arithmetic of the kernel is checked on the CPU (the TMA / mbarrier pipeline is not).
template <typename T>
static void emulate_mmf(const hmb200_flat &flat, char op, const Options &opt, T alpha, const T *in, T beta, T *out, int mu) {
    const int es = sizeof(T);
    constexpr int NW = 8;
    const int CC = es == 8 ? 8 : 4, PC = es == 8 ? 32 : 16;
    Plan plan;
    build_plan(flat, es, op, opt, plan);
    if (!opt.build_matmat) throw std::runtime_error("plan has no multi-RHS stream");
    const T nanv = T(std::numeric_limits<double>::quiet_NaN());
    const PhasePlan &ph1 = plan.phase1m, &ph2 = plan.phase2;
    std::vector<uint8_t> stream1(ph1.stream_bytes + 256, 0xAB), stream2(ph2.stream_bytes + 256, 0xAB);
    if (!ph1.stages.empty()) pack_stages(flat, plan, 3, 0, ph1.stages.size(), stream1.data());
    if (!ph2.stages.empty()) pack_stages(flat, plan, 2, 0, ph2.stages.size(), stream2.data());
    for (int c0 = 0; c0 < mu; c0 += PC) {
        const int nc = std::min(PC, mu - c0), nc8 = (nc + 7) & ~7;
        const int ld = es == 16 ? nc8 + 2 : nc8 + ((nc8 % 16 == 0) ? 4 : 12);
        const int NCOL = nc8 <= 8 ? 8 : ((nc8 <= 16 || es == 16) ? 16 : 32);
        if (NCOL > ld || NCOL < nc8) throw std::runtime_error("mmf: NCOL / ld mismatch");
        std::vector<T> ws((plan.mm_rows + 4) * (size_t)ld, nanv);
        for (size_t i = 0; i < (size_t)plan.n_in; i++)
            for (int j = 0; j < nc; j++) ws[i * ld + j] = in[i * mu + c0 + j]; // mm_load_rows_kernel: only nc columns are written
        // ---- phase 1 ----
        for (const WorkUnit &wu : ph1.units)
            for (uint32_t s = wu.first_stage; s < wu.first_stage + wu.n_stages; s++) {
                const StageEntry &se = ph1.stages[s];
                const uint32_t x0 = se.pad[0];
                const uint8_t *pk = stream1.data() + se.gmem_off;
                PacketHeader h;
                std::memcpy(&h, pk, 16);
                std::vector<int> taken(h.n_pieces, 0);
                for (int warp = 0; warp < NW; warp++) {
                    const int first = (int)((NW - (h.piece_base % NW) + warp) % NW);
                    for (uint32_t i = first; i < h.n_pieces; i += NW) {
                        taken[i]++;
                        PieceDesc d;
                        std::memcpy(&d, pk + 16 + 16 * (size_t)i, 16);
                        const T *A = (const T *)(pk + d.data_off);
                        const int nq = d.nc, lr = d.nr;
                        const T *xr0 = ws.data() + (size_t)(x0 + d.z_off) * ld; // = xres + z_off * ld in the kernel
                        for (int rb = 0; rb < lr; rb += 32) {
                            const int rows = std::min(32, lr - rb), G = 32 / rows;
                            std::vector<T> sacc((size_t)32 * NCOL, T(0));
                            for (int lane = 0; lane < 32; lane++) {
                                const int g = lane / rows, l = lane - g * rows;
                                if (g >= G) continue;
                                for (int k = g; k < nq; k += G)
                                    for (int j = 0; j < NCOL; j++) sacc[(size_t)lane * NCOL + j] += A[(size_t)k * lr + rb + l] * xr0[(size_t)k * ld + j];
                            }
                            for (int gg = 1; gg < G; gg++)
                                for (int lane = 0; lane < rows; lane++)
                                    for (int j = 0; j < NCOL; j++) sacc[(size_t)lane * NCOL + j] += sacc[(size_t)(lane + gg * rows) * NCOL + j];
                            for (int lane = 0; lane < rows; lane++)
                                for (int j = 0; j < NCOL; j++)
                                    if (j < nc8) ws[(size_t)(d.out_off + rb + lane) * ld + j] = sacc[(size_t)lane * NCOL + j];
                        }
                    }
                }
                for (int t : taken)
                    if (t != 1) throw std::runtime_error("mmf: a phase-1 piece is not taken exactly once");
            }
        // ---- reduce_rows_kernel (ncols8 columns) ----
        for (const ReduceSeg &rs : plan.reduce_mm)
            for (uint32_t l = 0; l < rs.len; l++)
                for (int j = 0; j < nc8; j++) {
                    T sacc = 0;
                    for (uint32_t q = 0; q < rs.n_parts; q++) sacc += ws[(size_t)(plan.reduce_mm_rows[rs.first_partial + q] + l) * ld + j];
                    ws[(size_t)(rs.final_off + l) * ld + j] = sacc;
                }
        // ---- phase 2 ----
        std::vector<T> zrows((size_t)opt.stage_z_elems * ld);
        std::vector<T> acc((size_t)NW * kMaxBinRows * NCOL);        // [warp][bin row][column]: acc0 = rows < 32, acc1 = rows >= 32
        std::vector<T> red((size_t)NW * kMaxBinRows * (CC + 1));
        for (const WorkUnit &wu : ph2.units) {
            std::fill(acc.begin(), acc.end(), T(0));
            for (uint32_t s = wu.first_stage; s < wu.first_stage + wu.n_stages; s++) {
                const StageEntry &se = ph2.stages[s];
                std::fill(zrows.begin(), zrows.end(), nanv);
                for (uint32_t i = 0; i < se.n_z; i++) {
                    const ZCopy &zc = ph2.zcopies[se.z_first + i];
                    // bulk copy of (bytes * ld) bytes to byte offset (dst_off * ld): whole padded rows
                    std::memcpy(zrows.data() + (size_t)(zc.dst_off / es) * ld, ws.data() + (size_t)zc.src_elem * ld, (size_t)(zc.bytes / es) * ld * es);
                }
                const uint8_t *pk = stream2.data() + se.gmem_off;
                PacketHeader h;
                std::memcpy(&h, pk, 16);
                for (int warp = 0; warp < NW; warp++) {
                    const int first = (int)((NW - (h.piece_base % NW) + warp) % NW);
                    for (uint32_t i = first; i < h.n_pieces; i += NW) {
                        PieceDesc d;
                        std::memcpy(&d, pk + 16 + 16 * (size_t)i, 16);
                        const T *A = (const T *)(pk + d.data_off);
                        for (int lane = 0; lane < 32; lane++)
                            for (int half = 0; half < 2; half++) {
                                const int ir = lane + 32 * half - d.ro;
                                if (ir < 0 || ir >= d.nr) continue;
                                for (int c = 0; c < d.nc; c++)
                                    for (int j = 0; j < NCOL; j++)
                                        acc[((size_t)warp * kMaxBinRows + lane + 32 * half) * NCOL + j] += A[(size_t)c * d.nr + ir] * zrows[(size_t)(d.z_off + c) * ld + j];
                            }
                    }
                }
                if (h.flags & kFlagLastOfBin) {
                    for (int cb = 0; cb < NCOL; cb += CC) {
                        if (cb >= nc8) continue;
                        for (int warp = 0; warp < NW; warp++)
                            for (int r = 0; r < kMaxBinRows; r++)
                                for (int j = 0; j < CC; j++) red[((size_t)warp * kMaxBinRows + r) * (CC + 1) + j] = acc[((size_t)warp * kMaxBinRows + r) * NCOL + cb + j];
                        for (int idx = 0; idx < kMaxBinRows * CC; idx++) {
                            const int r = idx / CC, j = idx % CC, col = cb + j;
                            if (r >= h.bin_rows || col >= nc) continue;
                            T sacc = red[(size_t)r * (CC + 1) + j];
                            for (int w = 1; w < NW; w++) sacc += red[((size_t)w * kMaxBinRows + r) * (CC + 1) + j];
                            T v = alpha * sacc;
                            T *yr = out + (size_t)(h.row0 + r) * mu + c0 + col;
                            if (beta != T(0)) v += beta * *yr;
                            *yr = v;
                        }
                    }
                    std::fill(acc.begin(), acc.end(), T(0));
                }
            }
        }
    }
}

static thread_local std::string g_err;

extern "C" int plan_emulate(const hmb200_flat *flat, int elem_bytes, char op, const hmb200_options *opt, const double *alpha, const void *in, const double *beta, void *out, int64_t *stats) {
    try {
        Options o = resolve_options(opt);
        if (elem_bytes == 8)
            emulate<double>(*flat, op, o, alpha[0], (const double *)in, beta[0], (double *)out, stats);
        else
            emulate<std::complex<double>>(*flat, op, o, {alpha[0], alpha[1]}, (const std::complex<double> *)in, {beta[0], beta[1]}, (std::complex<double> *)out, stats);
        return 0;
    } catch (const std::exception &e) {
        g_err = e.what();
        return 1;
    }
}
extern "C" const char *plan_emulate_error() { return g_err.c_str(); }

extern "C" int plan_emulate_mm(const hmb200_flat *flat, int elem_bytes, char op, const hmb200_options *opt, const double *alpha, const void *in, const double *beta, void *out, int mu, int64_t *stats) {
    try {
        Options o = resolve_options(opt);
        if (elem_bytes == 8)
            emulate_mm<double>(*flat, op, o, alpha[0], (const double *)in, beta[0], (double *)out, mu, stats);
        else
            emulate_mm<std::complex<double>>(*flat, op, o, {alpha[0], alpha[1]}, (const std::complex<double> *)in, {beta[0], beta[1]}, (std::complex<double> *)out, mu, stats);
        return 0;
    } catch (const std::exception &e) {
        g_err = e.what();
        return 1;
    }
}

extern "C" int plan_emulate_mmf(const hmb200_flat *flat, int elem_bytes, char op, const hmb200_options *opt, const double *alpha, const void *in, const double *beta, void *out, int mu) {
    try {
        Options o = resolve_options(opt);
        if (elem_bytes == 8)
            emulate_mmf<double>(*flat, op, o, alpha[0], (const double *)in, beta[0], (double *)out, mu);
        else
            emulate_mmf<std::complex<double>>(*flat, op, o, {alpha[0], alpha[1]}, (const std::complex<double> *)in, {beta[0], beta[1]}, (std::complex<double> *)out, mu);
        return 0;
    } catch (const std::exception &e) {
        g_err = e.what();
        return 1;
    }
}
