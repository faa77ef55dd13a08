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


// multi-RHS path: the two panel streams of kernels/mm_kernels.cuh walked on the CPU exactly as the kernel's lanes do -- packet
// header, fragment-major chunks decoded per (k-group, m-tile, lane), operand rows gathered through the ZCopy table into a
// stage-local area, accumulators per (panel row, column), row map / reduce segments / bin epilogue.  Rows the kernels never
// write hold NaN here, pad columns too, so that a wrong row, column or K-padding operand poisons the result.
template <typename T>
static void emulate_mm(const hmb200_flat &flat, char op, const Options &opt, T alpha, const T *in, T beta, T *out, int mu, int64_t *stats) {
    const int es = sizeof(T);
    constexpr int KG = sizeof(T) == 8 ? 8 : 4;
    Plan plan;
    build_plan(flat, es, op, opt, plan);
    if (!opt.build_matmat) throw std::runtime_error("plan has no multi-RHS stream");
    const double qnan = std::numeric_limits<double>::quiet_NaN();
    const T nanv = T(qnan);
    const int PC = es == 8 ? 64 : 32;
    for (int c0 = 0; c0 < mu; c0 += PC) {
        const int nc = std::min(PC, mu - c0), nc8 = (nc + 7) & ~7, ld = nc8 + 4;
        std::vector<T> ws((size_t)plan.mm_rows * ld, nanv);
        for (size_t i = 0; i < (size_t)plan.n_in; i++)
            for (int j = 0; j < nc; j++) ws[i * ld + j] = in[i * mu + c0 + j]; // mm_load_rows_kernel writes only the pass's columns
        for (int r = 0; r < kMmZeroRows; r++)
            for (int j = 0; j < ld; j++) ws[(size_t)(plan.mm_zero_row + r) * ld + j] = T(0);
        if ((uint64_t)plan.mm_zero_row + kMmZeroRows != plan.mm_rows) throw std::runtime_error("zero rows are not the last workspace rows");
        for (int phase = 1; phase <= 2; phase++) {
            const MmPhasePlan &ph = phase == 1 ? plan.mm1 : plan.mm2;
            const int rows_max = phase == 1 ? kMmRows1 : kMmRows2, kg_max = phase == 1 ? opt.mm1_kg : opt.mm2_kg;
            std::vector<uint8_t> stream(ph.stream_bytes + 256, 0xAB);
            if (!ph.stages.empty()) pack_stages(flat, plan, phase + 2, 0, ph.stages.size(), stream.data());
            if (ph.stages.size() != ph.packets.size()) throw std::runtime_error("stage / packet tables differ in length");
            if (phase == 2) // between the phases: reduce_rows_kernel
                for (const ReduceSeg &rs : plan.reduce_mm)
                    for (uint32_t l = 0; l < rs.len; l++)
                        for (int j = 0; j < nc8; j++) {
                            T sacc = T(0);
                            for (uint32_t q = 0; q < rs.n_parts; q++) sacc += ws[(size_t)(plan.reduce_mm_rows[rs.first_partial + q] + l) * ld + j];
                            ws[(size_t)(rs.final_off + l) * ld + j] = sacc;
                        }
            std::vector<char> row_written(plan.n_out, 0);
            std::vector<T> zarea((size_t)kg_max * KG * ld), acc((size_t)rows_max * nc8);
            for (const WorkUnit &wu : ph.units) {
                std::fill(acc.begin(), acc.end(), T(0));
                for (uint32_t s = wu.first_stage; s < wu.first_stage + wu.n_stages; s++) {
                    const StageEntry &se = ph.stages[s];
                    if ((se.bytes & 15) || (se.gmem_off & 127)) throw std::runtime_error("bad mm packet size/alignment");
                    const uint8_t *pk = stream.data() + se.gmem_off;
                    MmPacketHeader h;
                    std::memcpy(&h, pk, 16);
                    if (h.n_kg > (uint32_t)kg_max || h.rows > rows_max || h.mt_panel != (h.rows + 7) / 8) throw std::runtime_error("bad mm packet header");
                    if (((h.flags & kMmFlagFirst) != 0) != (s == wu.first_stage) || ((h.flags & kMmFlagLast) != 0) != (s + 1 == wu.first_stage + wu.n_stages)) throw std::runtime_error("first/last flags do not match the unit");
                    if (se.bytes != 16 + h.n_kg * h.mt_panel * 512u + ((phase == 1 && (h.flags & kMmFlagLast)) ? (h.rows * 4u + 15) / 16 * 16 : 0u)) throw std::runtime_error("mm packet bytes mismatch");
                    std::fill(zarea.begin(), zarea.end(), nanv);
                    uint32_t zrows = 0;
                    for (uint32_t i = 0; i < se.n_z; i++) {
                        const ZCopy &zc = ph.zcopies[se.z_first + i];
                        const uint32_t nr = zc.bytes / es, dr = zc.dst_off / es;
                        if (zc.bytes % es || zc.dst_off % es || dr != zrows) throw std::runtime_error("operand rows are not laid out consecutively");
                        if ((uint64_t)zc.src_elem + nr > plan.mm_rows) throw std::runtime_error("operand rows out of range");
                        if (dr + nr > (uint32_t)kg_max * KG) throw std::runtime_error("operand rows overflow the z-area");
                        std::memcpy(zarea.data() + (size_t)dr * ld, ws.data() + (size_t)zc.src_elem * ld, (size_t)nr * ld * es);
                        zrows += nr;
                    }
                    if (zrows != h.n_kg * KG || se.z_bytes != zrows * (uint32_t)es) throw std::runtime_error("operand row count mismatch");
                    const T *chunks = reinterpret_cast<const T *>(pk + 16);
                    for (uint32_t kg = 0; kg < h.n_kg; kg++)
                        for (int mt = 0; mt < h.mt_panel; mt++)
                            for (int lane = 0; lane < 32; lane++) {
                                const int gid = lane >> 2, tig = lane & 3, row = 8 * mt + gid;
                                const T *c = chunks + ((size_t)kg * h.mt_panel + mt) * (512 / sizeof(T));
                                for (int half = 0; half < (es == 8 ? 2 : 1); half++) {
                                    const T a = es == 8 ? c[2 * lane + half] : c[lane];
                                    const int k = (int)kg * KG + 4 * half + tig;
                                    if (row >= h.rows) {
                                        if (a != T(0)) throw std::runtime_error("non-zero coefficient in a pad row of a panel");
                                        continue;
                                    }
                                    if (a == T(0)) continue; // zero coefficients (K padding, rows outside a piece) skip their operand, as 0 * finite does
                                    for (int j = 0; j < nc8; j++) acc[(size_t)row * nc8 + j] += a * zarea[(size_t)k * ld + j];
                                }
                            }
                    // K padding must multiply finite rows: the kernel does NOT skip zeros
                    for (uint32_t k = 0; k < h.n_kg * KG; k++)
                        for (int j = 0; j < nc; j++)
                            if (zarea[(size_t)k * ld + j] != zarea[(size_t)k * ld + j]) throw std::runtime_error("NaN operand inside a packet's K range (columns of the pass)");
                    if (h.flags & kMmFlagLast) {
                        if (phase == 1) {
                            const uint32_t *map = reinterpret_cast<const uint32_t *>(pk + 16 * (size_t)h.map_off16);
                            if (16u * h.map_off16 != 16 + h.n_kg * h.mt_panel * 512u) throw std::runtime_error("row map offset");
                            for (int i = 0; i < h.rows; i++) {
                                if (map[i] < (uint32_t)plan.n_in || map[i] >= plan.mm_zero_row) throw std::runtime_error("row map target out of range");
                                for (int j = 0; j < nc8; j++) ws[(size_t)map[i] * ld + j] = acc[(size_t)i * nc8 + j];
                            }
                        } else {
                            for (int i = 0; i < h.rows; i++) {
                                const size_t r = (size_t)h.out0 + i;
                                if (r >= (size_t)plan.n_out || row_written[r]) throw std::runtime_error("output row written twice / out of range");
                                row_written[r] = 1;
                                for (int j = 0; j < nc; j++) {
                                    T v = alpha * acc[(size_t)i * nc8 + j];
                                    if (beta != T(0)) v += beta * out[r * mu + c0 + j];
                                    out[r * mu + c0 + j] = v;
                                }
                            }
                        }
                    }
                }
            }
            if (phase == 2)
                for (int r = 0; r < plan.n_out; r++)
                    if (!row_written[r]) throw std::runtime_error("output row never written");
        }
    }
    if (stats) {
        stats[0] = (int64_t)plan.mm1.stream_bytes; stats[1] = (int64_t)plan.mm2.stream_bytes;
        stats[2] = (int64_t)plan.mm1.pieces.size(); stats[3] = (int64_t)plan.mm2.pieces.size();
        stats[4] = (int64_t)plan.mm1.stages.size(); stats[5] = (int64_t)plan.mm2.stages.size();
        stats[6] = (int64_t)plan.reduce_mm.size(); stats[7] = (int64_t)plan.mm1.units.size();
        stats[8] = (int64_t)plan.mm_rows; stats[9] = (int64_t)plan.stored_coefficients;
        stats[10] = (int64_t)plan.mm1.zcopies.size(); stats[11] = (int64_t)plan.mm2.zcopies.size();
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


// plan only (no packing, the leaf data pointers are not touched): schedule statistics of large leaf tables
extern "C" int plan_stats(const hmb200_flat *flat, int elem_bytes, char op, const hmb200_options *opt, double *stats) {
    try {
        Options o = resolve_options(opt);
        Plan plan;
        build_plan(*flat, elem_bytes, op, o, plan);
        const int kgc = elem_bytes == 8 ? 8 : 4;
        double k1 = 0, k2 = 0, a1 = 0, a2 = 0;
        for (const auto &p : plan.mm1.packets) { k1 += (double)p.n_kg * kgc * p.rows; a1 += (double)p.n_kg * ((p.rows + 7) / 8) * 512.0 / elem_bytes; }
        for (const auto &p : plan.mm2.packets) { k2 += (double)p.n_kg * kgc * p.rows; a2 += (double)p.n_kg * ((p.rows + 7) / 8) * 512.0 / elem_bytes; }
        double s[] = {(double)plan.mm1.stream_bytes, (double)plan.mm2.stream_bytes, (double)plan.mm1.stages.size(), (double)plan.mm2.stages.size(),
                      (double)plan.mm1.units.size(), (double)plan.mm2.units.size(), (double)plan.mm1.zcopies.size(), (double)plan.mm2.zcopies.size(),
                      (double)plan.mm_rows, (double)plan.reduce_mm.size(), (double)plan.reduce_mm_rows.size(), k1, k2, a1, a2, plan.stored_coefficients,
                      (double)plan.phase1.stream_bytes, (double)plan.phase2.stream_bytes, (double)plan.phase1.zcopies.size(), (double)plan.phase2.zcopies.size(),
                      (double)plan.phase1.stages.size(), (double)plan.phase2.stages.size()};
        std::memcpy(stats, s, sizeof(s));
        return 0;
    } catch (const std::exception &e) {
        g_err = e.what();
        return 1;
    }
}
