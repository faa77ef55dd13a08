// planner.cpp -- builds the static schedule of the two-phase H-matrix product and packs the leaf data into it.
//
// What the schedule replaces in the reference: the per-call DFS get_leaves_from + one BLAS call pair per leaf
// (lib/htool/include/htool/hmatrix/linalg/add_hmatrix_vector_product.hpp:56-104) and the per-thread full-length
// temporaries + `omp critical` reduction of the OpenMP variant (:139-169).  Here every output row belongs to exactly
// one target bin, every bin's tiles are contiguous in HBM, and one CTA reduces a bin -- no atomics on y.
#include "planner.h"

#include <algorithm>
#include <complex>
#include <cstring>
#include <stdexcept>

namespace hmb200 {

static inline uint32_t align_up(uint32_t v, uint32_t a) { return (v + a - 1) / a * a; }
static inline uint64_t align_up64(uint64_t v, uint64_t a) { return (v + a - 1) / a * a; }

Options resolve_options(const hmb200_options *o) {
    Options r;
    if (o) {
        if (o->stage_bytes > 0) r.stage_bytes = o->stage_bytes;
        if (o->stage_z_bytes > 0) r.stage_z_bytes = o->stage_z_bytes;
        if (o->n_stages > 0) r.n_stages = o->n_stages;
        if (o->consumer_warps > 0) r.consumer_warps = o->consumer_warps;
        if (o->piece1_bytes > 0) r.piece1_bytes = o->piece1_bytes;
        if (o->piece2_bytes > 0) r.piece2_bytes = o->piece2_bytes;
        if (o->bin_rows > 0) r.bin_rows = o->bin_rows;
        if (o->unit_stages > 0) r.unit_stages = o->unit_stages;
        if (o->ctas > 0) r.ctas = o->ctas;
        r.build_transpose = o->build_transpose;
        r.build_matmat    = o->no_matmat ? 0 : 1;
        if (o->mm_task_cols > 0) r.mm_task_cols = o->mm_task_cols;
        if (o->mm1_kg > 0) r.mm1_kg = o->mm1_kg;
        if (o->mm2_kg > 0) r.mm2_kg = o->mm2_kg;
    }
    r.stage_bytes   = (int)align_up((uint32_t)std::max(r.stage_bytes, 2048), 128);
    r.stage_z_bytes = (int)align_up((uint32_t)std::min(std::max(r.stage_z_bytes, 1024), 65536 - 128), 128);
    r.bin_rows      = std::min(std::max(r.bin_rows, 1), kMaxBinRows);
    r.n_stages      = std::min(std::max(r.n_stages, 2), 16);
    r.consumer_warps = std::min(std::max(r.consumer_warps, 1), 16);
    int piece_cap   = r.stage_bytes - 16 - 16 - 64;
    r.piece1_bytes  = std::min(std::max(r.piece1_bytes, 256), piece_cap);
    r.piece2_bytes  = std::min(std::max(r.piece2_bytes, 256), piece_cap);
    r.unit_stages   = std::max(r.unit_stages, 1);
    // even: the kernels' K loop alternates two fragment sets and runs on across packet boundaries (mm_kernels.cuh)
    r.mm1_kg       = std::min(std::max(r.mm1_kg & ~1, 2), 6);   // 8 KiB per k-group of a 128-row panel
    r.mm2_kg       = std::min(std::max(r.mm2_kg & ~1, 2), 12);  // 4 KiB per k-group of a 64-row panel
    r.mm_task_cols = std::max(r.mm_task_cols, 8);
    return r;
}

// ------------------------------------------------------------------------------------------------------------
// virtual leaves: which stored leaves enter out = op(A) in, and in which orientation
// (add_hmatrix_vector_product.hpp:69,74-81 trans_sym logic; :90-103 the two loops)
// ------------------------------------------------------------------------------------------------------------
static void make_vleaves(const hmb200_flat &f, char op, Plan &plan) {
    const char sym = (char)f.symmetry_for_leaves;
    if ((op == 'T' && sym == 'H') || (op == 'C' && sym == 'S'))
        throw std::runtime_error(std::string("operation is not supported (trans=") + op + " with symmetry " + sym + ")"); // reference logs the same, :58-61
    plan.vleaves.clear();
    plan.vleaves.reserve((size_t)f.n_leaves);
    double W = 0;
    auto push = [&](int64_t l, bool transposed, bool conj) {
        VLeaf v;
        v.leaf = l;
        v.r    = f.rank[l];
        v.kind = (uint8_t)HMB200_LEAF_DENSE_KIND(f.flags[l]);
        if (v.kind != 0 && v.r < 0) { // symv/hemv leaves are their own (conjugate) transpose: reference ignores trans (:21-24,:41-46)
            // ... which is only right when op(leaf) == leaf: a symmetric ('S') leaf under 'C', or a Hermitian ('H') leaf under
            // 'T', would need conj(leaf).  The root-level test above excludes them when the whole block carries the symmetry;
            // a lone such leaf inside an 'N' block is refused like the root case rather than multiplied unconjugated.
            const bool herm = v.kind >= 3;
            if (plan.elem_bytes == 16 && ((op == 'C' && !herm) || (op == 'T' && herm)))
                throw std::runtime_error(std::string("operation is not supported (trans=") + op + " on a dense leaf stored with symmetry " + (herm ? "H" : "S") + ")");
            transposed = false;
            conj       = false;
        }
        v.transposed = transposed;
        v.conj       = conj;
        if (!transposed) {
            v.row_off = f.row_off[l]; v.m = f.nrows[l];
            v.col_off = f.col_off[l]; v.n = f.ncols[l];
        } else {
            v.row_off = f.col_off[l]; v.m = f.ncols[l];
            v.col_off = f.row_off[l]; v.n = f.nrows[l];
        }
        v.t_off = 0;
        if (v.m <= 0 || v.n <= 0 || v.r == 0)
            return; // empty blocks and rank-0 leaves contribute nothing (add_lrmat_vector_product.hpp:12)
        plan.vleaves.push_back(v);
    };
    for (int64_t l = 0; l < f.n_leaves; l++) {
        int m = f.nrows[l], n = f.ncols[l], r = f.rank[l];
        if (m < 0 || n < 0 || f.row_off[l] < 0 || f.col_off[l] < 0 || f.row_off[l] + m > f.n_rows || f.col_off[l] + n > f.n_cols)
            throw std::runtime_error("leaf " + std::to_string(l) + " lies outside the root block");
        if (r < 0)
            W += HMB200_LEAF_DENSE_KIND(f.flags[l]) ? m * (n + 1.) / 2. : double(m) * n;
        else
            W += double(r) * (m + n);
        push(l, op != 'N', op == 'C');
        if (sym != 'N' && (f.flags[l] & HMB200_LEAF_MIRROR))
            push(l, op == 'N', op == 'N' && sym == 'H');
    }
    plan.stored_coefficients = W;
    // order the views by output range: all views with the same target cluster become neighbours, so do their t slots
    // (assigned in this order) and the phase-2 pieces of a bin -- one operand copy then serves many tiles of a stage
    std::stable_sort(plan.vleaves.begin(), plan.vleaves.end(), [](const VLeaf &a, const VLeaf &b) {
        if (a.row_off != b.row_off) return a.row_off < b.row_off;
        if (a.m != b.m) return a.m > b.m;
        return a.col_off < b.col_off;
    });
}

// target bins: ranges of <= bin_rows output rows, aligned with the leaves' cluster boundaries where possible
static void make_bins(const Plan &plan, int n_out, int bin_rows, std::vector<int32_t> &bin_start) {
    std::vector<int32_t> cuts;
    cuts.reserve(plan.vleaves.size() * 2 + 2);
    cuts.push_back(0);
    cuts.push_back(n_out);
    for (const VLeaf &v : plan.vleaves) {
        cuts.push_back(v.row_off);
        cuts.push_back(v.row_off + v.m);
    }
    std::sort(cuts.begin(), cuts.end());
    cuts.erase(std::unique(cuts.begin(), cuts.end()), cuts.end());
    bin_start.clear();
    size_t i = 0;
    while (i + 1 < cuts.size()) {
        int s = cuts[i];
        size_t j = i + 1;
        while (j + 1 < cuts.size() && cuts[j + 1] - s <= bin_rows)
            j++;
        int e = cuts[j];
        if (e - s <= bin_rows) {
            bin_start.push_back(s);
        } else { // one cluster larger than a bin: equal split
            int parts = (e - s + bin_rows - 1) / bin_rows;
            for (int p = 0; p < parts; p++)
                bin_start.push_back(s + (int)((int64_t)(e - s) * p / parts));
        }
        i = j;
    }
    bin_start.push_back(n_out);
    if (n_out == 0) {
        bin_start.assign(1, 0);
    }
}

// element range [s, s+cnt) of xz -> 16-byte aligned bulk copy; returns the element offset of `s` inside the copy
static inline void aligned_span(int es, uint32_t s, uint32_t cnt, uint32_t &src_al, uint32_t &bytes, uint32_t &lead) {
    if (es == 16) {
        src_al = s; bytes = cnt * 16; lead = 0;
    } else {
        src_al = s & ~1u;
        uint32_t e_al = (s + cnt + 1) & ~1u;
        bytes = (e_al - src_al) * 8;
        lead  = s - src_al;
    }
}

struct StageBuilder {
    PhasePlan &ph;
    const Options &opt;
    int es;
    // current stage
    uint32_t first_piece = 0, n_pieces = 0, data_bytes = 0, z_bytes = 0, z_first = 0;
    std::vector<uint32_t> tile_bytes; // of the pieces of the current stage
    uint32_t piece_base = 0;
    uint32_t row0 = 0; uint16_t bin_rows = 0, flags = 0;

    StageBuilder(PhasePlan &p, const Options &o, int e) : ph(p), opt(o), es(e), stage_cap((uint32_t)o.stage_bytes), z_cap((uint32_t)o.stage_z_bytes) {}

    uint32_t stage_cap = 0; // packet bytes (opt.stage_bytes)
    uint32_t z_cap     = 0; // operand bytes a packet may reference
    bool fits(uint32_t tile, uint32_t zb) const {
        uint32_t total = 16 + 16 * (n_pieces + 1) + data_bytes + align_up(tile, 16);
        return total <= stage_cap && z_bytes + zb <= z_cap && n_pieces < 255;
    }
    void begin(uint32_t piece_index) {
        first_piece = piece_index; n_pieces = 0; data_bytes = 0; z_bytes = 0; z_first = (uint32_t)ph.zcopies.size();
        tile_bytes.clear();
    }
    // adds the next piece; its operand is the xz element span [z_src, z_src + z_cnt).  Operand copies are the expensive
    // part of a stage for the TMA unit (each small cp.async.bulk has a fixed cost), so a span is (1) served from a copy
    // this stage already makes if it lies inside one, (2) appended to the last copy if it overlaps / touches its end,
    // and only otherwise (3) becomes a new copy.  The planner orders pieces so that (1) and (2) are the common case.
    void add(uint32_t tile, uint32_t z_src, uint32_t z_cnt) {
        uint32_t src_al, zb, lead;
        aligned_span(es, z_src, z_cnt, src_al, zb, lead);
        const uint32_t end_al = src_al + zb / es; // element index one past the aligned span
        uint32_t zoff = UINT32_MAX;
        for (size_t c = ph.zcopies.size(); c-- > z_first;) { // (1) contained in an earlier copy of this stage
            const ZCopy &zc = ph.zcopies[c];
            if (src_al >= zc.src_elem && end_al <= zc.src_elem + zc.bytes / (uint32_t)es) {
                zoff = zc.dst_off / es + (z_src - zc.src_elem);
                break;
            }
        }
        if (zoff == UINT32_MAX && ph.zcopies.size() > z_first) { // (2) extend the last copy
            ZCopy &zc = ph.zcopies.back();
            const uint32_t last_end = zc.src_elem + zc.bytes / (uint32_t)es;
            if (src_al >= zc.src_elem && src_al <= last_end) {
                const uint32_t grow = (end_al - last_end) * es;
                if ((uint32_t)zc.bytes + grow <= 65520u) {
                    zoff = zc.dst_off / es + (z_src - zc.src_elem);
                    zc.bytes = (uint16_t)(zc.bytes + grow);
                    z_bytes += grow;
                }
            }
        }
        if (zoff == UINT32_MAX) { // (3) new copy
            ZCopy zc; zc.src_elem = src_al; zc.bytes = (uint16_t)zb; zc.dst_off = (uint16_t)z_bytes;
            ph.zcopies.push_back(zc);
            zoff = z_bytes / es + lead;
            z_bytes += zb;
        }
        ph.piece_zoff.push_back((uint16_t)zoff);
        tile_bytes.push_back(align_up(tile, 16));
        data_bytes += align_up(tile, 16);
        n_pieces++;
    }
    static uint32_t z_need(int es, uint32_t z_src, uint32_t z_cnt) {
        uint32_t a, b, c; aligned_span(es, z_src, z_cnt, a, b, c); return b;
    }
    void close() {
        StageEntry se{};
        uint32_t hdr = 16 + 16 * n_pieces;
        se.gmem_off = ph.stream_bytes;
        se.bytes    = hdr + data_bytes;
        se.z_first  = z_first;
        se.n_z      = (uint32_t)ph.zcopies.size() - z_first;
        se.z_bytes  = z_bytes;
        ph.stages.push_back(se);
        uint32_t off = hdr;
        for (uint32_t tb : tile_bytes) { ph.piece_dataoff.push_back(off); off += tb; }
        StageHost sh{first_piece, n_pieces, piece_base, row0, bin_rows, flags};
        ph.stage_host.push_back(sh);
        ph.stream_bytes = align_up64(ph.stream_bytes + se.bytes, 128);
        piece_base += n_pieces;
    }
};

// ------------------------------------------------------------------------------------------------------------
// multi-RHS panel streams (plan.h "mm streams"; reference: the leaf loop of
// hmatrix/linalg/add_hmatrix_matrix_product_row_major.hpp:57-178 with the two gemms of
// lrmat/linalg/add_lrmat_matrix_product_row_major.hpp:11-27 and the dense gemm of
// matrix/linalg/add_matrix_matrix_product_row_major.hpp:23-46)
// ------------------------------------------------------------------------------------------------------------
namespace {

struct MmBuilder {
    MmPhasePlan &ph;
    int es, kg_cols, kg_per_packet, panel_rows_max;
    uint32_t zero_row;
    // current unit
    uint32_t unit_first = 0;
    // current packet
    uint32_t first_piece = 0, z_first = 0, kfill = 0, z_rows = 0;
    uint16_t rows = 0, flags = 0;
    uint32_t out0 = 0;

    MmBuilder(MmPhasePlan &p, int e, int kgc, int kgpp, int prm, uint32_t zr) : ph(p), es(e), kg_cols(kgc), kg_per_packet(kgpp), panel_rows_max(prm), zero_row(zr) {}
    uint32_t packet_cols() const { return (uint32_t)kg_cols * kg_per_packet; }

    void begin_unit(uint16_t panel_rows, uint32_t out_row0) {
        unit_first = (uint32_t)ph.stages.size();
        rows = panel_rows; out0 = out_row0;
        flags = (uint16_t)kMmFlagFirst;
        begin_packet();
    }
    void begin_packet() {
        first_piece = (uint32_t)ph.pieces.size();
        z_first = (uint32_t)ph.zcopies.size();
        kfill = 0; z_rows = 0;
    }
    void add_rows(uint32_t src_row, uint32_t n) { // operand rows for the next n columns of the packet
        if (ph.zcopies.size() > z_first) {
            ZCopy &zc = ph.zcopies.back();
            if (zc.src_elem + zc.bytes / (uint32_t)es == src_row && (uint32_t)zc.bytes + n * es <= 65520u) { // contiguous in the workspace: extend
                zc.bytes = (uint16_t)(zc.bytes + n * es);
                z_rows += n;
                return;
            }
        }
        ZCopy zc; zc.src_elem = src_row; zc.bytes = (uint16_t)(n * es); zc.dst_off = (uint16_t)(z_rows * es);
        ph.zcopies.push_back(zc);
        z_rows += n;
    }
    // appends columns [c0, c0+nc) of a role matrix (rows [r0, r0+nr) at panel row ro), whose operand rows start at src_row
    void add_columns(int32_t vleaf, int32_t r0, int32_t nr, int32_t c0, int32_t nc, int32_t ro, uint32_t src_row) {
        while (nc > 0) {
            if (kfill == packet_cols()) { close_packet(false, UINT32_MAX); begin_packet(); }
            const int32_t take = std::min<int32_t>(nc, (int32_t)(packet_cols() - kfill));
            ph.pieces.push_back(MmPiece{vleaf, r0, nr, c0, take, ro, (int32_t)kfill});
            add_rows(src_row, (uint32_t)take);
            kfill += take; c0 += take; nc -= take; src_row += take;
        }
    }
    void close_packet(bool last, uint32_t map_first) {
        const uint32_t n_kg = (kfill + kg_cols - 1) / kg_cols;
        if (n_kg * kg_cols > kfill) add_rows(zero_row, n_kg * kg_cols - kfill); // K padding multiplies zero rows (never NaNs of a neighbour)
        if (last) flags |= (uint16_t)kMmFlagLast;
        const uint32_t mt = ((uint32_t)rows + 7) / 8;
        uint32_t bytes = 16 + n_kg * mt * 512;
        if (map_first != UINT32_MAX) bytes += align_up((uint32_t)rows * 4, 16);
        StageEntry se{};
        se.gmem_off = ph.stream_bytes; se.bytes = bytes; se.z_first = z_first; se.n_z = (uint32_t)ph.zcopies.size() - z_first; se.z_bytes = z_rows * es;
        ph.stages.push_back(se);
        ph.packets.push_back(MmPacketHost{first_piece, (uint32_t)ph.pieces.size() - first_piece, n_kg, out0, rows, flags, map_first});
        ph.stream_bytes = align_up64(ph.stream_bytes + bytes, 128);
        ph.max_packet_bytes = std::max(ph.max_packet_bytes, bytes);
        ph.max_z_rows = std::max(ph.max_z_rows, z_rows);
        flags = 0;
    }
    void end_unit(uint32_t map_first) {
        close_packet(true, map_first);
        ph.units.push_back(WorkUnit{unit_first, (uint32_t)ph.stages.size() - unit_first});
    }
};

} // namespace

static void build_mm_streams(const Options &opt, Plan &plan) {
    const int es = plan.elem_bytes, kgc = es == 8 ? 8 : 4;
    const size_t nv = plan.vleaves.size();
    // ---- workspace rows: x | T (tight, view order) | partials | zero rows ------------------------------------------------
    plan.mm_t_row.assign(nv, 0);
    uint64_t row = (uint64_t)plan.n_in;
    for (size_t vi = 0; vi < nv; vi++)
        if (plan.vleaves[vi].r > 0) { plan.mm_t_row[vi] = (uint32_t)row; row += (uint64_t)plan.vleaves[vi].r; }
    const uint64_t t_end = row;

    // ---- phase 1: panels of V rows sharing one source cluster, split into column chunks ----------------------------------------
    struct Entry { uint32_t vleaf; int32_t l0, lr, ro; };
    struct Panel { int32_t col_off, n; uint32_t first_entry, n_entries; int32_t rows; };
    std::vector<Entry> entries;
    std::vector<Panel> panels;
    {
        std::vector<uint32_t> order;
        for (size_t vi = 0; vi < nv; vi++)
            if (plan.vleaves[vi].r > 0) order.push_back((uint32_t)vi);
        std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) {
            const VLeaf &x = plan.vleaves[a], &y = plan.vleaves[b];
            if (x.col_off != y.col_off) return x.col_off < y.col_off;
            return x.n > y.n;
        });
        for (size_t i = 0; i < order.size();) {
            const VLeaf &v0 = plan.vleaves[order[i]];
            size_t j = i;
            Panel cur{v0.col_off, v0.n, (uint32_t)entries.size(), 0, 0};
            for (; j < order.size() && plan.vleaves[order[j]].col_off == v0.col_off && plan.vleaves[order[j]].n == v0.n; j++) {
                const VLeaf &v = plan.vleaves[order[j]];
                int l0 = 0;
                while (l0 < v.r) { // a view may straddle two panels: panels are filled to kMmRows1 rows
                    if (cur.rows == kMmRows1) {
                        panels.push_back(cur);
                        cur = Panel{v0.col_off, v0.n, (uint32_t)entries.size(), 0, 0};
                    }
                    const int take = std::min(v.r - l0, kMmRows1 - cur.rows);
                    entries.push_back(Entry{order[j], l0, take, cur.rows});
                    cur.n_entries++; cur.rows += take; l0 += take;
                }
            }
            if (cur.rows > 0) panels.push_back(cur);
            i = j;
        }
    }
    struct Task { uint32_t panel; int32_t k0, nk; int32_t chunk, n_chunks; uint32_t out_row; };
    std::vector<Task> tasks;
    std::vector<std::vector<uint32_t>> panel_part_rows(panels.size());
    for (size_t pi = 0; pi < panels.size(); pi++) {
        const Panel &pn = panels[pi];
        const int nch = (pn.n + opt.mm_task_cols - 1) / opt.mm_task_cols;
        for (int c = 0; c < nch; c++) {
            const int k0 = (int)((int64_t)pn.n * c / nch), k1 = (int)((int64_t)pn.n * (c + 1) / nch);
            uint32_t out_row = 0;
            if (nch > 1) { out_row = (uint32_t)row; panel_part_rows[pi].push_back(out_row); row += (uint64_t)pn.rows; }
            tasks.push_back(Task{(uint32_t)pi, k0, k1 - k0, c, nch, out_row});
        }
        if (nch > 1)
            for (uint32_t e = pn.first_entry; e < pn.first_entry + pn.n_entries; e++) {
                const Entry &en = entries[e];
                ReduceSeg rs{plan.mm_t_row[en.vleaf] + (uint32_t)en.l0, (uint32_t)plan.reduce_mm_rows.size(), (uint32_t)nch, (uint32_t)en.lr};
                for (uint32_t base : panel_part_rows[pi]) plan.reduce_mm_rows.push_back(base + (uint32_t)en.ro);
                plan.reduce_mm.push_back(rs);
            }
    }
    plan.mm_zero_row = (uint32_t)row;
    row += kMmZeroRows;
    if (row >= (1ull << 32)) throw std::runtime_error("multi-RHS workspace exceeds 2^32 rows; split the operator into slabs");
    plan.mm_rows = row;
    (void)t_end;
    // units are visited by the first source column they read: concurrently running CTAs then share the rows of X in L2
    std::stable_sort(tasks.begin(), tasks.end(), [&](const Task &a, const Task &b) {
        return panels[a.panel].col_off + a.k0 < panels[b.panel].col_off + b.k0;
    });
    {
        MmBuilder mb(plan.mm1, es, kgc, opt.mm1_kg, kMmRows1, plan.mm_zero_row);
        for (const Task &t : tasks) {
            const Panel &pn = panels[t.panel];
            mb.begin_unit((uint16_t)pn.rows, 0);
            // the packet columns of a unit are the unit's source columns in order; every panel entry contributes its rows
            for (int k = 0; k < t.nk;) {
                if (mb.kfill == mb.packet_cols()) { mb.close_packet(false, UINT32_MAX); mb.begin_packet(); }
                const int take = std::min<int>(t.nk - k, (int)(mb.packet_cols() - mb.kfill));
                for (uint32_t e = pn.first_entry; e < pn.first_entry + pn.n_entries; e++) {
                    const Entry &en = entries[e];
                    plan.mm1.pieces.push_back(MmPiece{(int32_t)en.vleaf, en.l0, en.lr, t.k0 + k, take, en.ro, (int32_t)mb.kfill});
                }
                mb.add_rows((uint32_t)(pn.col_off + t.k0 + k), (uint32_t)take);
                mb.kfill += take;
                k += take;
            }
            const uint32_t map_first = (uint32_t)plan.mm_row_map.size();
            for (uint32_t e = pn.first_entry; e < pn.first_entry + pn.n_entries; e++) {
                const Entry &en = entries[e];
                for (int l = 0; l < en.lr; l++)
                    plan.mm_row_map.push_back(t.n_chunks > 1 ? t.out_row + (uint32_t)(en.ro + l) : plan.mm_t_row[en.vleaf] + (uint32_t)(en.l0 + l));
            }
            mb.end_unit(map_first);
        }
    }
    // ---- phase 2: one unit per target bin, its views concatenated along K --------------------------------------------------
    {
        MmBuilder mb(plan.mm2, es, kgc, opt.mm2_kg, kMmRows2, plan.mm_zero_row);
        const int n_bins = (int)plan.bin_start.size() - 1;
        auto bin_of = [&](int r) { return (int)(std::upper_bound(plan.bin_start.begin(), plan.bin_start.end(), r) - plan.bin_start.begin()) - 1; };
        // views per bin, in view order (views are sorted by target range, so the T rows of a bin's views form few runs)
        std::vector<uint64_t> first(n_bins + 1, 0);
        for (size_t vi = 0; vi < nv; vi++) {
            const VLeaf &v = plan.vleaves[vi];
            for (int b = bin_of(v.row_off); b <= bin_of(v.row_off + v.m - 1); b++) first[b + 1]++;
        }
        for (int b = 0; b < n_bins; b++) first[b + 1] += first[b];
        std::vector<uint32_t> bin_views(first[n_bins]);
        {
            std::vector<uint64_t> fill(first.begin(), first.end() - 1);
            for (size_t vi = 0; vi < nv; vi++) {
                const VLeaf &v = plan.vleaves[vi];
                for (int b = bin_of(v.row_off); b <= bin_of(v.row_off + v.m - 1); b++) bin_views[fill[b]++] = (uint32_t)vi;
            }
        }
        for (int b = 0; b < n_bins; b++) {
            const int bs = plan.bin_start[b], be = plan.bin_start[b + 1];
            mb.begin_unit((uint16_t)(be - bs), (uint32_t)bs);
            for (uint64_t q = first[b]; q < first[b + 1]; q++) {
                const uint32_t vi = bin_views[q];
                const VLeaf &v = plan.vleaves[vi];
                const int rs = std::max(v.row_off, bs), re = std::min(v.row_off + v.m, be);
                if (re <= rs) continue;
                const int K = v.r < 0 ? v.n : v.r;
                const uint32_t src = v.r < 0 ? (uint32_t)v.col_off : plan.mm_t_row[vi];
                mb.add_columns((int32_t)vi, rs - v.row_off, re - rs, 0, K, rs - bs, src);
            }
            mb.end_unit(UINT32_MAX); // a bin without views still gets one (empty) packet: beta must be applied
        }
    }
}

// ------------------------------------------------------------------------------------------------------------
// the plan
// ------------------------------------------------------------------------------------------------------------
void build_plan(const hmb200_flat &f, int es, char op, const Options &opt, Plan &plan) {
    if (es != 8 && es != 16) throw std::runtime_error("elem_bytes must be 8 or 16");
    if (op != 'N' && op != 'T' && op != 'C') throw std::runtime_error("trans must be N, T or C");
    plan            = Plan();
    plan.op         = op;
    plan.elem_bytes = es;
    plan.n_out      = (op == 'N') ? f.n_rows : f.n_cols;
    plan.n_in       = (op == 'N') ? f.n_cols : f.n_rows;
    make_vleaves(f, op, plan);
    make_bins(plan, plan.n_out, opt.bin_rows, plan.bin_start);
    const int n_bins = (int)plan.bin_start.size() - 1;

    // ---- t slots of the low-rank views ------------------------------------------------------------------
    uint64_t cur = align_up64((uint64_t)plan.n_in + 1, 2); // x, padded so aligned over-reads stay inside
    plan.t_begin = (uint32_t)cur;
    for (VLeaf &v : plan.vleaves)
        if (v.r > 0) {
            v.t_off = (uint32_t)cur;
            cur += align_up64((uint64_t)v.r, 2);
        }

    // ---- phase 1 pieces: V-role tiles --------------------------------------------------------------------
    const uint32_t zmax_elems = (uint32_t)(opt.stage_z_bytes / es) - 2;
    // phase-1 tiles are visited by input range (source cluster): V tiles of different leaves that multiply the same slice
    // of x land in the same stage and share one operand copy
    std::vector<uint32_t> order1;
    order1.reserve(plan.vleaves.size());
    for (size_t vi = 0; vi < plan.vleaves.size(); vi++)
        if (plan.vleaves[vi].r > 0) order1.push_back((uint32_t)vi);
    std::stable_sort(order1.begin(), order1.end(), [&](uint32_t a, uint32_t b) {
        const VLeaf &x = plan.vleaves[a], &y = plan.vleaves[b];
        if (x.col_off != y.col_off) return x.col_off < y.col_off;
        if (x.n != y.n) return x.n > y.n;
        return x.row_off < y.row_off;
    });
    for (uint32_t vi : order1) {
        const VLeaf &v = plan.vleaves[vi];
        int nl = (v.r + kMaxPhase1Rows - 1) / kMaxPhase1Rows;
        for (int li = 0; li < nl; li++) {
            int l0 = (int)((int64_t)v.r * li / nl), l1 = (int)((int64_t)v.r * (li + 1) / nl);
            int lr = l1 - l0;
            uint32_t nq_max = std::max<uint32_t>(1, (uint32_t)opt.piece1_bytes / (uint32_t)(lr * es));
            nq_max          = std::min<uint32_t>(nq_max, std::min<uint32_t>(zmax_elems, 65535u));
            int nk          = (int)((v.n + nq_max - 1) / nq_max);
            uint32_t first_partial = 0;
            if (nk > 1) {
                first_partial = (uint32_t)cur;
                ReduceSeg rs{v.t_off + (uint32_t)l0, first_partial, (uint32_t)nk, (uint32_t)lr};
                plan.reduce.push_back(rs);
                cur += (uint64_t)nk * lr;
            }
            for (int ki = 0; ki < nk; ki++) {
                int k0 = (int)((int64_t)v.n * ki / nk), k1 = (int)((int64_t)v.n * (ki + 1) / nk);
                Piece1 p{(int32_t)vi, l0, lr, k0, k1 - k0, nk > 1 ? first_partial + (uint32_t)ki * lr : v.t_off + (uint32_t)l0};
                plan.pieces1.push_back(p);
            }
        }
    }
    cur += 2;
    if (cur >= (1ull << 32)) throw std::runtime_error("workspace exceeds 2^32 elements; split the operator into slabs");
    plan.xz_elems = cur;

    // ---- phase 2 pieces: U-role / dense tiles, sorted by target bin ------------------------------------------
    auto bin_of = [&](int row) { // bin containing `row`
        return (int)(std::upper_bound(plan.bin_start.begin(), plan.bin_start.end(), row) - plan.bin_start.begin()) - 1;
    };
    auto for_each_piece2 = [&](auto &&emit) {
        for (size_t vi = 0; vi < plan.vleaves.size(); vi++) {
            const VLeaf &v = plan.vleaves[vi];
            int K          = v.r < 0 ? v.n : v.r;
            uint32_t zbase = v.r < 0 ? (uint32_t)v.col_off : v.t_off;
            int b0 = bin_of(v.row_off), b1 = bin_of(v.row_off + v.m - 1);
            for (int b = b0; b <= b1; b++) {
                int rs = std::max(v.row_off, plan.bin_start[b]), re = std::min(v.row_off + v.m, plan.bin_start[b + 1]);
                int nr = re - rs;
                if (nr <= 0) continue;
                uint32_t nc_max = std::max<uint32_t>(1, (uint32_t)opt.piece2_bytes / (uint32_t)(nr * es));
                nc_max          = std::min<uint32_t>(nc_max, std::min<uint32_t>(zmax_elems, 65535u));
                int nch         = (int)((K + nc_max - 1) / nc_max);
                for (int c = 0; c < nch; c++) {
                    int c0 = (int)((int64_t)K * c / nch), c1 = (int)((int64_t)K * (c + 1) / nch);
                    Piece2 p{(int32_t)vi, rs - v.row_off, nr, c0, c1 - c0, rs - plan.bin_start[b], zbase + (uint32_t)c0};
                    emit(b, p);
                }
            }
        }
    };
    std::vector<uint64_t> bin_first(n_bins + 1, 0);
    for_each_piece2([&](int b, const Piece2 &) { bin_first[b + 1]++; });
    for (int b = 0; b < n_bins; b++) bin_first[b + 1] += bin_first[b];
    plan.pieces2.resize(bin_first[n_bins]);
    {
        std::vector<uint64_t> fill(bin_first.begin(), bin_first.end() - 1);
        for_each_piece2([&](int b, const Piece2 &p) { plan.pieces2[fill[b]++] = p; });
    }

    // ---- phase 1 stages / units -------------------------------------------------------------------------------
    {
        PhasePlan &ph = plan.phase1;
        StageBuilder sb(ph, opt, es);
        sb.begin(0);
        uint32_t unit_first = 0;
        auto close_stage = [&](uint32_t next_piece) {
            sb.close();
            if ((uint32_t)ph.stages.size() - unit_first >= (uint32_t)opt.unit_stages) {
                ph.units.push_back(WorkUnit{unit_first, (uint32_t)ph.stages.size() - unit_first});
                unit_first    = (uint32_t)ph.stages.size();
                sb.piece_base = 0;
            }
            sb.begin(next_piece);
        };
        for (size_t pi = 0; pi < plan.pieces1.size(); pi++) {
            const Piece1 &p = plan.pieces1[pi];
            const VLeaf &v  = plan.vleaves[p.vleaf];
            uint32_t tile   = (uint32_t)p.lr * p.nq * es;
            uint32_t zsrc   = (uint32_t)(v.col_off + p.k0);
            if (sb.n_pieces && !sb.fits(tile, StageBuilder::z_need(es, zsrc, p.nq))) close_stage((uint32_t)pi);
            sb.add(tile, zsrc, p.nq);
        }
        if (sb.n_pieces) sb.close();
        if ((uint32_t)ph.stages.size() > unit_first)
            ph.units.push_back(WorkUnit{unit_first, (uint32_t)ph.stages.size() - unit_first});
    }
    // ---- phase 2 stages / units (one unit per bin; a packet never spans bins) --------------------------------
    {
        PhasePlan &ph = plan.phase2;
        StageBuilder sb(ph, opt, es);
        for (int b = 0; b < n_bins; b++) {
            uint32_t unit_first = (uint32_t)ph.stages.size();
            sb.piece_base       = 0;
            sb.row0             = (uint32_t)plan.bin_start[b];
            sb.bin_rows         = (uint16_t)(plan.bin_start[b + 1] - plan.bin_start[b]);
            sb.flags            = kFlagFirstOfBin;
            sb.begin((uint32_t)bin_first[b]);
            for (uint64_t pi = bin_first[b]; pi < bin_first[b + 1]; pi++) {
                const Piece2 &p = plan.pieces2[pi];
                uint32_t tile   = (uint32_t)p.nr * p.nc * es;
                if (sb.n_pieces && !sb.fits(tile, StageBuilder::z_need(es, p.z_src, p.nc))) {
                    sb.close();
                    sb.flags = 0;
                    sb.begin((uint32_t)pi);
                }
                sb.add(tile, p.z_src, p.nc);
            }
            sb.flags |= kFlagLastOfBin; // a bin without pieces still gets one (empty) packet: beta must be applied
            sb.close();
            ph.units.push_back(WorkUnit{unit_first, (uint32_t)ph.stages.size() - unit_first});
        }
    }
    if (opt.build_matmat) build_mm_streams(opt, plan);
}

// ------------------------------------------------------------------------------------------------------------
// packing
// ------------------------------------------------------------------------------------------------------------
template <typename T> static inline T conj_if(T v, bool) { return v; }
template <> inline std::complex<double> conj_if(std::complex<double> v, bool c) { return c ? std::conj(v) : v; }
template <typename T> static inline T real_part(T v) { return v; }
template <> inline std::complex<double> real_part(std::complex<double> v) { return std::complex<double>(v.real(), 0.); }

// element (i, j) of a dense leaf stored with symv / hemv semantics: only the `uplo` triangle is referenced
// (matrix/linalg/add_matrix_vector_product.hpp:26-52)
template <typename T>
static inline T sym_elem(const T *D, int n, int kind, int i, int j) {
    const bool lower = (kind == 1 || kind == 3), herm = (kind >= 3);
    const bool in_tri = lower ? (i >= j) : (i <= j);
    T v = in_tri ? D[(size_t)j * n + i] : conj_if(D[(size_t)i * n + j], herm);
    if (herm && i == j) v = real_part(v);
    return v;
}

template <typename T>
static void fill_tile2(const hmb200_flat &f, const VLeaf &v, const Piece2 &p, T *dst, int ld = 0) {
    const int64_t l = v.leaf;
    const int lm = f.nrows[l];
    if (ld <= 0) ld = p.nr;
    if (v.r < 0) {
        const T *D = (const T *)f.a[l];
        if (v.kind != 0) {
            for (int j = 0; j < p.nc; j++)
                for (int i = 0; i < p.nr; i++)
                    dst[(size_t)j * ld + i] = sym_elem(D, lm, v.kind, p.r0 + i, p.c0 + j);
        } else if (!v.transposed) {
            for (int j = 0; j < p.nc; j++)
                std::memcpy(dst + (size_t)j * ld, D + (size_t)(p.c0 + j) * lm + p.r0, sizeof(T) * p.nr);
        } else { // view = D^T (or D^H): view(i, j) = D(j, i)
            for (int j = 0; j < p.nc; j++)
                for (int i = 0; i < p.nr; i++)
                    dst[(size_t)j * ld + i] = conj_if(D[(size_t)(p.r0 + i) * lm + (p.c0 + j)], v.conj);
        }
    } else if (!v.transposed) { // U: lm x r
        const T *U = (const T *)f.a[l];
        for (int j = 0; j < p.nc; j++)
            std::memcpy(dst + (size_t)j * ld, U + (size_t)(p.c0 + j) * lm + p.r0, sizeof(T) * p.nr);
    } else { // (U V)^T = V^T U^T : the U-role is V^T (n_leaf x r), V is r x n_leaf with ld r
        const T *V = (const T *)f.b[l];
        const int r = v.r;
        for (int j = 0; j < p.nc; j++)
            for (int i = 0; i < p.nr; i++)
                dst[(size_t)j * ld + i] = conj_if(V[(size_t)(p.r0 + i) * r + (p.c0 + j)], v.conj);
    }
}

template <typename T>
static void fill_tile1(const hmb200_flat &f, const VLeaf &v, const Piece1 &p, T *dst, int ld) {
    const int64_t l = v.leaf;
    const int r = v.r;
    if (!v.transposed) { // V: r x n, ld r
        const T *V = (const T *)f.b[l];
        if (p.lr == r && p.l0 == 0 && ld == r) {
            std::memcpy(dst, V + (size_t)p.k0 * r, sizeof(T) * (size_t)r * p.nq);
        } else {
            for (int k = 0; k < p.nq; k++)
                std::memcpy(dst + (size_t)k * ld, V + (size_t)(p.k0 + k) * r + p.l0, sizeof(T) * p.lr);
        }
    } else { // V-role is U^T (r x m_leaf): view(l, k) = U(k, l)
        const T *U = (const T *)f.a[l];
        const int lm = f.nrows[l];
        for (int k = 0; k < p.nq; k++)
            for (int ll = 0; ll < p.lr; ll++)
                dst[(size_t)k * ld + ll] = conj_if(U[(size_t)(p.l0 + ll) * lm + (p.k0 + k)], v.conj);
    }
}

template <typename T>
static void pack_stages_t(const hmb200_flat &f, const Plan &plan, int phase, size_t s0, size_t s1, uint8_t *dst) {
    const PhasePlan &ph = phase == 1 ? plan.phase1 : plan.phase2;
    const uint64_t base = ph.stages[s0].gmem_off;
#pragma omp parallel for schedule(dynamic, 16)
    for (int64_t s = (int64_t)s0; s < (int64_t)s1; s++) {
        const StageEntry &se = ph.stages[s];
        const StageHost &sh  = ph.stage_host[s];
        uint8_t *pk          = dst + (se.gmem_off - base);
        PacketHeader hdr;
        hdr.n_pieces   = sh.n_pieces;
        hdr.piece_base = sh.piece_base;
        hdr.row0       = sh.row0;
        hdr.bin_rows   = sh.bin_rows;
        hdr.flags      = sh.flags;
        std::memcpy(pk, &hdr, 16);
        for (uint32_t i = 0; i < sh.n_pieces; i++) {
            const uint32_t pi = sh.first_piece + i;
            PieceDesc d{};
            d.data_off = ph.piece_dataoff[pi];
            d.z_off    = ph.piece_zoff[pi];
            T *tile    = (T *)(pk + d.data_off);
            if (phase == 1) {
                const Piece1 &p = plan.pieces1[pi];
                d.out_off = p.out_off; d.nc = (uint16_t)p.nq; d.nr = (uint8_t)p.lr;
                d.ro  = (uint8_t)(32 / p.lr);                      // column groups G
                d.pad = (uint16_t)((32768 + p.lr - 1) / p.lr);     // lane / lr == (lane * pad) >> 15 for lane < 32, lr <= 32
                fill_tile1<T>(f, plan.vleaves[p.vleaf], p, tile, p.lr);
                if (((size_t)p.lr * p.nq * sizeof(T)) & 15) std::memset(tile + (size_t)p.lr * p.nq, 0, 8);
            } else {
                const Piece2 &p = plan.pieces2[pi];
                d.out_off = 0; d.nc = (uint16_t)p.nc; d.nr = (uint8_t)p.nr; d.ro = (uint8_t)p.ro;
                fill_tile2<T>(f, plan.vleaves[p.vleaf], p, tile);
                if (((size_t)p.nr * p.nc * sizeof(T)) & 15) std::memset(tile + (size_t)p.nr * p.nc, 0, 8);
            }
            std::memcpy(pk + 16 + 16 * (size_t)i, &d, 16);
        }
        // zero the alignment gap up to the next packet so the staging buffer never ships uninitialised bytes
        uint64_t end = se.gmem_off + se.bytes;
        uint64_t nxt = (s + 1 < (int64_t)ph.stages.size()) ? ph.stages[s + 1].gmem_off : ph.stream_bytes;
        if (nxt > end && s + 1 <= (int64_t)s1) std::memset(pk + se.bytes, 0, (size_t)std::min<uint64_t>(nxt - end, 128));
    }
}

// multi-RHS panel packets: the packet's pieces are laid into a zero-filled column-major panel, which is then written out
// fragment-major (plan.h): chunk (kg, mt), lane = 4 gid + tig  <-  A[8 mt + gid][kg_cols kg + tig] (+ 4 for the second half, real)
template <typename T>
static void pack_mm_t(const hmb200_flat &f, const Plan &plan, int phase, size_t s0, size_t s1, uint8_t *dst) {
    const MmPhasePlan &ph = phase == 3 ? plan.mm1 : plan.mm2;
    const uint64_t base   = ph.stages[s0].gmem_off;
    constexpr int KG      = sizeof(T) == 8 ? 8 : 4;
#pragma omp parallel
    {
        std::vector<T> panel;
#pragma omp for schedule(dynamic, 16)
        for (int64_t s = (int64_t)s0; s < (int64_t)s1; s++) {
            const StageEntry &se   = ph.stages[s];
            const MmPacketHost &pk = ph.packets[s];
            uint8_t *out           = dst + (se.gmem_off - base);
            const int mt_panel = (pk.rows + 7) / 8, ldp = mt_panel * 8, ncol = (int)pk.n_kg * KG;
            MmPacketHeader h{};
            h.n_kg = pk.n_kg; h.out0 = pk.out0; h.rows = pk.rows; h.mt_panel = (uint16_t)mt_panel; h.flags = pk.flags;
            const uint32_t a_bytes = pk.n_kg * (uint32_t)mt_panel * 512u;
            h.map_off16 = pk.map_first != UINT32_MAX ? (uint16_t)((16 + a_bytes) / 16) : 0;
            std::memcpy(out, &h, 16);
            panel.assign((size_t)ldp * std::max(ncol, 1), T(0));
            for (uint32_t i = 0; i < pk.n_pieces; i++) {
                const MmPiece &mp = ph.pieces[pk.first_piece + i];
                const VLeaf &v    = plan.vleaves[mp.vleaf];
                T *at             = panel.data() + (size_t)mp.kcol * ldp + mp.ro;
                if (phase == 3) fill_tile1<T>(f, v, Piece1{mp.vleaf, mp.r0, mp.nr, mp.c0, mp.nc, 0}, at, ldp);
                else fill_tile2<T>(f, v, Piece2{mp.vleaf, mp.r0, mp.nr, mp.c0, mp.nc, mp.ro, 0}, at, ldp);
            }
            T *chunks = reinterpret_cast<T *>(out + 16);
            for (uint32_t kg = 0; kg < pk.n_kg; kg++)
                for (int mt = 0; mt < mt_panel; mt++) {
                    T *c = chunks + ((size_t)kg * mt_panel + mt) * (512 / sizeof(T));
                    for (int lane = 0; lane < 32; lane++) {
                        const int gid = lane >> 2, tig = lane & 3;
                        const T *src  = panel.data() + (size_t)(kg * KG + tig) * ldp + 8 * mt + gid;
                        if (sizeof(T) == 8) {
                            c[2 * lane]     = src[0];
                            c[2 * lane + 1] = src[(size_t)4 * ldp];
                        } else {
                            c[lane] = src[0];
                        }
                    }
                }
            if (pk.map_first != UINT32_MAX) {
                uint8_t *m = out + 16 + a_bytes;
                std::memset(m, 0, align_up((uint32_t)pk.rows * 4, 16));
                std::memcpy(m, plan.mm_row_map.data() + pk.map_first, (size_t)pk.rows * 4);
            }
            uint64_t end = se.gmem_off + se.bytes;
            uint64_t nxt = (s + 1 < (int64_t)ph.stages.size()) ? ph.stages[s + 1].gmem_off : ph.stream_bytes;
            if (nxt > end && s + 1 <= (int64_t)s1) std::memset(out + se.bytes, 0, (size_t)std::min<uint64_t>(nxt - end, 128));
        }
    }
}

void pack_stages(const hmb200_flat &f, const Plan &plan, int phase, size_t s0, size_t s1, uint8_t *dst) {
    if (s1 <= s0) return;
    if (phase >= 3) {
        if (plan.elem_bytes == 8) pack_mm_t<double>(f, plan, phase, s0, s1, dst);
        else pack_mm_t<std::complex<double>>(f, plan, phase, s0, s1, dst);
        return;
    }
    if (plan.elem_bytes == 8)
        pack_stages_t<double>(f, plan, phase, s0, s1, dst);
    else
        pack_stages_t<std::complex<double>>(f, plan, phase, s0, s1, dst);
}

} // namespace hmb200
