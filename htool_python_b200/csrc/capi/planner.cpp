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
        if (o->stage_z_elems > 0) r.stage_z_elems = o->stage_z_elems;
        if (o->mm_stage_bytes > 0) r.mm_stage_bytes = o->mm_stage_bytes;
        if (o->mm_bin_cols > 0) r.mm_bin_cols = o->mm_bin_cols;
        if (o->mm_variant > 0) r.mm_variant = o->mm_variant - 1;
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
    r.stage_z_elems  = std::min(std::max(r.stage_z_elems, 8), 4096);
    r.mm_stage_bytes = (int)align_up((uint32_t)std::max(r.mm_stage_bytes, 2048), 128);
    r.mm_bin_cols    = std::min(std::max(r.mm_bin_cols, 4), 4096);
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

    uint32_t stage_cap = 0; // packet bytes (opt.stage_bytes, or opt.mm_stage_bytes for the multi-RHS phase-1 stream)
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
    // piece whose operand is already resident in shared memory (multi-RHS phase 1): only its row offset is recorded
    void add_resident(uint32_t tile, uint32_t z_row) {
        ph.piece_zoff.push_back((uint16_t)z_row);
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
// multi-RHS phase 1: T_j (r x mu) = V_j (r x n) X[cols_j] (n x mu), V tiles grouped by source bin
// (reference: the first gemm of hmatrix/lrmat/linalg/add_lrmat_matrix_product_row_major.hpp:11-27)
// ------------------------------------------------------------------------------------------------------------
static void build_mm_phase1(const Options &opt, Plan &plan) {
    const int es = plan.elem_bytes;
    // source bins: <= mm_bin_cols input entries, cut at the leaves' source-cluster boundaries where possible
    {
        std::vector<int32_t> cuts{0, plan.n_in};
        for (const VLeaf &v : plan.vleaves)
            if (v.r > 0) { cuts.push_back(v.col_off); cuts.push_back(v.col_off + v.n); }
        std::sort(cuts.begin(), cuts.end());
        cuts.erase(std::unique(cuts.begin(), cuts.end()), cuts.end());
        auto &bs = plan.src_bin_start;
        bs.clear();
        size_t i = 0;
        while (i + 1 < cuts.size()) {
            int s0 = cuts[i];
            size_t j = i + 1;
            while (j + 1 < cuts.size() && cuts[j + 1] - s0 <= opt.mm_bin_cols) j++;
            int e0 = cuts[j];
            if (e0 - s0 <= opt.mm_bin_cols) bs.push_back(s0);
            else {
                int parts = (e0 - s0 + opt.mm_bin_cols - 1) / opt.mm_bin_cols;
                for (int q = 0; q < parts; q++) bs.push_back(s0 + (int)((int64_t)(e0 - s0) * q / parts));
            }
            i = j;
        }
        bs.push_back(plan.n_in);
        if (plan.n_in == 0) bs.assign(1, 0);
    }
    const auto &bs   = plan.src_bin_start;
    const int n_sbin = (int)bs.size() - 1;
    auto sbin_of = [&](int col) { return (int)(std::upper_bound(bs.begin(), bs.end(), col) - bs.begin()) - 1; };
    const uint32_t piece_cap = (uint32_t)(opt.mm_stage_bytes - 16 - 16 - 64);
    // rows a tile of nq columns may have: a multiple of 16 (one DMMA row pair per warp item) that fits a packet
    auto rows_cap = [&](int nq) { return std::min(128, std::max(8, (int)(piece_cap / ((uint32_t)nq * es)) / 16 * 16)); };

    // ---- sub-pieces: (leaf rows l0..l0+lr) x (columns of one source bin, possibly sub-chunked) -----------------------
    struct Sub { int b; int k_abs; Piece1 p; uint32_t seg; };
    std::vector<Sub> subs;
    for (size_t vi = 0; vi < plan.vleaves.size(); vi++) {
        const VLeaf &v = plan.vleaves[vi];
        if (v.r <= 0) continue;
        const int b0 = sbin_of(v.col_off), b1 = sbin_of(v.col_off + v.n - 1);
        struct Part { int b, k0, nq; };
        std::vector<Part> parts;
        for (int b = b0; b <= b1; b++) {
            int cs = std::max(v.col_off, bs[b]), ce = std::min(v.col_off + v.n, bs[b + 1]);
            if (ce <= cs) continue;
            int nq_max = std::max(4, (int)(piece_cap / (8u * es)));
            int nsub   = (ce - cs + nq_max - 1) / nq_max;
            for (int q = 0; q < nsub; q++) {
                int k0 = cs + (int)((int64_t)(ce - cs) * q / nsub), k1 = cs + (int)((int64_t)(ce - cs) * (q + 1) / nsub);
                parts.push_back(Part{b, k0 - v.col_off, k1 - k0});
            }
        }
        int nq_big = 0;
        for (const Part &pt : parts) nq_big = std::max(nq_big, pt.nq);
        const int lmax = rows_cap(nq_big);
        const int nl   = (v.r + lmax - 1) / lmax;
        for (int li = 0; li < nl; li++) {
            int l0 = (int)((int64_t)v.r * li / nl), l1 = (int)((int64_t)v.r * (li + 1) / nl), lr = l1 - l0;
            const uint32_t seg = (uint32_t)plan.reduce_mm.size();
            plan.reduce_mm.push_back(ReduceSeg{v.t_off + (uint32_t)l0, 0, 0, (uint32_t)lr});
            for (const Part &pt : parts) subs.push_back(Sub{pt.b, v.col_off + pt.k0, Piece1{(int32_t)vi, l0, lr, pt.k0, pt.nq, 0}, seg});
        }
    }
    // same (bin, column range) => neighbours: they are stacked into one tile
    std::stable_sort(subs.begin(), subs.end(), [](const Sub &x, const Sub &y) {
        if (x.b != y.b) return x.b < y.b;
        if (x.k_abs != y.k_abs) return x.k_abs < y.k_abs;
        return x.p.nq > y.p.nq;
    });
    // ---- tiles: stacks of sub-pieces sharing their columns; every tile writes a contiguous block of workspace rows ---
    uint64_t rows = plan.xz_elems;
    plan.pieces1m.resize(subs.size());
    std::vector<int> tile_bin;
    std::vector<std::vector<uint32_t>> seg_rows(plan.reduce_mm.size());
    for (size_t i = 0; i < subs.size();) {
        MmTile t{};
        t.first_sub = (uint32_t)i;
        t.nq        = (uint16_t)subs[i].p.nq;
        t.zrow      = (uint16_t)(subs[i].k_abs - bs[subs[i].b]);
        t.out_row   = (uint32_t)rows;
        const int cap = rows_cap(subs[i].p.nq);
        int r = 0;
        size_t j = i;
        while (j < subs.size() && subs[j].b == subs[i].b && subs[j].k_abs == subs[i].k_abs && subs[j].p.nq == subs[i].p.nq && (r == 0 || r + subs[j].p.lr <= cap)) {
            Piece1 p  = subs[j].p;
            p.out_off = (uint32_t)(rows + r);
            plan.pieces1m[j] = p;
            seg_rows[subs[j].seg].push_back(p.out_off);
            r += p.lr;
            j++;
        }
        t.n_sub = (uint32_t)(j - i);
        t.rows  = (uint16_t)r;
        rows += (uint64_t)r;
        plan.mm_tiles.push_back(t);
        tile_bin.push_back(subs[i].b);
        i = j;
    }
    if (rows >= (1ull << 32)) throw std::runtime_error("multi-RHS workspace exceeds 2^32 rows; split the operator into slabs");
    plan.mm_rows = rows;
    for (size_t sg = 0; sg < plan.reduce_mm.size(); sg++) {
        plan.reduce_mm[sg].first_partial = (uint32_t)plan.reduce_mm_rows.size();
        plan.reduce_mm[sg].n_parts       = (uint32_t)seg_rows[sg].size();
        plan.reduce_mm_rows.insert(plan.reduce_mm_rows.end(), seg_rows[sg].begin(), seg_rows[sg].end());
    }
    // ---- stages: one work unit per source bin; StageEntry.pad = (first x row, rows) of the bin's resident X slice ----
    PhasePlan &ph = plan.phase1m;
    StageBuilder sb(ph, opt, es);
    sb.stage_cap = (uint32_t)opt.mm_stage_bytes;
    sb.z_cap     = 0xffffffffu;
    size_t ti = 0;
    for (int b = 0; b < n_sbin; b++) {
        size_t te = ti;
        while (te < plan.mm_tiles.size() && tile_bin[te] == b) te++;
        if (te == ti) continue; // no tile reads this slice of x
        const uint32_t unit_first = (uint32_t)ph.stages.size();
        sb.piece_base = 0;
        sb.row0       = (uint32_t)bs[b];
        sb.bin_rows   = (uint16_t)(bs[b + 1] - bs[b]);
        sb.flags      = kFlagFirstOfBin;
        sb.begin((uint32_t)ti);
        for (; ti < te; ti++) {
            const MmTile &t = plan.mm_tiles[ti];
            uint32_t tile   = (uint32_t)t.rows * t.nq * es;
            if (sb.n_pieces && !sb.fits(tile, 0)) {
                sb.close();
                sb.flags = 0;
                sb.begin((uint32_t)ti);
            }
            sb.add_resident(tile, t.zrow);
        }
        sb.flags |= kFlagLastOfBin;
        sb.close();
        for (uint32_t s2 = unit_first; s2 < ph.stages.size(); s2++) {
            ph.stages[s2].pad[0] = (uint32_t)bs[b];
            ph.stages[s2].pad[1] = (uint32_t)(bs[b + 1] - bs[b]);
        }
        ph.units.push_back(WorkUnit{unit_first, (uint32_t)ph.stages.size() - unit_first});
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
                nc_max          = std::min<uint32_t>(nc_max, (uint32_t)opt.stage_z_elems - 2);
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
        sb.z_cap = std::min<uint32_t>(sb.z_cap, (uint32_t)opt.stage_z_elems * es); // the multi-RHS kernel stages operand ROWS
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
    plan.mm_rows = plan.xz_elems;
    if (opt.build_matmat) build_mm_phase1(opt, plan);
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
static void fill_tile2(const hmb200_flat &f, const VLeaf &v, const Piece2 &p, T *dst) {
    const int64_t l = v.leaf;
    const int lm = f.nrows[l];
    if (v.r < 0) {
        const T *D = (const T *)f.a[l];
        if (v.kind != 0) {
            for (int j = 0; j < p.nc; j++)
                for (int i = 0; i < p.nr; i++)
                    dst[(size_t)j * p.nr + i] = sym_elem(D, lm, v.kind, p.r0 + i, p.c0 + j);
        } else if (!v.transposed) {
            for (int j = 0; j < p.nc; j++)
                std::memcpy(dst + (size_t)j * p.nr, D + (size_t)(p.c0 + j) * lm + p.r0, sizeof(T) * p.nr);
        } else { // view = D^T (or D^H): view(i, j) = D(j, i)
            for (int j = 0; j < p.nc; j++)
                for (int i = 0; i < p.nr; i++)
                    dst[(size_t)j * p.nr + i] = conj_if(D[(size_t)(p.r0 + i) * lm + (p.c0 + j)], v.conj);
        }
    } else if (!v.transposed) { // U: lm x r
        const T *U = (const T *)f.a[l];
        for (int j = 0; j < p.nc; j++)
            std::memcpy(dst + (size_t)j * p.nr, U + (size_t)(p.c0 + j) * lm + p.r0, sizeof(T) * p.nr);
    } else { // (U V)^T = V^T U^T : the U-role is V^T (n_leaf x r), V is r x n_leaf with ld r
        const T *V = (const T *)f.b[l];
        const int r = v.r;
        for (int j = 0; j < p.nc; j++)
            for (int i = 0; i < p.nr; i++)
                dst[(size_t)j * p.nr + i] = conj_if(V[(size_t)(p.r0 + i) * r + (p.c0 + j)], v.conj);
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
    const PhasePlan &ph = phase == 1 ? plan.phase1 : (phase == 2 ? plan.phase2 : plan.phase1m);
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
            } else if (phase == 3) { // multi-RHS phase 1: a stack of sub-pieces sharing their columns, ld = total rows
                const MmTile &t = plan.mm_tiles[pi];
                d.out_off = t.out_row; d.nc = t.nq; d.nr = (uint8_t)t.rows; d.ro = 0; d.pad = 0;
                int roff = 0;
                for (uint32_t q = 0; q < t.n_sub; q++) {
                    const Piece1 &p = plan.pieces1m[t.first_sub + q];
                    fill_tile1<T>(f, plan.vleaves[p.vleaf], p, tile + roff, t.rows);
                    roff += p.lr;
                }
                if (((size_t)t.rows * t.nq * sizeof(T)) & 15) std::memset(tile + (size_t)t.rows * t.nq, 0, 8);
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

void pack_stages(const hmb200_flat &f, const Plan &plan, int phase, size_t s0, size_t s1, uint8_t *dst) {
    if (s1 <= s0) return;
    if (plan.elem_bytes == 8)
        pack_stages_t<double>(f, plan, phase, s0, s1, dst);
    else
        pack_stages_t<std::complex<double>>(f, plan, phase, s0, s1, dst);
}

} // namespace hmb200
