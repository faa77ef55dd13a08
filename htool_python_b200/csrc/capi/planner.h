// planner.h -- host-side schedule ("plan") of one packed operator and the packer that fills its streams.
// Pure host C++ (no CUDA, no htool headers): input is the flat leaf table of include/hmb200.h.
#pragma once
#include "hmb200.h"
#include "plan.h"

#include <cstdint>
#include <string>
#include <vector>

namespace hmb200 {

struct Options {
    int stage_bytes    = 49152;
    int stage_z_bytes  = 8192;
    int n_stages       = 3;
    int consumer_warps = 16;
    int piece1_bytes   = 8192;
    int piece2_bytes   = 8192;
    int bin_rows       = 64;
    int unit_stages    = 32;
    int ctas           = 0;
    int build_transpose = 0;
    // multi-RHS (DMMA) path
    int build_matmat    = 1;     // also pack the phase-1 stream of the multi-RHS kernels
    int stage_z_elems   = 112;   // max operand elements (x / t entries) a phase-2 packet may reference: the multi-RHS kernel
                                 // stages stage_z_elems x (mu + pad) operand ROWS per packet
    int mm_stage_bytes  = 40960; // packet size of the multi-RHS phase-1 stream
    int mm_bin_cols     = 128;
    int mm_variant      = 1;     // bit 0 / bit 1: register-double-buffered DMMA fragments in multi-RHS phase 1 / phase 2   // columns of a source bin (the slice of X kept resident in shared memory)
};
Options resolve_options(const hmb200_options *opt);

// a leaf as it enters the product for a given op: possibly the transposed (mirrored) view of a stored leaf
struct VLeaf {
    int64_t leaf;       // index into the flat table
    int32_t row_off, m; // output range of the view
    int32_t col_off, n; // input range of the view
    int32_t r;          // -1 dense
    uint32_t t_off;     // low-rank: element offset of its final t in xz
    uint8_t transposed, conj, kind;
};

struct Piece1 { // V-role tile rows [l0,l0+lr) x cols [k0,k0+nq)
    int32_t vleaf;
    int32_t l0, lr, k0, nq;
    uint32_t out_off;
};
struct Piece2 { // U-role / dense tile rows [r0,r0+nr) x cols [c0,c0+nc) of the view, placed at bin row ro
    int32_t vleaf;
    int32_t r0, nr, c0, nc;
    int32_t ro;
    uint32_t z_src; // xz element the tile's first column multiplies
};

struct MmTile { // multi-RHS phase-1 tile: sub-pieces [first_sub, first_sub + n_sub) of pieces1m stacked row-wise
    uint32_t first_sub, n_sub;
    uint32_t out_row; // first workspace row of its `rows` output rows
    uint16_t rows, nq, zrow;
};

struct StageHost { // planner-side description of a packet
    uint32_t first_piece, n_pieces;
    uint32_t piece_base;
    uint32_t row0;
    uint16_t bin_rows, flags;
};

struct PhasePlan {
    std::vector<StageEntry> stages;
    std::vector<StageHost> stage_host;
    std::vector<ZCopy> zcopies;
    std::vector<uint16_t> piece_zoff;   // per piece: element offset in the stage z-area
    std::vector<uint32_t> piece_dataoff; // per piece: byte offset in its packet
    std::vector<WorkUnit> units;
    uint64_t stream_bytes = 0;
};

struct Plan {
    char op = 'N';
    int elem_bytes = 8;
    int32_t n_in = 0, n_out = 0;
    std::vector<int32_t> bin_start; // n_bins + 1
    std::vector<VLeaf> vleaves;
    std::vector<Piece1> pieces1;
    std::vector<Piece2> pieces2;
    PhasePlan phase1, phase2;
    std::vector<ReduceSeg> reduce;
    // multi-RHS phase 1 (T_j = V_j X): V tiles grouped by SOURCE bin; the bin's rows of X stay resident in shared memory
    // while its tiles stream.  Leaves wider than a bin produce partial T rows (workspace rows >= xz_elems) summed by
    // reduce_mm; piece_zoff = first row of the tile's columns inside the resident X slice.
    std::vector<int32_t> src_bin_start; // n_src_bins + 1
    std::vector<Piece1> pieces1m;        // sub-pieces; out_off = workspace row of the sub-piece's first output row
    std::vector<MmTile> mm_tiles;        // the packed tiles (the `pieces` of phase1m)
    std::vector<uint32_t> reduce_mm_rows; // part rows of the segments: seg.first_partial indexes this array
    PhasePlan phase1m;
    std::vector<ReduceSeg> reduce_mm;   // in workspace ROWS
    uint64_t mm_rows = 0;               // workspace rows of the multi-RHS path: [x | t | (matvec partials) | mm partials]
    uint64_t xz_elems = 0; // x | t | partials (+ padding)
    uint32_t t_begin  = 0; // first element of the t region
    double stored_coefficients = 0;
};

// builds the schedule for out = op(A) in ; throws std::runtime_error on invalid input
void build_plan(const hmb200_flat &flat, int elem_bytes, char op, const Options &opt, Plan &plan);

// fills packets [stage_begin, stage_end) of phase `phase` (1 or 2) into dst, where dst corresponds to stream offset
// plan.phaseN.stages[stage_begin].gmem_off.  T is double or std::complex<double> (selected by plan.elem_bytes).
void pack_stages(const hmb200_flat &flat, const Plan &plan, int phase, size_t stage_begin, size_t stage_end, uint8_t *dst);

} // namespace hmb200
