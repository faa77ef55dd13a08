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
    int build_matmat    = 1;     // also pack the two panel streams of the multi-RHS kernels
    int mm1_kg          = 6;     // k-groups (8 columns real / 4 complex) per phase-1 packet (even): 128 rows x 48 columns = 48 KiB (measured best at N=1M)
    int mm2_kg          = 8;     // k-groups per phase-2 packet (even): 64 rows x 64 columns = 32 KiB
    int mm_task_cols    = 1024;  // columns of a phase-1 unit; wider source clusters are split and their partial T rows summed
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

struct MmPiece { // part of a view placed in a panel packet: panel rows [ro, ro+nr) x packet columns [kcol, kcol+nc)
    int32_t vleaf;
    int32_t r0, nr; // rows of the role matrix (phase 1: rows of V, phase 2: rows of U / of the dense view)
    int32_t c0, nc; // columns of the role matrix
    int32_t ro;     // first panel row
    int32_t kcol;   // first column inside the packet
};
struct MmPacketHost { // planner-side description of a multi-RHS packet
    uint32_t first_piece, n_pieces;
    uint32_t n_kg, out0;
    uint16_t rows, flags;
    uint32_t map_first; // phase 1, last packet: first entry of mm_row_map (rows entries)
};
struct MmPhasePlan {
    std::vector<StageEntry> stages;
    std::vector<MmPacketHost> packets;
    std::vector<MmPiece> pieces;
    std::vector<ZCopy> zcopies; // src_elem = workspace row, bytes = rows * elem_bytes, dst_off = z-area row * elem_bytes
    std::vector<WorkUnit> units;
    uint64_t stream_bytes = 0;
    uint32_t max_packet_bytes = 0, max_z_rows = 0;
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
    // multi-RHS path (plan.h, "mm streams"): workspace ROWS [ x (n_in) | T (sum of ranks, tight, view order) | partial T rows
    // of split phase-1 units | kMmZeroRows zero rows ]
    MmPhasePlan mm1, mm2;
    std::vector<uint32_t> mm_t_row;       // per view: first workspace row of its T (low-rank views)
    std::vector<uint32_t> mm_row_map;     // phase 1: destination workspace row of every panel row of every unit
    std::vector<uint32_t> reduce_mm_rows; // part rows of the segments: seg.first_partial indexes this array
    std::vector<ReduceSeg> reduce_mm;     // in workspace ROWS: final row, first entry of reduce_mm_rows, parts, rows
    uint64_t mm_rows = 0;                 // workspace rows in total
    uint32_t mm_zero_row = 0;             // first of the zero rows
    uint64_t xz_elems = 0; // x | t | partials (+ padding)
    uint32_t t_begin  = 0; // first element of the t region
    double stored_coefficients = 0;
};

// builds the schedule for out = op(A) in ; throws std::runtime_error on invalid input
void build_plan(const hmb200_flat &flat, int elem_bytes, char op, const Options &opt, Plan &plan);

// fills packets [stage_begin, stage_end) of phase `phase` (1 or 2; 3 / 4 = the multi-RHS panel streams mm1 / mm2) into dst, where dst corresponds to stream offset
// plan.phaseN.stages[stage_begin].gmem_off.  T is double or std::complex<double> (selected by plan.elem_bytes).
void pack_stages(const hmb200_flat &flat, const Plan &plan, int phase, size_t stage_begin, size_t stage_end, uint8_t *dst);

} // namespace hmb200
