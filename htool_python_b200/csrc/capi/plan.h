// plan.h -- data layout shared by the host planner/packer (planner.cpp) and the device kernels (kernels/*.cuh).
//
// HBM layout of one packed operator ("engine"):
//   stream1 : phase-1 packets, t = V x      (V tiles, lr x nq col-major, lr <= 32)
//   stream2 : phase-2 packets, y += U t / y += D x  (tiles nr x nc col-major, nr <= 64 rows of ONE target bin)
//   each packet = [PacketHeader 16 B][PieceDesc 16 B x n_pieces][tile data, each tile padded to 16 B]
//   and is fetched by ONE cp.async.bulk into a shared-memory stage; the operand vectors the tiles multiply
//   (slices of x, or of t) are fetched into the stage's z-area by small bulk copies listed in the ZCopy table.
//   xz      : workspace of T: [ x (n_in) | t (final per low-rank leaf) | partial t (one slot per phase-1 piece) ]
#pragma once
#include <stdint.h>

namespace hmb200 {

constexpr int kMaxBinRows    = 64; // rows of a target bin: lane l owns bin rows l and l+32
constexpr int kMaxPhase1Rows = 32; // l-range of a phase-1 tile
constexpr uint32_t kFlagFirstOfBin = 1u;
constexpr uint32_t kFlagLastOfBin  = 2u;
constexpr uint32_t kFlagTerminate  = 4u;

struct PacketHeader { // 16 B
    uint32_t n_pieces;
    uint32_t piece_base; // running piece index of the first piece (for round-robin warp assignment)
    uint32_t row0;       // phase 2: first output row of the bin
    uint16_t bin_rows;   // phase 2: rows of the bin
    uint16_t flags;
};

struct PieceDesc { // 16 B
    uint32_t data_off; // byte offset of the tile inside the packet (16 B aligned)
    uint32_t out_off;  // phase 1: element offset into xz where the lr results go
    uint16_t z_off;    // element offset inside the stage z-area of the operand slice
    uint16_t nc;       // phase 2: columns; phase 1: nq (columns of the V tile)
    uint8_t nr;        // phase 2: rows (<= 64); phase 1: lr (<= 32)
    uint8_t ro;        // phase 2: first row inside the bin; phase 1: G = 32 / lr column groups
    uint16_t pad;      // phase 1: ceil(32768 / lr), so that lane / lr == (lane * pad) >> 15
};

struct StageEntry { // 32 B, read by the producer warp
    uint64_t gmem_off; // byte offset of the packet in the stream (16 B aligned)
    uint32_t bytes;    // packet bytes (multiple of 16)
    uint32_t z_first;  // first ZCopy of this stage
    uint32_t n_z;      // number of ZCopies
    uint32_t z_bytes;  // total bytes of the ZCopies
    uint32_t pad[2];
};

struct ZCopy { // 8 B
    uint32_t src_elem; // element offset in xz (16 B aligned address)
    uint16_t bytes;    // multiple of 16
    uint16_t dst_off;  // byte offset in the stage z-area (multiple of 16)
};

struct WorkUnit { // 8 B
    uint32_t first_stage;
    uint32_t n_stages;
};

struct ReduceSeg { // 16 B: t[final_off + l] = sum_q xz[first_partial + q*len + l], l < len
    uint32_t final_off;
    uint32_t first_partial;
    uint32_t n_parts;
    uint32_t len;
};

// ---- multi-RHS ("mm") streams: every unit is one panel GEMM  C (rows x mu) = A (rows x K) B (K x mu) -------------------------
//   phase 1: A = V rows of the low-rank leaves that share one source cluster, stacked (<= 128 rows), K = a chunk of that
//            cluster's columns, B = the rows of X for those columns, C = T rows (or partial T rows when K is split)
//   phase 2: A = [U_1 | U_2 | ... | D_1 | ...] of ONE target bin (<= 64 rows), concatenated along K, B = the T rows / X rows of
//            those leaves in the same order, C = the bin's rows of Y
// A is packed FRAGMENT-MAJOR for mma.sync.m8n8k4.f64: the panel is cut into m-tiles of 8 rows and k-groups of kMmKg columns
// (8 real / 4 complex); chunk (kg, mt) is 32 lanes x 16 bytes, lane = 4*gid + tig holding A[8 mt + gid][kMmKg kg + tig] (and
// ... + 4 + tig for real), so that one 128-bit shared load per lane yields the A fragments of the whole k-group, conflict-free.
// Packet = [MmPacketHeader 16 B][chunks, kg-major then mt][phase 1, last packet of a unit: uint32 row map].  The operand rows of
// a packet are gathered into the stage's z-area by bulk copies (ZCopy: src_elem = workspace ROW, bytes / dst_off in units of
// elem_bytes per row), in the order of the packet's K columns.
constexpr uint32_t kMmFlagFirst = 1u, kMmFlagLast = 2u, kMmFlagTerminate = 4u;
constexpr int kMmRows1 = 128; // rows of a phase-1 panel (16 m-tiles)
constexpr int kMmRows2 = 64;  // rows of a phase-2 panel = a target bin (8 m-tiles)
constexpr int kMmZeroRows = 8; // zero rows at the end of the workspace: the operand of K padding

struct MmPacketHeader { // 16 B
    uint32_t n_kg;      // k-groups in this packet
    uint32_t out0;      // phase 2: first output row of the bin
    uint16_t rows;      // rows of the panel
    uint16_t mt_panel;  // m-tiles of the panel = ceil(rows / 8)
    uint16_t flags;     // kMmFlag*
    uint16_t map_off16; // phase 1, last packet: offset of the row map in 16-byte units (one destination workspace row per panel row)
};

} // namespace hmb200
