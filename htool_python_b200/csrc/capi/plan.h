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

} // namespace hmb200
