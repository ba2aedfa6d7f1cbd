// The flattened index as it lives in HBM, and the host-side exporter that produces it from the
// decoded builder files (index_files.hpp).  Plain structs shared by host and device code.
//
// Pass 1 (MS pointers; reference include/ms/ms_pointers.hpp:455-511) runs on a "move table":
// one 32-byte row per BWT run (long runs and runs whose LF image would cover many rows are cut
// into pieces).  A row carries everything one backward step needs, so a step costs ONE random
// 32-byte sector instead of the reference's chain of Elias-Fano / wavelet-tree probes:
//   * the LF image of the row's first position as (destination row, offset)   -> rle_string::rank
//     (thirdparty/r-index/internal/rle_string.hpp:230-260) becomes dest + offset, then a short
//     forward scan over consecutive rows;
//   * for each other frequent symbol c: the threshold of the enclosing c-gap clipped to the row
//     (thr_bv::rank, include/ms/thresholds_ds.hpp:494-497, becomes one compare) and the distance
//     in rows to the previous / next c-run (run_and_head_rank + run_head_select,
//     include/ms/ms_rle_string.hpp:152-167, become a row index);
//   * samples_start / samples_last of the run as 40-bit integers (ms_pointers.hpp:48,493,501).
// Anything the rows cannot express (symbols outside the 4 most frequent run heads, pos == n,
// absent symbols, neighbours further than 254 rows) goes through an exact general path over the
// per-symbol tables below (binary searches; rare by construction).
//
// Pass 2 (MS lengths / MEMs; reference src/matching_statistics.cpp:246-255, src/mems.cpp:247-259)
// runs on a linearised grammar: PlainSlp rules (thirdparty/ShapedSlp/PlainSlp.hpp:101-126) whose
// expansion is longer than `leaf_len` become 16-byte nodes in breadth-first order (top levels
// contiguous and L2 resident); shorter variables are expanded once into a byte pool, so a
// descent ends in a literal block that is compared sequentially.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace moni {

constexpr uint32_t kMaxPiece = 65535;        // longest row (pieces of longer runs share a head)
constexpr uint8_t kFar = 255;                // neighbour distance that does not fit: general path
constexpr uint8_t kRowSlow = 1;              // head is not one of the 4 table symbols (or sentinel)
constexpr uint64_t kThrInf = ~0ULL;

struct alignas(32) MoveRow {
    uint32_t dest_row;      // row containing LF(first position of this row)
    uint16_t dest_off;      // offset of that position inside dest_row
    uint16_t len;           // 1..65535 (0 only in the sentinel row at index n_rows)
    uint16_t t_rel[3];      // per other table symbol: offsets < t_rel jump up, >= t_rel jump down
    uint8_t head;           // the run head byte
    uint8_t flags;          // kRowSlow
    uint8_t s_start[5];     // samples_start of the run (valid in its first piece)
    uint8_t s_last[5];      // samples_last of the run (valid in its last piece)
    uint8_t dist_up[3];     // rows back to the last piece of the previous c-run (kFar = general path)
    uint8_t dist_dn[3];     // rows forward to the first piece of the next c-run
};
static_assert(sizeof(MoveRow) == 32, "MoveRow must be one 32-byte sector");

// General path, per c-run boundary j in [0, nruns_c]: everything a threshold jump between the
// (j-1)-th and j-th c-run needs (0-based).  thr = the j-th smallest stored threshold of c
// (0 for j == 0: always down; kThrInf when there is none: always up).
struct alignas(32) GapSlot {
    uint64_t thr;
    uint64_t s_start;       // samples_start of c-run j        (unused for j == nruns_c)
    uint64_t s_last_prev;   // samples_last of c-run j-1        (unused for j == 0)
    uint64_t cum;           // # of c in BWT before c-run j  (j == nruns_c: count[c])
};
static_assert(sizeof(GapSlot) == 32, "GapSlot must be one 32-byte sector");

struct alignas(16) SlpNode {
    uint32_t left;          // child reference: bit 31 set = leaf (low 31 bits = pool offset >> pool_shift)
    uint32_t right;
    uint64_t left_len;      // expansion length of the left child
};
static_assert(sizeof(SlpNode) == 16, "SlpNode must be 16 bytes");
constexpr uint32_t kLeafBit = 0x80000000u;

// One entry of a Phi / Phi^-1 table (`moni align` indexes only): the key is a text position marked in r_index::pred
// (resp. ms_pointers::pred_start), the value is the sample the reference adds the distance to --
//   Phi      (thirdparty/r-index/internal/r_index.hpp:225-243):  samples_last [pred_to_run[jr] - 1]
//   Phi^-1   (include/ms/moni.hpp:338-354):                       samples_start[pred_start_to_run[jr] + 1]
// so a step is one predecessor search over the keys plus arithmetic.
struct alignas(16) PhiEntry {
    uint64_t key;
    uint64_t val;           // kPhiNone where the reference would index outside its arrays (Phi of SA[0], Phi^-1 of SA[n-1])
};
constexpr uint64_t kPhiNone = ~0ULL;

// Per-symbol constants (device: __constant__ memory).
struct SymbolTables {
    uint64_t F[256];
    uint64_t count[256];        // 0 = symbol absent from the index
    uint64_t slot_base[256];    // first GapSlot of the symbol
    uint64_t crun_base[256];    // first entry of the symbol in crun_row
    uint32_t nruns[256];
    uint8_t tcode[256];         // 0..3 for the table symbols, 0xFF otherwise
};

struct ImageHeader {
    uint64_t n = 0, R = 0, n_rows = 0;          // rows exclude the sentinel
    uint64_t text_len = 0;                       // N = PlainSlp::getLen()
    uint64_t init_sample = 0;                    // (samples_last[R-1] + 1) % n   (r_index.hpp:470-472)
    uint64_t n_nodes = 0, pool_bytes = 0;
    uint32_t pool_shift = 0, leaf_len = 0;
    uint32_t root_ref = 0;                       // node 0, or a leaf reference for tiny texts
    uint32_t max_scan = 0;                       // longest forward scan left after balancing
    uint8_t table_syms[4] = {0, 0, 0, 0};
    uint32_t n_table_syms = 0;
    uint32_t slp_depth_max = 0;
    uint32_t init_row = 0, init_off = 0;         // pos = n - 1 as (row, offset)
    uint64_t phi_n = 0;                          // entries of each Phi table (R), 0 without the `moni align` index
    uint32_t phi_shift = 0;                      // bucket b of a Phi directory covers keys in [b << phi_shift, (b + 1) << phi_shift)
    uint64_t first_run_sample = 0;               // (samples_start[0] + 1) % n   (include/ms/moni.hpp:330-332)
};

// Host copy of the image.
struct HostImage {
    ImageHeader hdr;
    SymbolTables sym;
    std::vector<MoveRow> rows;          // n_rows + 1 (sentinel)
    std::vector<uint64_t> row_pos;      // n_rows + 1 : first BWT position of each row; row_pos[n_rows] = n
    std::vector<uint32_t> crun_row;     // R : per symbol (ascending byte order), first-piece row of each c-run
    std::vector<GapSlot> slots;         // R + (#symbols)
    std::vector<SlpNode> nodes;
    std::vector<uint8_t> pool;
    std::vector<PhiEntry> phi, phi_inv; // R each, keys increasing (empty without the `moni align` index)
    std::vector<uint32_t> phi_dir, phi_inv_dir;   // (n >> phi_shift) + 2 each: dir[b] = # keys < (b << phi_shift)
    uint64_t bytes() const {
        return rows.size() * sizeof(MoveRow) + row_pos.size() * 8 + crun_row.size() * 4 +
               slots.size() * sizeof(GapSlot) + nodes.size() * sizeof(SlpNode) + pool.size() +
               (phi.size() + phi_inv.size()) * sizeof(PhiEntry) + (phi_dir.size() + phi_inv_dir.size()) * 4;
    }
};

struct FlatRlbwt;
struct FlatSlp;

struct ExportOptions {
    uint64_t leaf_len = 128;     // variables expanding to <= leaf_len bytes become literal blocks
    uint32_t max_scan = 8;       // balancing target: rows covered by one row's LF image
    uint32_t max_rounds = 24;
    bool verbose = false;
    uint32_t threads = 0;        // host threads of the exporter's row passes; 0 = one per core, at most 32 (same image for any value)
};

// Throws std::runtime_error on anything the image cannot represent.
void build_move_table(const FlatRlbwt& bwt, const ExportOptions& opt, HostImage& img);
void build_grammar_table(const FlatSlp& slp, const ExportOptions& opt, HostImage& img);
void build_phi_tables(const FlatRlbwt& bwt, HostImage& img);     // no-op when the index has no Phi structures

}  // namespace moni
