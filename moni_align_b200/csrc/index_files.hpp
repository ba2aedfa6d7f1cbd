// Host-side reader of the two index files the unmodified CPU builder produces:
//   <prefix>.thrbv.ms   = ms_pointers<>::serialize      (reference include/ms/ms_pointers.hpp:279-297)
//   <prefix>.plain.slp  = PlainSlp<u32,Fblc,Fblc>::serialize (reference thirdparty/ShapedSlp/PlainSlp.hpp:322-335)
// Written from the on-disk layouts; it does not link or include sdsl.
// The succinct containers (Elias-Fano sd_vector, Huffman-shaped wavelet tree, bit-packed
// int_vector) are decoded ONCE into plain per-run arrays: the "flat per-run model" that
// SURVEY.md Appendix A shows is sufficient for bit-exact results.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace moni {

struct FlatRlbwt {
    uint64_t n = 0;                    // BWT length (= |T| + 1)
    uint64_t R = 0;                    // number of runs
    uint64_t terminator_position = 0;
    uint64_t F[256] = {0};             // F[c] = # BWT symbols < c (as stored by the builder)
    uint64_t count[256] = {0};         // occurrences of c            (runs_per_letter[c].size())
    uint64_t nruns[256] = {0};         // number of c-runs
    std::vector<uint8_t> heads;        // R      run head symbols   (wt_huff bulk decode)
    std::vector<uint64_t> starts;      // R + 1  first BWT position of each run, starts[R] = n
    std::vector<uint64_t> samples_start;  // R   text position of the BWT char at the run's first row
    std::vector<uint64_t> samples_last;   // R   ... at the run's last row
    std::vector<uint64_t> thr[256];    // per letter: the stored (non-zero) threshold positions, increasing
    // `moni align` index only (.thrbv.full[.lcp].ms): the Phi / Phi^-1 structures, else empty
    std::vector<uint64_t> pred, pred_to_run;              // r_index::pred (set bits, increasing), pred_to_run   (R each)
    std::vector<uint64_t> pred_start, pred_start_to_run;  // ms_pointers::pred_start, pred_start_to_run          (R each)
};

struct FlatSlp {
    uint64_t text_len = 0;
    uint64_t num_rules = 0;
    uint32_t start_var = 0;
    uint64_t alph_size = 0;
    std::vector<uint8_t> alph;
    std::vector<uint32_t> left, right;   // num_rules  (variable ids; id < alph_size is a terminal)
    std::vector<uint64_t> explen;        // num_rules
};

// Both throw std::runtime_error with a message naming the file and the failing field.
void read_thrbv_ms(const std::string& path, FlatRlbwt& out, bool full = false);
void read_plain_slp(const std::string& path, FlatSlp& out);
// <prefix>.slp = SelfShapedSlp<u32,DagcSd,DagcSd,SelSd>::serialize (reference thirdparty/ShapedSlp/SelfShapedSlp.hpp:399-418),
// the grammar `ms -q` / `mems -q` name (src/matching_statistics.cpp:164-168,178-182): decoded into the same flat rule
// table (children by variable id, the root sequence folded into a balanced tree of extra rules).
void read_shaped_slp(const std::string& path, FlatSlp& out);

}  // namespace moni
