// See index_files.hpp.  Layout citations are to the reference checkout (read-only, never linked):
//   sdsl int_vector        thirdparty/sdsl-lite/include/sdsl/int_vector.hpp:593-609,1546-1595
//   sdsl sd_vector         thirdparty/sdsl-lite/include/sdsl/sd_vector.hpp:404-427
//   sdsl select_support    thirdparty/sdsl-lite/include/sdsl/select_support_mcl.hpp:424-459
//   sdsl wt_pc / tree      thirdparty/sdsl-lite/include/sdsl/wt_pc.hpp:656-683, wt_helper.hpp:74-127,264-288
//   sparse_sd_vector       thirdparty/r-index/internal/sparse_sd_vector.hpp:213-241
//   rle_string             thirdparty/r-index/internal/rle_string.hpp:347-374
//   ms_pointers            include/ms/ms_pointers.hpp:279-320 ; thr_bv include/ms/thresholds_ds.hpp:502-523
//   PlainSlp               thirdparty/ShapedSlp/PlainSlp.hpp:305-335
#include "index_files.hpp"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <cstring>
#include <stdexcept>

namespace moni {
namespace {

struct MappedFile {
    const uint8_t* p = nullptr;
    size_t size = 0;
    int fd = -1;
    explicit MappedFile(const std::string& path) {
        fd = ::open(path.c_str(), O_RDONLY);
        if (fd < 0) throw std::runtime_error("cannot open " + path);
        struct stat st;
        if (fstat(fd, &st) != 0) { ::close(fd); throw std::runtime_error("cannot stat " + path); }
        size = (size_t)st.st_size;
        if (size) {
            void* m = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
            if (m == MAP_FAILED) { ::close(fd); throw std::runtime_error("cannot mmap " + path); }
            p = (const uint8_t*)m;
            madvise(m, size, MADV_SEQUENTIAL);
        }
    }
    ~MappedFile() {
        if (p) munmap((void*)p, size);
        if (fd >= 0) ::close(fd);
    }
    MappedFile(const MappedFile&) = delete;
    MappedFile& operator=(const MappedFile&) = delete;
};

struct Cursor {
    const uint8_t* base;
    size_t size;
    size_t off = 0;
    std::string name;
    void need(size_t k, const char* what) const {
        if (k > size - off) throw std::runtime_error(name + ": truncated while reading " + what);
    }
    uint64_t u64(const char* what) { need(8, what); uint64_t v; memcpy(&v, base + off, 8); off += 8; return v; }
    uint32_t u32(const char* what) { need(4, what); uint32_t v; memcpy(&v, base + off, 4); off += 4; return v; }
    uint16_t u16(const char* what) { need(2, what); uint16_t v; memcpy(&v, base + off, 2); off += 2; return v; }
    uint8_t u8(const char* what) { need(1, what); return base[off++]; }
    const uint8_t* bytes(size_t k, const char* what) { need(k, what); const uint8_t* r = base + off; off += k; return r; }
};

// A bit-packed vector viewed in place (words are 8-byte aligned in none of these files, so
// every word access goes through memcpy).
struct PackedView {
    uint64_t bits = 0;
    uint8_t width = 1;
    const uint8_t* words = nullptr;
    uint64_t size() const { return width ? bits / width : 0; }
    uint64_t word(uint64_t i) const { uint64_t v; memcpy(&v, words + 8 * i, 8); return v; }
    uint64_t get(uint64_t i) const {
        const uint64_t b = i * width, w = b >> 6, s = b & 63;
        uint64_t v = word(w) >> s;
        if (s + width > 64) v |= word(w + 1) << (64 - s);
        return width == 64 ? v : (v & ((1ULL << width) - 1));
    }
};

PackedView read_int_vector(Cursor& c, int fixed_width, const char* what) {
    PackedView v;
    v.bits = c.u64(what);
    v.width = fixed_width ? (uint8_t)fixed_width : c.u8(what);
    const uint64_t nwords = (v.bits + 63) >> 6;
    v.words = c.bytes(nwords * 8, what);
    return v;
}

void skip_select_support(Cursor& c, const char* what) {
    const uint64_t arg_cnt = c.u64(what);
    if (!arg_cnt) return;
    read_int_vector(c, 0, what);                         // superblock
    read_int_vector(c, 1, what);                         // mini_or_long
    const uint64_t sb = (arg_cnt + 4095) >> 12;
    for (uint64_t i = 0; i < sb; ++i) read_int_vector(c, 0, what);  // one long- or mini-block each
}

// Elias-Fano sd_vector -> positions of its 1 bits.  Returns the universe size.
uint64_t read_sd_vector(Cursor& c, std::vector<uint64_t>& ones, const char* what) {
    const uint64_t size = c.u64(what);
    const uint8_t wl = c.u8(what);
    PackedView low = read_int_vector(c, 0, what);
    PackedView high = read_int_vector(c, 1, what);
    skip_select_support(c, what);
    skip_select_support(c, what);
    const uint64_t m = low.size();
    ones.clear();
    ones.reserve(m);
    if (wl != low.width && m) throw std::runtime_error(c.name + ": sd_vector wl/low width mismatch in " + what);
    uint64_t i = 0;
    const uint64_t nwords = (high.bits + 63) >> 6;
    for (uint64_t w = 0; w < nwords && i < m; ++w) {
        uint64_t x = high.word(w);
        while (x && i < m) {
            const uint64_t p = (w << 6) + (uint64_t)__builtin_ctzll(x);
            x &= x - 1;
            ones.push_back(((p - i) << wl) | (wl ? low.get(i) : 0));
            ++i;
        }
    }
    if (i != m) throw std::runtime_error(c.name + ": sd_vector high has fewer 1s than low entries in " + what);
    return size;
}

// ri::sparse_sd_vector: u64 u, then (if u != 0) the sd_vector.
uint64_t read_sparse_sd(Cursor& c, std::vector<uint64_t>& ones, const char* what) {
    ones.clear();
    const uint64_t u = c.u64(what);
    if (u == 0) return 0;
    read_sd_vector(c, ones, what);
    return u;
}

void unpack(const PackedView& v, std::vector<uint64_t>& out) {
    const uint64_t m = v.size();
    out.resize(m);
    for (uint64_t i = 0; i < m; ++i) out[i] = v.get(i);
}

// wt_huff<> -> the symbol sequence.  No rank structure is needed for a bulk decode: one cursor
// per inner node, consumed in sequence order.
void read_wt_huff(Cursor& c, std::vector<uint8_t>& seq) {
    const uint64_t size = c.u64("wt.size");
    const uint64_t sigma = c.u64("wt.sigma");
    (void)sigma;
    PackedView bv = read_int_vector(c, 1, "wt.bv");
    read_int_vector(c, 64, "wt.bv_rank");
    skip_select_support(c, "wt.bv_select1");
    skip_select_support(c, "wt.bv_select0");
    const uint64_t n_nodes = c.u64("wt.tree.n_nodes");
    if (n_nodes > 1024) throw std::runtime_error(c.name + ": implausible wavelet tree node count");
    struct Node { uint64_t bv_pos; uint16_t child[2]; };
    std::vector<Node> nodes(n_nodes);
    for (uint64_t i = 0; i < n_nodes; ++i) {
        nodes[i].bv_pos = c.u64("wt.node.bv_pos");
        c.u64("wt.node.bv_pos_rank");
        c.u16("wt.node.parent");
        nodes[i].child[0] = c.u16("wt.node.child0");
        nodes[i].child[1] = c.u16("wt.node.child1");
    }
    uint16_t c_to_leaf[256];
    memcpy(c_to_leaf, c.bytes(512, "wt.c_to_leaf"), 512);
    c.bytes(256 * 8, "wt.path");
    std::vector<int> leaf_sym(n_nodes, -1);
    for (int s = 0; s < 256; ++s)
        if (c_to_leaf[s] != 0xFFFF) {
            if (c_to_leaf[s] >= n_nodes) throw std::runtime_error(c.name + ": wt leaf id out of range");
            leaf_sym[c_to_leaf[s]] = s;
        }
    seq.resize(size);
    if (size == 0) return;
    if (n_nodes == 0) throw std::runtime_error(c.name + ": empty wavelet tree for a non-empty sequence");
    std::vector<uint64_t> cur(n_nodes);
    for (uint64_t i = 0; i < n_nodes; ++i) cur[i] = nodes[i].bv_pos;
    for (uint64_t i = 0; i < size; ++i) {
        uint32_t v = 0;
        while (nodes[v].child[0] != 0xFFFF) {
            const uint64_t b = cur[v]++;
            if (b >= bv.bits) throw std::runtime_error(c.name + ": wavelet tree bit cursor out of range");
            const uint64_t bit = (bv.word(b >> 6) >> (b & 63)) & 1;
            v = nodes[v].child[bit];
            if (v >= n_nodes) throw std::runtime_error(c.name + ": wavelet tree child out of range");
        }
        if (leaf_sym[v] < 0) throw std::runtime_error(c.name + ": wavelet tree leaf without a symbol");
        seq[i] = (uint8_t)leaf_sym[v];
    }
}

}  // namespace

// `full`: the `moni align` index (include/ms/moni.hpp:359-382 / include/aligner/moni_lcp.hpp:208-225): the same members
// in the same order with the Phi structures in between -- after the BWT: pred (sparse_sd_vector) and pred_to_run
// (int_vector<>), after the thresholds: pred_start and pred_start_to_run -- and, in the .lcp variant, one trailing
// int_vector<> slcp.  The matching-statistics query (moni.hpp:568-626) reads none of them; the Phi members are kept for
// occurrence enumeration, slcp is parsed and skipped.
void read_thrbv_ms(const std::string& path, FlatRlbwt& o, bool full) {
    MappedFile mf(path);
    Cursor c{mf.p, mf.size, 0, path};
    o.terminator_position = c.u64("terminator_position");
    const uint64_t fsz = c.u64("F.size");
    if (fsz != 256) throw std::runtime_error(path + ": F has " + std::to_string(fsz) + " entries, expected 256");
    for (int i = 0; i < 256; ++i) o.F[i] = c.u64("F");
    // rle_string
    o.n = c.u64("bwt.n");
    o.R = c.u64("bwt.R");
    const uint64_t B = c.u64("bwt.B");
    if (o.n == 0 || o.R == 0 || o.R > o.n || B == 0) throw std::runtime_error(path + ": implausible n/R/B header");
    if (o.n >= (1ULL << 40)) throw std::runtime_error(path + ": BWT length >= 2^40 is not supported");
    if (o.R >= 0xFFFFFFF0ULL) throw std::runtime_error(path + ": run count >= 2^32 is not supported");
    std::vector<uint64_t> run_marks;
    read_sparse_sd(c, run_marks, "bwt.runs");
    std::vector<std::vector<uint64_t>> per_letter(256);
    for (int s = 0; s < 256; ++s) {
        o.count[s] = read_sparse_sd(c, per_letter[s], "bwt.runs_per_letter");
        o.nruns[s] = per_letter[s].size();
    }
    read_wt_huff(c, o.heads);
    if (o.heads.size() != o.R) throw std::runtime_error(path + ": run_heads length != R");
    // run lengths = gaps of runs_per_letter[head] (ms_rle_string.hpp:284-285, sparse_sd_vector.hpp:179-187)
    o.starts.resize(o.R + 1);
    {
        uint64_t seen[256] = {0};
        uint64_t pos = 0;
        for (uint64_t k = 0; k < o.R; ++k) {
            const uint8_t h = o.heads[k];
            const uint64_t j = seen[h]++;
            if (j >= per_letter[h].size()) throw std::runtime_error(path + ": more runs of a letter than run marks");
            const uint64_t last = per_letter[h][j];
            const uint64_t len = j ? last - per_letter[h][j - 1] : last + 1;
            if (len == 0 || len > o.n) throw std::runtime_error(path + ": bad run length");
            o.starts[k] = pos;
            pos += len;
        }
        o.starts[o.R] = pos;
        if (pos != o.n) throw std::runtime_error(path + ": run lengths do not sum to n");
        for (int s = 0; s < 256; ++s)
            if (seen[s] != o.nruns[s]) throw std::runtime_error(path + ": per-letter run count mismatch");
        // integrity: `runs` marks the last position of every B-th run (and of the last run)
        for (uint64_t i = 0; i < run_marks.size(); ++i) {
            const uint64_t k = (i + 1) * B;
            if (k > o.R || run_marks[i] != o.starts[k] - 1) throw std::runtime_error(path + ": runs bitvector disagrees with run lengths");
        }
    }
    per_letter.clear();
    per_letter.shrink_to_fit();
    if (full) {
        read_sparse_sd(c, o.pred, "pred");
        unpack(read_int_vector(c, 0, "pred_to_run"), o.pred_to_run);
        if (o.pred_to_run.size() != o.R || o.pred.size() != o.R)
            throw std::runtime_error(path + ": Phi structures (pred) do not have R entries");
    }
    unpack(read_int_vector(c, 0, "samples_last"), o.samples_last);
    for (int s = 0; s < 256; ++s) {
        read_sparse_sd(c, o.thr[s], "thresholds_per_letter");
        for (size_t i = 1; i < o.thr[s].size(); ++i)
            if (o.thr[s][i] <= o.thr[s][i - 1]) throw std::runtime_error(path + ": thresholds not increasing");
    }
    if (full) {
        read_sparse_sd(c, o.pred_start, "pred_start");
        unpack(read_int_vector(c, 0, "pred_start_to_run"), o.pred_start_to_run);
        if (o.pred_start_to_run.size() != o.R || o.pred_start.size() != o.R)
            throw std::runtime_error(path + ": Phi structures (pred_start) do not have R entries");
    }
    unpack(read_int_vector(c, 0, "samples_start"), o.samples_start);
    if (full && c.off != c.size) {                    // .thrbv.full.lcp.ms: sampled LCP values, one per run
        if (read_int_vector(c, 0, "slcp").size() != o.R) throw std::runtime_error(path + ": slcp does not have R entries");
    }
    if (o.samples_last.size() != o.R || o.samples_start.size() != o.R)
        throw std::runtime_error(path + ": sample arrays do not have R entries");
    if (c.off != c.size) throw std::runtime_error(path + ": " + std::to_string(c.size - c.off) + " trailing bytes");
}

void read_plain_slp(const std::string& path, FlatSlp& o) {
    MappedFile mf(path);
    Cursor c{mf.p, mf.size, 0, path};
    o.text_len = c.u64("len");
    o.num_rules = c.u64("numRules");
    o.start_var = c.u32("startVar");
    o.alph_size = c.u64("alphSize");
    if (o.alph_size == 0 || o.alph_size > 256) throw std::runtime_error(path + ": implausible alphabet size");
    const uint8_t* a = c.bytes(o.alph_size, "alph");
    o.alph.assign(a, a + o.alph_size);
    PackedView l = read_int_vector(c, 0, "left");
    PackedView r = read_int_vector(c, 0, "right");
    PackedView e = read_int_vector(c, 0, "expLen");
    if (l.size() != o.num_rules || r.size() != o.num_rules || e.size() != o.num_rules)
        throw std::runtime_error(path + ": rule arrays do not have numRules entries");
    if (o.num_rules + o.alph_size >= 0xFFFFFFF0ULL) throw std::runtime_error(path + ": too many variables");
    o.left.resize(o.num_rules);
    o.right.resize(o.num_rules);
    o.explen.resize(o.num_rules);
    const uint64_t nv = o.num_rules + o.alph_size;
    for (uint64_t i = 0; i < o.num_rules; ++i) {
        const uint64_t lv = l.get(i), rv = r.get(i);
        if (lv >= nv || rv >= nv) throw std::runtime_error(path + ": rule child out of range");
        o.left[i] = (uint32_t)lv;
        o.right[i] = (uint32_t)rv;
        o.explen[i] = e.get(i);
    }
    if (o.num_rules) {
        if (o.start_var < o.alph_size || o.start_var - o.alph_size >= o.num_rules)
            throw std::runtime_error(path + ": start variable out of range");
        if (o.explen[o.start_var - o.alph_size] != o.text_len)
            throw std::runtime_error(path + ": expLen[startVar] != len");
    }
    if (c.off != c.size) throw std::runtime_error(path + ": " + std::to_string(c.size - c.off) + " trailing bytes");
}

}  // namespace moni
