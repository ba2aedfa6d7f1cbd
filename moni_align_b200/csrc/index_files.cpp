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
#include <atomic>
#include <cmath>
#include <cstring>
#include <functional>
#include <stdexcept>
#include <thread>

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

// Decoding work found while walking a file: the walk itself only reads sizes, the decodes run afterwards on all cores.
struct Tasks {
    std::vector<std::function<void()>> list;
    void add(std::function<void()> fn) { list.push_back(std::move(fn)); }
    void run() {
        const unsigned T = (unsigned)std::max<size_t>(1, std::min<size_t>({list.size(), (size_t)32, (size_t)std::max(1u, std::thread::hardware_concurrency())}));
        std::atomic<size_t> next{0};
        std::vector<std::exception_ptr> err(T);
        auto work = [&](unsigned t) {
            try {
                for (size_t i; (i = next.fetch_add(1)) < list.size();) list[i]();
            } catch (...) { err[t] = std::current_exception(); }
        };
        std::vector<std::thread> th;
        for (unsigned t = 1; t < T; ++t) th.emplace_back(work, t);
        work(0);
        for (auto& x : th) x.join();
        list.clear();
        for (auto& e : err) if (e) std::rethrow_exception(e);
    }
};

// Elias-Fano sd_vector -> positions of its 1 bits.  Returns the universe size; `ones` is filled by a task.
uint64_t read_sd_vector(Cursor& c, std::vector<uint64_t>& ones, const char* what, Tasks* tasks = nullptr) {
    const uint64_t size = c.u64(what);
    const uint8_t wl = c.u8(what);
    PackedView low = read_int_vector(c, 0, what);
    PackedView high = read_int_vector(c, 1, what);
    skip_select_support(c, what);
    skip_select_support(c, what);
    const uint64_t m = low.size();
    ones.clear();
    if (wl != low.width && m) throw std::runtime_error(c.name + ": sd_vector wl/low width mismatch in " + what);
    const std::string name = c.name;
    auto decode = [&ones, low, high, wl, m, name, what]() {
        ones.reserve(m);
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
        if (i != m) throw std::runtime_error(name + ": sd_vector high has fewer 1s than low entries in " + what);
    };
    if (tasks) tasks->add(decode);
    else decode();
    return size;
}

// ri::sparse_sd_vector: u64 u, then (if u != 0) the sd_vector.
uint64_t read_sparse_sd(Cursor& c, std::vector<uint64_t>& ones, const char* what, Tasks* tasks = nullptr) {
    ones.clear();
    const uint64_t u = c.u64(what);
    if (u == 0) return 0;
    read_sd_vector(c, ones, what, tasks);
    return u;
}

void unpack(const PackedView& v, std::vector<uint64_t>& out) {
    const uint64_t m = v.size();
    out.resize(m);
    for (uint64_t i = 0; i < m; ++i) out[i] = v.get(i);
}

// wt_huff<> -> the symbol sequence.  No rank structure is needed for a bulk decode: one cursor
// per inner node, consumed in sequence order.
void read_wt_huff(Cursor& c, std::vector<uint8_t>& seq, Tasks& tasks) {
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
    if (size && n_nodes == 0) throw std::runtime_error(c.name + ": empty wavelet tree for a non-empty sequence");
    const std::string name = c.name;
    tasks.add([&seq, size, n_nodes, nodes, leaf_sym, bv, name]() {
        seq.resize(size);
        std::vector<uint64_t> cur(n_nodes);
        for (uint64_t i = 0; i < n_nodes; ++i) cur[i] = nodes[i].bv_pos;
        for (uint64_t i = 0; i < size; ++i) {
            uint32_t v = 0;
            while (nodes[v].child[0] != 0xFFFF) {
                const uint64_t b = cur[v]++;
                if (b >= bv.bits) throw std::runtime_error(name + ": wavelet tree bit cursor out of range");
                const uint64_t bit = (bv.word(b >> 6) >> (b & 63)) & 1;
                v = nodes[v].child[bit];
                if (v >= n_nodes) throw std::runtime_error(name + ": wavelet tree child out of range");
            }
            if (leaf_sym[v] < 0) throw std::runtime_error(name + ": wavelet tree leaf without a symbol");
            seq[i] = (uint8_t)leaf_sym[v];
        }
    });
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
    // the walk below only reads sizes and records where every member lies; the decodes run on all cores afterwards
    Tasks tasks;
    auto unpack_later = [&tasks](PackedView v, std::vector<uint64_t>& out) { tasks.add([v, &out]() { unpack(v, out); }); };
    std::vector<uint64_t> run_marks;
    read_sparse_sd(c, run_marks, "bwt.runs", &tasks);
    std::vector<std::vector<uint64_t>> per_letter(256);
    for (int s = 0; s < 256; ++s) o.count[s] = read_sparse_sd(c, per_letter[s], "bwt.runs_per_letter", &tasks);
    read_wt_huff(c, o.heads, tasks);
    if (full) {
        read_sparse_sd(c, o.pred, "pred", &tasks);
        unpack_later(read_int_vector(c, 0, "pred_to_run"), o.pred_to_run);
    }
    unpack_later(read_int_vector(c, 0, "samples_last"), o.samples_last);
    for (int s = 0; s < 256; ++s) read_sparse_sd(c, o.thr[s], "thresholds_per_letter", &tasks);
    if (full) {
        read_sparse_sd(c, o.pred_start, "pred_start", &tasks);
        unpack_later(read_int_vector(c, 0, "pred_start_to_run"), o.pred_start_to_run);
    }
    unpack_later(read_int_vector(c, 0, "samples_start"), o.samples_start);
    if (full && c.off != c.size) {                    // .thrbv.full.lcp.ms: sampled LCP values, one per run
        if (read_int_vector(c, 0, "slcp").size() != o.R) throw std::runtime_error(path + ": slcp does not have R entries");
    }
    if (c.off != c.size) throw std::runtime_error(path + ": " + std::to_string(c.size - c.off) + " trailing bytes");
    tasks.run();
    for (int s = 0; s < 256; ++s) o.nruns[s] = per_letter[s].size();
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
    if (full && (o.pred_to_run.size() != o.R || o.pred.size() != o.R))
        throw std::runtime_error(path + ": Phi structures (pred) do not have R entries");
    for (int s = 0; s < 256; ++s)
        for (size_t i = 1; i < o.thr[s].size(); ++i)
            if (o.thr[s][i] <= o.thr[s][i - 1]) throw std::runtime_error(path + ": thresholds not increasing");
    if (full && (o.pred_start_to_run.size() != o.R || o.pred_start.size() != o.R))
        throw std::runtime_error(path + ": Phi structures (pred_start) do not have R entries");
    if (o.samples_last.size() != o.R || o.samples_start.size() != o.R)
        throw std::runtime_error(path + ": sample arrays do not have R entries");
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


// ---- shaped grammar, <prefix>.slp --------------------------------------------------------------------------------
// SelfShapedSlp<u32, DagcSd, DagcSd, SelSd>::serialize (thirdparty/ShapedSlp/SelfShapedSlp.hpp:399-418): alphabet, a
// RecSplit minimal perfect hash from "expansion length" to "group of the rules with that length"
// (external/sux/sux/function/RecSplit.hpp:806-831), the prefix sums of the root sequence, the group boundaries, the
// left-length table and three gamma-coded arrays (DirectAccessibleGammaCode.hpp:135-143).  A child is named by
// (its expansion length, its offset inside the group of that length): the hash has to be evaluated to follow an edge.
namespace {

inline uint64_t rot64(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }

// SpookyHash V2 (Bob Jenkins, public domain), 128-bit, both seeds 0, of the 8 little-endian bytes of `key`: a message
// shorter than 16 bytes only meets the final mix of the short variant (eleven xor-rotate-add rounds over a rotating
// pair of the four lanes).  RecSplit takes the second word for the bucket and the first for the tree
// (RecSplit.hpp:116-120).
void spooky_u64(uint64_t key, uint64_t& bucket_hash, uint64_t& tree_hash) {
    const uint64_t kc = 0xdeadbeefdeadbeefULL;
    uint64_t h[4] = {0, 0, kc + key, kc + (8ULL << 56)};
    static const int r[11] = {15, 52, 26, 51, 28, 9, 47, 54, 32, 25, 63};
    for (int s = 0; s < 11; ++s) {
        uint64_t& x = h[(3 + s) & 3];
        uint64_t& y = h[(2 + s) & 3];
        x ^= y;
        y = rot64(y, r[s]);
        x += y;
    }
    bucket_hash = h[1];
    tree_hash = h[0];
}

inline uint64_t mix64(uint64_t z) {          // Stafford's 13th finalizer (RecSplit.hpp:88-92)
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
}
inline uint64_t scale48(uint64_t x, uint64_t n) { return ((x & ((1ULL << 48) - 1)) * n) >> 48; }

// Evaluation of a serialized RecSplit<8> (bucket size as stored).  Buckets come from a double Elias-Fano list
// (DoubleEF.hpp:93-127), decoded here once into plain arrays; every bucket is a splitting tree whose nodes store a
// Golomb-Rice coded seed: all fixed parts first, then all unary parts (RiceBitVector.hpp:118-176).
struct LengthHash {
    static constexpr uint64_t kLeaf = 8, kLower = 32, kUpper = 96;   // SplittingStrategy<8> (RecSplit.hpp:156-158)
    uint64_t keys = 0, nbuckets = 0;
    std::vector<uint64_t> words;            // descriptors + two zero words
    std::vector<uint64_t> cum, pos;         // nbuckets + 1 each
    std::vector<uint32_t> memo;             // per subtree size: rice parameter << 27 | nodes << 16 | fixed bits
    std::string name;

    void fail(const char* what) const { throw std::runtime_error(name + ": length hash: " + what); }

    void fill_memo(uint64_t m_max) {
        static const uint8_t leaf_rice[9] = {0, 0, 0, 1, 3, 4, 5, 7, 8};     // RecSplit.hpp:130
        memo.assign(std::max<uint64_t>(m_max, kLeaf) + 1, 0);
        for (uint64_t s = 0; s <= kLeaf; ++s) memo[s] = (uint32_t)leaf_rice[s] << 27 | (uint32_t)(s > 1) << 16 | leaf_rice[s];
        for (uint64_t m = kLeaf + 1; m <= m_max; ++m) {                         // RecSplit.hpp:211-243
            uint64_t fanout, unit;
            if (m > kUpper) { unit = kUpper * ((m / 2 + kUpper - 1) / kUpper); fanout = 2; }
            else if (m > kLower) { unit = kLower; fanout = (m + kLower - 1) / kLower; }
            else { unit = kLeaf; fanout = (m + kLeaf - 1) / kLeaf; }
            const uint64_t last = m - (fanout - 1) * unit;
            double sqrt_prod = 1;
            for (uint64_t i = 0; i + 1 < fanout; ++i) sqrt_prod *= sqrt((double)unit);
            sqrt_prod *= sqrt((double)last);
            const double p = sqrt((double)m) / (pow(2 * M_PI, (fanout - 1.) / 2) * sqrt_prod);
            uint32_t rice = (uint32_t)ceil(log2(-log((sqrt(5) + 1) / 2) / log1p(-p)));
            uint32_t bits = rice + (uint32_t)(fanout - 1) * (memo[unit] & 0xFFFF) + (memo[last] & 0xFFFF);
            uint32_t nodes = 1 + (uint32_t)(fanout - 1) * ((memo[unit] >> 16) & 0x7FF) + ((memo[last] >> 16) & 0x7FF);
            if (rice > 31 || bits > 0xFFFF || nodes > 0x7FF) fail("splitting table overflow");
            memo[m] = rice << 27 | nodes << 16 | bits;
        }
    }
    uint32_t rice(uint64_t m) const { return memo[m] >> 27; }
    uint32_t fixed_bits(uint64_t m) const { return memo[m] & 0xFFFF; }
    uint32_t nodes(uint64_t m) const { return (memo[m] >> 16) & 0x7FF; }

    uint64_t bits_at(uint64_t bit, uint32_t width) const {
        if (!width) return 0;
        const uint64_t w = bit >> 6, s = bit & 63;
        if (w + 1 >= words.size()) fail("descriptor read out of range");
        uint64_t v = words[w] >> s;
        if (s + width > 64) v |= words[w + 1] << (64 - s);
        return v & ((1ULL << width) - 1);
    }
    uint64_t next_one(uint64_t bit) const {      // position of the first 1 at or after `bit`
        uint64_t w = bit >> 6;
        if (w >= words.size()) fail("descriptor read out of range");
        uint64_t x = words[w] & (~0ULL << (bit & 63));
        while (!x) {
            if (++w >= words.size()) fail("descriptor read out of range");
            x = words[w];
        }
        return (w << 6) + (uint64_t)__builtin_ctzll(x);
    }
    struct Reader {
        const LengthHash& h;
        uint64_t fix, un;
        uint64_t next(uint32_t log2g) {
            const uint64_t one = h.next_one(un);
            const uint64_t q = one - un;
            un = one + 1;
            const uint64_t v = q << log2g | h.bits_at(fix, log2g);
            fix += log2g;
            return v;
        }
        void skip(uint64_t n_nodes, uint64_t n_fixed) {
            for (uint64_t k = 0; k < n_nodes; ++k) un = h.next_one(un) + 1;
            fix += n_fixed;
        }
    };

    uint64_t operator()(uint64_t key) const;      // RecSplit.hpp:377-426
};


uint64_t LengthHash::operator()(uint64_t key) const {
    static const uint64_t seed[20] = {                        // RecSplit.hpp:76-78, one per tree level
        0x106393c187cae21aULL, 0x6453cec3f7376937ULL, 0x643e521ddbd2be98ULL, 0x3740c6412f6572cbULL, 0x717d47562f1ce470ULL,
        0x4cd6eb4c63befb7cULL, 0x9bfd8c5e18c8da73ULL, 0x082f20e10092a9a3ULL, 0x2ada2ce68d21defcULL, 0xe33cb4f3e7c6466bULL,
        0x3980be458c509c59ULL, 0xc466fd9584828e8cULL, 0x45f0aabe1a61ede6ULL, 0xf6e7b8b33ad9b98dULL, 0x4ef95e25f4b4983dULL,
        0x81175195173b92d3ULL, 0x4e50927d8dd15978ULL, 0x1ea2099d1fafae7fULL, 0x425c8a06fbaaa815ULL, 0xcd4216006c74052aULL};
    uint64_t hb, ht;
    spooky_u64(key, hb, ht);
    const uint64_t b = (uint64_t)(((unsigned __int128)hb * nbuckets) >> 64);
    uint64_t base = cum[b];
    uint64_t m = cum[b + 1] - base;
    if (m >= memo.size()) fail("bucket larger than the splitting table");
    Reader rd{*this, pos[b], pos[b] + fixed_bits(m)};
    int level = 0;
    while (m > kUpper) {                                      // two-way splits on multiples of kUpper
        const uint64_t d = rd.next(rice(m));
        const uint64_t x = scale48(mix64(ht + d + seed[level]), m);
        const uint64_t split = ((m / 2 + kUpper - 1) / kUpper) * kUpper;
        if (x < split) {
            m = split;
        } else {
            rd.skip(nodes(split), fixed_bits(split));
            m -= split;
            base += split;
        }
        ++level;
    }
    if (m > kLower) {                                         // parts of kLower keys
        const uint64_t d = rd.next(rice(m));
        const uint64_t part = scale48(mix64(ht + d + seed[level]), m) / kLower;
        m = std::min<uint64_t>(kLower, m - part * kLower);
        base += kLower * part;
        if (part) rd.skip((uint64_t)nodes(kLower) * part, (uint64_t)fixed_bits(kLower) * part);
        ++level;
    }
    if (m > kLeaf) {                                          // parts of kLeaf keys
        const uint64_t d = rd.next(rice(m));
        const uint64_t part = scale48(mix64(ht + d + seed[level]), m) / kLeaf;
        m = std::min<uint64_t>(kLeaf, m - part * kLeaf);
        base += kLeaf * part;
        if (part) rd.skip(part, (uint64_t)fixed_bits(kLeaf) * part);
        ++level;
    }
    const uint64_t d = rd.next(rice(m));                      // the leaf's bijection
    return base + scale48(mix64(ht + d + seed[level]), m);
}

void read_length_hash(Cursor& c, LengthHash& h) {
    h.name = c.name;
    if (c.u64("mph leaf size") != LengthHash::kLeaf) h.fail("leaf size is not 8");
    const uint64_t bucket_size = c.u64("mph bucket size");
    h.keys = c.u64("mph key count");
    if (!bucket_size) h.fail("bucket size 0");
    h.nbuckets = std::max<uint64_t>(1, (h.keys + bucket_size - 1) / bucket_size);
    auto words_of = [&](std::vector<uint64_t>& v, const char* what, size_t pad) {
        const uint64_t n = c.u64(what);
        if (n > c.size / 8) h.fail("implausible vector size");
        const uint8_t* p = c.bytes(n * 8, what);
        v.assign(n + pad, 0);
        if (n) memcpy(v.data(), p, n * 8);
    };
    words_of(h.words, "mph descriptors", 2);
    // double Elias-Fano list: (keys before bucket i, first descriptor bit of bucket i), i in [0, nbuckets]
    const uint64_t nb = c.u64("ef buckets");
    const uint64_t u_cum = c.u64("ef u_cum_keys"), u_pos = c.u64("ef u_position");
    const int64_t min_delta = (int64_t)c.u64("ef cum_keys_min_delta"), min_diff = (int64_t)c.u64("ef min_diff");
    const uint64_t bits_per_key = c.u64("ef bits_per_key");
    if (nb != h.nbuckets) h.fail("bucket count mismatch");
    auto log2_floor = [](uint64_t x) { return 63 - __builtin_clzll(x); };
    const uint32_t l_pos = u_pos / (nb + 1) == 0 ? 0 : log2_floor(u_pos / (nb + 1));
    const uint32_t l_cum = u_cum / (nb + 1) == 0 ? 0 : log2_floor(u_cum / (nb + 1));
    std::vector<uint64_t> lower, up_cum, up_pos, jump;
    words_of(lower, "ef lower bits", 2);
    words_of(up_cum, "ef upper bits (keys)", 0);
    words_of(up_pos, "ef upper bits (positions)", 0);
    words_of(jump, "ef jump table", 0);                       // only needed for random access: not used
    auto ones = [&](const std::vector<uint64_t>& bits, std::vector<uint64_t>& out) {
        out.clear();
        for (uint64_t w = 0; w < bits.size() && out.size() <= nb; ++w)
            for (uint64_t x = bits[w]; x && out.size() <= nb; x &= x - 1) out.push_back((w << 6) + (uint64_t)__builtin_ctzll(x));
        if (out.size() != nb + 1) h.fail("Elias-Fano upper bits are short");
    };
    std::vector<uint64_t> oc, op;
    ones(up_cum, oc);
    ones(up_pos, op);
    auto low_at = [&](uint64_t bit, uint32_t width) -> uint64_t {
        if (!width) return 0;
        const uint64_t w = bit >> 6, s = bit & 63;
        uint64_t v = lower[w] >> s;
        if (s + width > 64) v |= lower[w + 1] << (64 - s);
        return v & ((1ULL << width) - 1);
    };
    if (((nb + 1) * (uint64_t)(l_cum + l_pos) + 63) / 64 > lower.size()) h.fail("Elias-Fano lower bits are short");
    h.cum.resize(nb + 1);
    h.pos.resize(nb + 1);
    uint64_t m_max = 0;
    for (uint64_t i = 0; i <= nb; ++i) {
        const uint64_t at = i * (l_cum + l_pos);
        h.cum[i] = (((oc[i] - i) << l_cum) | low_at(at, l_cum)) + (uint64_t)((int64_t)i * min_delta);
        h.pos[i] = (uint64_t)((int64_t)((((op[i] - i) << l_pos) | low_at(at + l_cum, l_pos))) + (int64_t)i * min_diff +
                              (int64_t)(bits_per_key * h.cum[i] >> 20));
        if (i) {
            if (h.cum[i] < h.cum[i - 1]) h.fail("bucket sizes are not monotone");
            m_max = std::max(m_max, h.cum[i] - h.cum[i - 1]);
        }
    }
    if (h.cum[nb] != h.keys) h.fail("bucket sizes do not add up to the key count");
    if (m_max >= 3000) h.fail("bucket larger than RecSplit allows");        // MAX_BUCKET_SIZE, RecSplit.hpp:48
    h.fill_memo(m_max);
}

// DirectAccessibleGammaCode<SelectSdvec<>>: value + 1 = an implicit leading 1 followed by `hi` stored bits
void read_gamma_array(Cursor& c, std::vector<uint64_t>& out, const char* what) {
    const uint64_t num = c.u64(what), nwords = c.u64(what);
    if (nwords > c.size / 8) throw std::runtime_error(c.name + ": implausible word count in " + what);
    const uint8_t* arr = c.bytes(nwords * 8, what);
    std::vector<uint64_t> starts;
    read_sd_vector(c, starts, what);                          // a 1 where every codeword starts + a sentinel
    if (starts.size() != num + 1) throw std::runtime_error(c.name + ": gamma array marks do not match its size in " + what);
    auto word = [&](uint64_t i) -> uint64_t { uint64_t v = 0; if (i < nwords) memcpy(&v, arr + 8 * i, 8); return v; };
    out.resize(num);
    for (uint64_t i = 0; i < num; ++i) {
        const uint64_t hi = starts[i + 1] - starts[i] - 1;
        if (hi > 63) throw std::runtime_error(c.name + ": gamma codeword too long in " + what);
        uint64_t val = 1ULL << hi;
        if (hi) {
            const uint64_t bit = starts[i] - i, w = bit >> 6, s = bit & 63;
            uint64_t x = word(w) >> s;
            if (s + hi > 64) x |= word(w + 1) << (64 - s);
            val += x & ((1ULL << hi) - 1);
        }
        out[i] = val - 1;
    }
}

}  // namespace


void read_shaped_slp(const std::string& path, FlatSlp& o) {
    MappedFile mf(path);
    Cursor c{mf.p, mf.size, 0, path};
    o = FlatSlp();
    o.alph_size = c.u64("alphSize");
    if (o.alph_size == 0 || o.alph_size > 256) throw std::runtime_error(path + ": implausible alphabet size");
    const uint8_t* a = c.bytes(o.alph_size, "alph");
    o.alph.assign(a, a + o.alph_size);
    const uint64_t A = o.alph_size;
    LengthHash mph;
    read_length_hash(c, mph);
    std::vector<uint64_t> seq_end, group_start;
    const uint64_t seq_universe = read_sd_vector(c, seq_end, "seqSBV");          // prefix sums of the root lengths
    const uint64_t NR = read_sd_vector(c, group_start, "slpDivSel");            // 1 = first rule of a group
    PackedView bal_bv = read_int_vector(c, 1, "balBv");                         // per non-first rule: left length changes
    read_int_vector(c, 64, "balBvRank");
    std::vector<uint64_t> seq_off, child_off, left_len_tab;
    read_gamma_array(c, seq_off, "vlcSeq");
    read_gamma_array(c, child_off, "vlc");
    read_gamma_array(c, left_len_tab, "bal");
    if (c.off != c.size) throw std::runtime_error(path + ": " + std::to_string(c.size - c.off) + " trailing bytes");
    if (seq_universe == 0 || seq_end.empty() || seq_end.back() + 1 != seq_universe)
        throw std::runtime_error(path + ": root prefix sums do not end at the text length");
    o.text_len = seq_universe - 1;
    const uint64_t G = group_start.size(), NS = seq_end.size();
    if (seq_off.size() != NS || child_off.size() != 2 * NR) throw std::runtime_error(path + ": array sizes do not match the rule counts");
    if (NR == 0 || G == 0 || group_start[0] != 0 || G != mph.keys) throw std::runtime_error(path + ": groups do not match the length hash");
    if (NR + NS + A >= 0xFFFFFFF0ULL) throw std::runtime_error(path + ": too many variables");
    group_start.push_back(NR);

    // left length and group of every rule (SelfShapedSlp.hpp:113-116 read sequentially: a new table entry at every
    // group start and wherever balBv is set)
    std::vector<uint64_t> left_len(NR);
    std::vector<uint32_t> group_of(NR);
    {
        uint64_t g = 0, z = 0, e = 0;
        for (uint64_t r = 0; r < NR; ++r) {
            if (r == group_start[g + 1]) ++g;
            if (r == group_start[g]) { if (r) ++e; }
            else {
                if (z >= bal_bv.size()) throw std::runtime_error(path + ": balBv is short");
                if (bal_bv.get(z++)) ++e;
            }
            if (e >= left_len_tab.size()) throw std::runtime_error(path + ": left-length table is short");
            left_len[r] = left_len_tab[e];
            group_of[r] = (uint32_t)g;
        }
    }
    // expansion length of every group: walk down from the roots; only lengths that occur are ever hashed
    std::vector<uint64_t> group_len(G, 0);
    std::vector<uint32_t> work;
    auto group_of_len = [&](uint64_t len) -> uint64_t {
        const uint64_t g = mph(len);
        if (g >= G) throw std::runtime_error(path + ": length hash out of range");
        return g;
    };
    auto visit = [&](uint64_t len) {
        if (len <= 1) return;
        const uint64_t g = group_of_len(len);
        if (group_len[g] == len) return;
        if (group_len[g]) throw std::runtime_error(path + ": two expansion lengths hash to one group");
        group_len[g] = len;
        work.push_back((uint32_t)g);
    };
    for (uint64_t i = 0; i < NS; ++i) {
        const uint64_t len = seq_end[i] - (i ? seq_end[i - 1] : 0);
        if (len == 0) throw std::runtime_error(path + ": empty root");
        visit(len);
    }
    while (!work.empty()) {
        const uint32_t g = work.back();
        work.pop_back();
        for (uint64_t r = group_start[g]; r < group_start[g + 1]; ++r) {
            if (left_len[r] == 0 || left_len[r] >= group_len[g]) throw std::runtime_error(path + ": left length out of range");
            visit(left_len[r]);
            visit(group_len[g] - left_len[r]);
        }
    }
    auto var_of = [&](uint64_t len, uint64_t off) -> uint32_t {
        if (len == 1) {
            if (off >= A) throw std::runtime_error(path + ": terminal out of range");
            return (uint32_t)off;
        }
        const uint64_t g = group_of_len(len);
        if (group_len[g] != len || off >= group_start[g + 1] - group_start[g]) throw std::runtime_error(path + ": child out of range");
        return (uint32_t)(A + group_start[g] + off);
    };
    o.left.assign(NR, 0);
    o.right.assign(NR, 0);
    o.explen.assign(NR, 2);          // rules of a group no root reaches keep a harmless dummy body
    for (uint64_t r = 0; r < NR; ++r) {
        const uint64_t len = group_len[group_of[r]];
        if (!len) continue;
        o.explen[r] = len;
        o.left[r] = var_of(left_len[r], child_off[2 * r]);
        o.right[r] = var_of(len - left_len[r], child_off[2 * r + 1]);
    }
    // the root sequence becomes a balanced binary tree of extra rules (the text is the concatenation of the roots)
    std::vector<uint32_t> level(NS);
    std::vector<uint64_t> level_len(NS);
    for (uint64_t i = 0; i < NS; ++i) {
        level_len[i] = seq_end[i] - (i ? seq_end[i - 1] : 0);
        level[i] = var_of(level_len[i], seq_off[i]);
    }
    while (level.size() > 1) {
        std::vector<uint32_t> up;
        std::vector<uint64_t> up_len;
        for (size_t i = 0; i + 1 < level.size(); i += 2) {
            o.left.push_back(level[i]);
            o.right.push_back(level[i + 1]);
            o.explen.push_back(level_len[i] + level_len[i + 1]);
            up.push_back((uint32_t)(A + o.left.size() - 1));
            up_len.push_back(o.explen.back());
        }
        if (level.size() & 1) { up.push_back(level.back()); up_len.push_back(level_len.back()); }
        level.swap(up);
        level_len.swap(up_len);
    }
    o.num_rules = o.left.size();
    o.start_var = level[0];
    if (level_len[0] != o.text_len) throw std::runtime_error(path + ": roots do not add up to the text length");
}

}  // namespace moni
