// Input preparation (not on the measured path): writes a straight-line grammar for a FASTA text
// in the "Bigrepair" .R/.C layout that the reference's SlpEncBuild consumes
// (format: /root/reference/thirdparty/ShapedSlp/NaiveSlp.hpp:151-213 — .R = int32 alphabet size
// followed by {u32 left,u32 right} rules, terminals are byte values < 256; .C = u32 start sequence).
//
// The reference builds this grammar with bigrepair, which is absent from the checkout
// (SURVEY.md §2), so the harness makes its own: content-defined chunks of the text are
// de-duplicated (exact compare), every distinct chunk becomes a balanced subtree whose
// (left,right) pairs are hash-consed, and the chunk sequence is the start sequence C.
// MS/MEM results do not depend on the grammar, only on the text it expands to.
//
// usage: mkgrammar <in.fa> <out-prefix> [--trivial]
//   --trivial : no rules, C = the text (one u32 per character)
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cctype>
#include <string>
#include <unordered_map>
#include <vector>

static std::vector<uint8_t> read_fasta_text(const char* path) {
    FILE* f = fopen(path, "rb");
    if (!f) { fprintf(stderr, "mkgrammar: cannot open %s\n", path); exit(1); }
    std::vector<uint8_t> text;
    std::vector<char> buf(1 << 22);
    bool in_header = false, at_line_start = true;
    size_t got;
    while ((got = fread(buf.data(), 1, buf.size(), f)) > 0) {
        for (size_t i = 0; i < got; ++i) {
            char c = buf[i];
            if (c == '\n') { in_header = false; at_line_start = true; continue; }
            if (at_line_start && c == '>') { in_header = true; at_line_start = false; continue; }
            at_line_start = false;
            if (in_header || c == '\r') continue;
            text.push_back((uint8_t)toupper((unsigned char)c));
        }
    }
    fclose(f);
    return text;
}

struct PairHash {
    size_t operator()(uint64_t k) const {
        k ^= k >> 33; k *= 0xff51afd7ed558ccdULL; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ULL; k ^= k >> 33;
        return (size_t)k;
    }
};

int main(int argc, char** argv) {
    if (argc < 3) { fprintf(stderr, "usage: mkgrammar <in.fa> <out-prefix> [--trivial]\n"); return 2; }
    bool trivial = argc > 3 && !strcmp(argv[3], "--trivial");
    std::vector<uint8_t> T = read_fasta_text(argv[1]);
    const size_t n = T.size();
    std::vector<uint32_t> rules;  // left,right interleaved
    std::vector<uint32_t> C;
    if (trivial) {
        C.resize(n);
        for (size_t i = 0; i < n; ++i) C[i] = T[i];
    } else {
        const int W = 12; const uint64_t P = 48; const size_t MINC = 16, MAXC = 1024;
        std::unordered_map<uint64_t, uint32_t, PairHash> pair_id;
        struct Chunk { size_t off; uint32_t len; uint32_t var; };
        std::unordered_multimap<uint64_t, Chunk, PairHash> chunk_id;
        pair_id.reserve(1 << 20);
        auto var_of_pair = [&](uint32_t l, uint32_t r) -> uint32_t {
            uint64_t key = ((uint64_t)l << 32) | r;
            auto it = pair_id.find(key);
            if (it != pair_id.end()) return it->second;
            uint64_t id = 256 + rules.size() / 2;
            if (id >= 0xFFFFFFF0ULL) { fprintf(stderr, "mkgrammar: too many rules\n"); exit(1); }
            rules.push_back(l); rules.push_back(r);
            pair_id.emplace(key, (uint32_t)id);
            return (uint32_t)id;
        };
        std::vector<uint32_t> lvl;
        auto var_of_chunk = [&](size_t off, size_t len) -> uint32_t {
            uint64_t h = 1469598103934665603ULL;
            for (size_t i = 0; i < len; ++i) { h ^= T[off + i]; h *= 1099511628211ULL; }
            auto range = chunk_id.equal_range(h);
            for (auto it = range.first; it != range.second; ++it)
                if (it->second.len == len && !memcmp(&T[it->second.off], &T[off], len)) return it->second.var;
            lvl.assign(T.begin() + off, T.begin() + off + len);
            while (lvl.size() > 1) {
                size_t m = lvl.size() / 2, odd = lvl.size() & 1;
                for (size_t i = 0; i < m; ++i) lvl[i] = var_of_pair(lvl[2 * i], lvl[2 * i + 1]);
                if (odd) lvl[m] = lvl[lvl.size() - 1];
                lvl.resize(m + odd);
            }
            uint32_t v = lvl[0];
            chunk_id.emplace(h, Chunk{off, (uint32_t)len, v});
            return v;
        };
        // content-defined boundaries: polynomial rolling hash of the last W characters
        uint64_t pw = 1; const uint64_t B = 0x100000001b3ULL;
        for (int i = 0; i < W; ++i) pw *= B;
        uint64_t h = 0; size_t start = 0;
        for (size_t i = 0; i < n; ++i) {
            h = h * B + T[i] + 1;
            if (i >= (size_t)W) h -= pw * (T[i - W] + 1);
            size_t len = i + 1 - start;
            bool cut = (len >= MINC && ((h >> 17) % P) == 0) || len >= MAXC;
            if (cut || i + 1 == n) { C.push_back(var_of_chunk(start, len)); start = i + 1; }
        }
    }
    std::string pre(argv[2]);
    FILE* fr = fopen((pre + ".R").c_str(), "wb");
    FILE* fc = fopen((pre + ".C").c_str(), "wb");
    if (!fr || !fc) { fprintf(stderr, "mkgrammar: cannot write outputs\n"); return 1; }
    int32_t alph = 256;
    fwrite(&alph, 4, 1, fr);
    if (!rules.empty()) fwrite(rules.data(), 4, rules.size(), fr);
    if (!C.empty()) fwrite(C.data(), 4, C.size(), fc);
    fclose(fr); fclose(fc);
    fprintf(stderr, "mkgrammar: text %zu chars, %zu rules, |C| = %zu\n", n, rules.size() / 2, C.size());
    return 0;
}
