// Exporter: decoded builder files -> HBM image.  See device_image.hpp for the layout and the
// reference functions each table replaces.
#include "device_image.hpp"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <thread>

#include <sys/mman.h>

#include "index_files.hpp"

namespace moni {
namespace {

inline void put40(uint8_t* d, uint64_t v) {
    d[0] = (uint8_t)v; d[1] = (uint8_t)(v >> 8); d[2] = (uint8_t)(v >> 16); d[3] = (uint8_t)(v >> 24); d[4] = (uint8_t)(v >> 32);
}

double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// fn(t) for t in [0, T) on T threads (the caller's thread takes t = 0); the first exception is rethrown
template <typename F>
void run_parts(unsigned T, F&& fn) {
    std::vector<std::thread> th;
    std::vector<std::exception_ptr> err(T);
    auto guarded = [&](unsigned t) {
        try { fn(t); } catch (...) { err[t] = std::current_exception(); }
    };
    for (unsigned t = 1; t < T; ++t) th.emplace_back(guarded, t);
    guarded(0);
    for (auto& x : th) x.join();
    for (auto& e : err) if (e) std::rethrow_exception(e);
}

// v.assign(n, val) for the image's large tables: the cost of a fresh multi-GB vector is its page faults, so the pages are
// touched from P threads first (and offered to the kernel as huge pages); the assign then runs at memory speed.
template <typename T>
void assign_large(std::vector<T>& v, size_t n, const T& val, unsigned P) {
    v.clear();
    v.reserve(n);
    const size_t bytes = n * sizeof(T);
    char* base = reinterpret_cast<char*>(v.data());
    if (P > 1 && bytes >= (32u << 20) && base) {
        const uintptr_t lo = ((uintptr_t)base + (2u << 20) - 1) & ~(uintptr_t)((2u << 20) - 1), hi = ((uintptr_t)base + bytes) & ~(uintptr_t)((2u << 20) - 1);
        if (hi > lo) madvise((void*)lo, hi - lo, MADV_HUGEPAGE);           // best effort
        run_parts(P, [&](unsigned t) {
            for (size_t o = bytes / P * t, e = t + 1 == P ? bytes : bytes / P * (t + 1); o < e; o += 4096) base[o] = 0;
        });
    }
    v.assign(n, val);
}

// The row passes below are scans with a carried state (characters of each table symbol seen so far, runs of each symbol
// closed so far, ...).  Every such state is a sum or a last-index over the rows before: the rows are cut into T contiguous
// parts, one cheap pass gives each part's contribution, an exclusive scan over the parts gives each part's initial state, and
// the parts then run the same loop side by side.  The LF cursor (first row that starts after an image start) is monotone and
// is found by one binary search per part and symbol.
struct PartSums {
    uint64_t len[4] = {0, 0, 0, 0};              // characters of table symbol tc in rows of the part
    uint64_t first_piece[4];                     // first row of the part that starts a run of tc   (~0 = none)
    uint64_t last_piece[4];                      // last row of the part that ends a run of tc      (~0 = none)
    std::vector<uint64_t> runs, chars;           // per byte value: runs closed / characters in closed runs
    PartSums() : runs(256, 0), chars(256, 0) {
        for (int c = 0; c < 4; ++c) first_piece[c] = last_piece[c] = ~0ULL;
    }
};

}  // namespace

void build_move_table(const FlatRlbwt& b, const ExportOptions& opt, HostImage& img) {
    const double t0 = now_s();
    const uint64_t R = b.R, n = b.n;
    if (opt.max_scan < 2) throw std::runtime_error("exporter: max_scan must be >= 2");
    for (uint64_t k = 0; k < R; ++k)
        if (b.samples_start[k] >= (1ULL << 40) || b.samples_last[k] >= (1ULL << 40))
            throw std::runtime_error("exporter: SA sample does not fit 40 bits");

    // ---- symbol tables -----------------------------------------------------------------------
    SymbolTables& S = img.sym;
    memset(&S, 0, sizeof(S));
    memset(S.tcode, 0xFF, sizeof(S.tcode));
    {
        std::vector<int> present;
        for (int c = 0; c < 256; ++c) {
            S.F[c] = b.F[c];
            S.count[c] = b.count[c];
            if (b.nruns[c] >= 0xFFFFFFFFULL) throw std::runtime_error("exporter: too many runs of one symbol");
            S.nruns[c] = (uint32_t)b.nruns[c];
            if ((b.count[c] == 0) != (b.nruns[c] == 0)) throw std::runtime_error("exporter: count/run mismatch");
            if (b.count[c]) {
                present.push_back(c);
                // thr_bv::rank can never exceed the number of gaps between c-runs in a usable index
                // (otherwise the reference itself selects past the last c-run).
                if (b.thr[c].size() > b.nruns[c] - 1)
                    throw std::runtime_error("exporter: more stored thresholds than gaps between runs of a symbol");
            } else if (!b.thr[c].empty()) {
                throw std::runtime_error("exporter: thresholds for an absent symbol");
            }
        }
        std::vector<int> by_runs = present;
        std::stable_sort(by_runs.begin(), by_runs.end(), [&](int x, int y) { return b.nruns[x] > b.nruns[y]; });
        if (by_runs.size() > 4) by_runs.resize(4);
        std::sort(by_runs.begin(), by_runs.end());
        img.hdr.n_table_syms = (uint32_t)by_runs.size();
        for (size_t i = 0; i < by_runs.size(); ++i) {
            S.tcode[by_runs[i]] = (uint8_t)i;
            img.hdr.table_syms[i] = (uint8_t)by_runs[i];
        }
        uint64_t sb = 0, cb = 0;
        for (int c = 0; c < 256; ++c) {
            S.slot_base[c] = sb;
            S.crun_base[c] = cb;
            if (b.count[c]) { sb += b.nruns[c] + 1; cb += b.nruns[c]; }
        }
        const unsigned Pz = opt.threads ? opt.threads : std::min(32u, std::max(1u, std::thread::hardware_concurrency()));
        assign_large(img.slots, sb, GapSlot{0, 0, 0, 0}, Pz);
        assign_large(img.crun_row, cb, (uint32_t)0, Pz);
    }
    const uint8_t* tcode = S.tcode;

    // ---- rows: cut runs longer than kMaxPiece, then balance ------------------------------------
    std::vector<uint64_t> pos;      // row starts
    std::vector<uint32_t> run;      // run of each row
    pos.reserve(R + R / 4 + 16);
    run.reserve(R + R / 4 + 16);
    for (uint64_t k = 0; k < R; ++k)
        for (uint64_t s = b.starts[k]; s < b.starts[k + 1]; s += kMaxPiece) { pos.push_back(s); run.push_back((uint32_t)k); }

    unsigned T = opt.threads ? opt.threads : std::min(32u, std::max(1u, std::thread::hardware_concurrency()));
    auto parts_for = [&](uint64_t rows) -> unsigned {      // explicit thread counts are honoured down to one row per part (tests)
        if (opt.threads) return (unsigned)std::max<uint64_t>(1, std::min<uint64_t>(opt.threads, rows));
        return (unsigned)std::max<uint64_t>(1, std::min<uint64_t>(T, rows / 65536));
    };

    std::vector<uint64_t> cuts;     // absolute positions of new row starts found in one round
    uint32_t worst = 0;
    for (uint32_t round = 0;; ++round) {
        const uint64_t nr = pos.size();
        const unsigned P = parts_for(nr);
        std::vector<PartSums> part(P);
        std::vector<std::vector<uint64_t>> part_cuts(P);
        std::vector<uint32_t> part_worst(P, 0);
        auto row_len = [&](uint64_t i) { return (i + 1 < nr ? pos[i + 1] : n) - pos[i]; };
        run_parts(P, [&](unsigned t) {
            for (uint64_t i = nr * t / P, e = nr * (t + 1) / P; i < e; ++i) {
                const uint8_t tc = tcode[b.heads[run[i]]];
                if (tc != 0xFF) part[t].len[tc] += row_len(i);
            }
        });
        run_parts(P, [&](unsigned t) {
            uint64_t seen[4] = {0, 0, 0, 0};   // # of the table symbol before the current row
            for (unsigned u = 0; u < t; ++u)
                for (int c = 0; c < 4; ++c) seen[c] += part[u].len[c];
            uint64_t q[4] = {~0ULL, ~0ULL, ~0ULL, ~0ULL};      // first row whose start is > the image start of the current row
            std::vector<uint64_t>& my_cuts = part_cuts[t];
            uint32_t my_worst = 0;
            for (uint64_t i = nr * t / P, e = nr * (t + 1) / P; i < e; ++i) {
                const uint8_t tc = tcode[b.heads[run[i]]];
                if (tc == 0xFF) continue;
                const uint64_t len = row_len(i);
                const uint64_t pi = S.F[b.heads[run[i]]] + seen[tc];
                seen[tc] += len;
                uint64_t tt = q[tc];
                if (tt == ~0ULL) tt = (uint64_t)(std::upper_bound(pos.begin(), pos.end(), pi) - pos.begin());
                while (tt < nr && pos[tt] <= pi) ++tt;
                uint32_t inside = 0;
                while (tt < nr && pos[tt] < pi + len) {
                    ++inside;
                    if (inside % opt.max_scan == 0 && round < opt.max_rounds) my_cuts.push_back(pos[i] + (pos[tt] - pi));
                    ++tt;
                }
                my_worst = std::max(my_worst, inside);
                q[tc] = tt;
            }
            part_worst[t] = my_worst;
        });
        cuts.clear();
        worst = 0;
        for (unsigned t = 0; t < P; ++t) {
            cuts.insert(cuts.end(), part_cuts[t].begin(), part_cuts[t].end());
            worst = std::max(worst, part_worst[t]);
        }
        if (opt.verbose)
            fprintf(stderr, "[moni-gpu] balance round %u: rows %llu, new cuts %llu, widest image %u rows\n", round,
                    (unsigned long long)nr, (unsigned long long)cuts.size(), worst);
        if (cuts.empty()) break;
        // merge the (sorted, distinct, strictly interior) cuts into the rows
        std::vector<uint64_t> npos;
        std::vector<uint32_t> nrun;
        npos.reserve(nr + cuts.size());
        nrun.reserve(nr + cuts.size());
        size_t ci = 0;
        for (uint64_t i = 0; i < nr; ++i) {
            npos.push_back(pos[i]);
            nrun.push_back(run[i]);
            const uint64_t end = (i + 1 < nr ? pos[i + 1] : n);
            while (ci < cuts.size() && cuts[ci] < end) {
                if (cuts[ci] > npos.back()) { npos.push_back(cuts[ci]); nrun.push_back(run[i]); }
                ++ci;
            }
        }
        pos.swap(npos);
        run.swap(nrun);
    }
    const uint64_t nr = pos.size();
    if (nr >= 0xFFFFFFF0ULL) throw std::runtime_error("exporter: too many rows");
    img.hdr.n = n;
    img.hdr.R = R;
    img.hdr.n_rows = nr;
    img.hdr.max_scan = worst;
    img.hdr.init_sample = (b.samples_last[R - 1] + 1) % n;
    img.hdr.init_row = (uint32_t)(nr - 1);
    img.hdr.init_off = (uint32_t)(n - 1 - pos[nr - 1]);

    // ---- fill rows -----------------------------------------------------------------------------
    assign_large(img.rows, nr + 1, MoveRow{}, T);
    img.row_pos.assign(pos.begin(), pos.end());
    img.row_pos.push_back(n);
    const int nts = (int)img.hdr.n_table_syms;
    const unsigned P = parts_for(nr);
    auto part_lo = [&](unsigned t) { return nr * t / P; };
    std::vector<PartSums> part(P);
    run_parts(P, [&](unsigned t) {                 // each part's contribution to the carried states
        PartSums& ps = part[t];
        for (uint64_t i = part_lo(t), e = part_lo(t + 1); i < e; ++i) {
            const uint32_t k = run[i];
            const uint8_t h = b.heads[k];
            const uint8_t th = tcode[h];
            if (th != 0xFF) {
                ps.len[th] += img.row_pos[i + 1] - pos[i];
                if ((i == 0 || run[i - 1] != k) && ps.first_piece[th] == ~0ULL) ps.first_piece[th] = i;
            }
            if (i + 1 == nr || run[i + 1] != k) {
                ps.runs[h] += 1;
                ps.chars[h] += b.starts[k + 1] - b.starts[k];
                if (th != 0xFF) ps.last_piece[th] = i;
            }
        }
    });
    run_parts(P, [&](unsigned t) {
        // backward pass: row of the first piece of the next c-run after row i (for dist_dn)
        uint64_t nxt[4] = {~0ULL, ~0ULL, ~0ULL, ~0ULL};
        for (int c = 0; c < 4; ++c)
            for (unsigned u = t + 1; u < P && nxt[c] == ~0ULL; ++u) nxt[c] = part[u].first_piece[c];
        // store the clipped distance directly into the row while walking backwards
        for (uint64_t ii = part_lo(t + 1), lo = part_lo(t); ii-- > lo;) {
            MoveRow& r = img.rows[ii];
            const uint8_t h = b.heads[run[ii]];
            const uint8_t th = tcode[h];
            if (th != 0xFF) {
                for (int c = 0, s = 0; c < nts; ++c) {
                    if (c == th) continue;
                    const uint64_t d = nxt[c] == ~0ULL ? kFar : nxt[c] - ii;
                    r.dist_dn[s++] = (uint8_t)std::min<uint64_t>(d, kFar);
                }
                const bool first_piece = (ii == 0 || run[ii - 1] != run[ii]);
                if (first_piece) nxt[th] = ii;
            }
        }
    });
    std::vector<uint64_t> total_runs(256, 0), total_chars(256, 0);
    for (unsigned u = 0; u < P; ++u)
        for (int c = 0; c < 256; ++c) { total_runs[c] += part[u].runs[c]; total_chars[c] += part[u].chars[c]; }
    run_parts(P, [&](unsigned t) {
        uint64_t seen[4] = {0, 0, 0, 0};
        uint64_t q[4] = {~0ULL, ~0ULL, ~0ULL, ~0ULL};
        uint64_t jc[256];                 // # c-runs that ended before the current run
        uint64_t cum[256];                // # c before the current run
        uint64_t last_piece[4] = {~0ULL, ~0ULL, ~0ULL, ~0ULL};
        memset(jc, 0, sizeof(jc));
        memset(cum, 0, sizeof(cum));
        for (unsigned u = 0; u < t; ++u) {
            for (int c = 0; c < 4; ++c) {
                seen[c] += part[u].len[c];
                if (part[u].last_piece[c] != ~0ULL) last_piece[c] = part[u].last_piece[c];
            }
            for (int c = 0; c < 256; ++c) { jc[c] += part[u].runs[c]; cum[c] += part[u].chars[c]; }
        }
        for (uint64_t i = part_lo(t), e = part_lo(t + 1); i < e; ++i) {
            MoveRow& r = img.rows[i];
            const uint32_t k = run[i];
            const uint8_t h = b.heads[k];
            const uint8_t th = tcode[h];
            const uint64_t len = img.row_pos[i + 1] - pos[i];
            const bool first_piece = (i == 0 || run[i - 1] != k);
            const bool last_piece_of_run = (i + 1 == nr || run[i + 1] != k);
            r.len = (uint16_t)len;
            r.head = h;
            r.flags = th == 0xFF ? kRowSlow : 0;
            put40(r.s_start, b.samples_start[k]);
            put40(r.s_last, b.samples_last[k]);
            if (first_piece) {
                // general-path tables: c-run j of symbol h starts here
                const uint64_t j = jc[h];
                img.crun_row[S.crun_base[h] + j] = (uint32_t)i;
                GapSlot& g = img.slots[S.slot_base[h] + j];
                g.thr = j == 0 ? 0 : (j - 1 < b.thr[h].size() ? b.thr[h][j - 1] : kThrInf);
                g.s_start = b.samples_start[k];
                g.cum = cum[h];
                GapSlot& g1 = img.slots[S.slot_base[h] + j + 1];
                g1.s_last_prev = b.samples_last[k];
            }
            if (th != 0xFF) {
                const uint64_t pi = S.F[h] + seen[th];
                seen[th] += len;
                uint64_t tt = q[th];
                if (tt == ~0ULL) tt = (uint64_t)(std::upper_bound(pos.begin(), pos.end(), pi) - pos.begin());
                while (tt < nr && pos[tt] <= pi) ++tt;
                q[th] = tt;                   // images of later h-rows start further right
                r.dest_row = (uint32_t)(tt - 1);
                r.dest_off = (uint16_t)(pi - pos[tt - 1]);
                for (int c = 0, s = 0; c < nts; ++c) {
                    if (c == th) continue;
                    const uint8_t sym = img.hdr.table_syms[c];
                    const uint64_t j = jc[sym];
                    const uint64_t T = j == 0 ? 0 : (j - 1 < b.thr[sym].size() ? b.thr[sym][j - 1] : kThrInf);
                    const uint64_t rel = T <= pos[i] ? 0 : std::min<uint64_t>(T - pos[i], len);
                    r.t_rel[s] = (uint16_t)rel;
                    const uint64_t d = last_piece[c] == ~0ULL ? kFar : i - last_piece[c];
                    r.dist_up[s] = (uint8_t)std::min<uint64_t>(d, kFar);
                    ++s;
                }
            }
            if (last_piece_of_run) {
                jc[h] += 1;
                cum[h] = cum[h] + (b.starts[k + 1] - b.starts[k]);
                if (th != 0xFF) last_piece[th] = i;
            }
        }
    });
    for (int c = 0; c < 256; ++c) {
        if (!b.count[c]) continue;
        if (total_runs[c] != b.nruns[c] || total_chars[c] != b.count[c]) throw std::runtime_error("exporter: per-symbol totals mismatch");
        GapSlot& g = img.slots[S.slot_base[c] + b.nruns[c]];
        g.thr = kThrInf;              // after the last c-run the reference can only jump up
        g.cum = b.count[c];
    }
    // sentinel row = "pos == n": never scanned past (len 1), always general path
    MoveRow& z = img.rows[nr];
    z.len = 1;
    z.flags = kRowSlow;
    if (opt.verbose)
        fprintf(stderr, "[moni-gpu] move table: n %llu, runs %llu, rows %llu (+%.1f%%), widest image %u rows, %.2f s\n",
                (unsigned long long)n, (unsigned long long)R, (unsigned long long)nr, 100.0 * (double)(nr - R) / (double)R,
                worst, now_s() - t0);
}

void build_phi_tables(const FlatRlbwt& b, HostImage& img) {
    img.phi.clear();
    img.phi_inv.clear();
    img.phi_dir.clear();
    img.phi_inv_dir.clear();
    img.hdr.phi_n = 0;
    img.hdr.first_run_sample = b.R ? (b.samples_start[0] + 1) % b.n : 0;
    if (b.pred.empty() || b.pred_start.empty()) return;
    const uint64_t R = b.R;
    if (b.pred.size() != R || b.pred_to_run.size() != R || b.pred_start.size() != R || b.pred_start_to_run.size() != R)
        throw std::runtime_error("exporter: Phi structures do not have R entries");
    img.phi.resize(R);
    img.phi_inv.resize(R);
    for (uint64_t j = 0; j < R; ++j) {
        if (j && (b.pred[j] <= b.pred[j - 1] || b.pred_start[j] <= b.pred_start[j - 1]))
            throw std::runtime_error("exporter: Phi keys not increasing");
        const uint64_t run = b.pred_to_run[j], run_s = b.pred_start_to_run[j];
        if (run >= R || run_s >= R) throw std::runtime_error("exporter: Phi run index out of range");
        img.phi[j].key = b.pred[j];
        img.phi[j].val = run > 0 ? b.samples_last[run - 1] : kPhiNone;
        img.phi_inv[j].key = b.pred_start[j];
        img.phi_inv[j].val = run_s + 1 < R ? b.samples_start[run_s + 1] : kPhiNone;
    }
    img.hdr.phi_n = R;
    // bucket directories: about four keys per bucket, so a predecessor search is one directory word pair plus one or two
    // lines of entries instead of log2(R) dependent probes
    uint32_t shift = 0;
    while (shift < 40 && ((b.n >> shift) > R / 4 + 1)) ++shift;
    img.hdr.phi_shift = shift;
    const uint64_t nb = (b.n >> shift) + 2;
    auto fill = [&](const std::vector<PhiEntry>& tab, std::vector<uint32_t>& dir) {
        dir.assign(nb, 0);
        uint64_t j = 0;
        for (uint64_t bk = 0; bk < nb; ++bk) {
            const uint64_t lim = bk << shift;                 // first key of bucket bk
            while (j < R && tab[j].key < lim) ++j;
            dir[bk] = (uint32_t)j;
        }
    };
    fill(img.phi, img.phi_dir);
    fill(img.phi_inv, img.phi_inv_dir);
}

void build_grammar_table(const FlatSlp& g, const ExportOptions& opt, HostImage& img) {
    const double t0 = now_s();
    const uint64_t A = g.alph_size, NR = g.num_rules;
    const uint64_t L = std::max<uint64_t>(opt.leaf_len, 1);
    if (g.text_len == 0) throw std::runtime_error("exporter: empty grammar");
    img.hdr.text_len = g.text_len;
    img.hdr.leaf_len = (uint32_t)std::min<uint64_t>(L, 0xFFFFFFFFull);
    img.hdr.pool_shift = 3;
    const uint64_t align = 1ULL << img.hdr.pool_shift;
    auto explen = [&](uint32_t v) -> uint64_t { return v < A ? 1 : g.explen[v - A]; };
    if (NR == 0 && (g.text_len != 1 || g.start_var >= A)) throw std::runtime_error("exporter: grammar without rules");

    std::vector<uint32_t> ref(NR + A, 0xFFFFFFFFu);   // child reference of every variable once assigned
    std::vector<uint32_t> order;                        // internal variables in breadth-first order
    std::vector<uint8_t>& pool = img.pool;
    pool.clear();
    std::vector<uint32_t> stack;
    constexpr uint64_t kCopyMin = 24;                   // shorter expansions are cheaper to walk than to look up
    std::vector<uint64_t> first_at(NR, ~0ULL);          // pool offset of the first complete expansion of a rule
    auto leaf_ref = [&](uint32_t v) -> uint32_t {
        if (ref[v] != 0xFFFFFFFFu) return ref[v];
        while (pool.size() % align) pool.push_back(0);
        const uint64_t off = pool.size() >> img.hdr.pool_shift;
        if (off >= kLeafBit) throw std::runtime_error("exporter: literal pool exceeds the addressable range; lower leaf_len");
        // depth-first expansion; a variable met again is copied from where its expansion was first written (the grammar of
        // a pangenome reuses long variables: this turns ~|T| stack operations into block copies)
        stack.clear();
        stack.push_back(v);
        while (!stack.empty()) {
            const uint32_t x = stack.back();
            stack.pop_back();
            if (x < A) { pool.push_back(g.alph[x]); continue; }
            const uint64_t len = g.explen[x - A];
            if (len >= kCopyMin) {
                uint64_t& fp = first_at[x - A];
                if (fp != ~0ULL) {
                    const size_t old = pool.size();
                    pool.resize(old + len);
                    memcpy(pool.data() + old, pool.data() + fp, len);     // [fp, fp + len) was complete before `old`
                    continue;
                }
                fp = pool.size();
            }
            stack.push_back(g.right[x - A]);
            stack.push_back(g.left[x - A]);
        }
        ref[v] = kLeafBit | (uint32_t)off;
        return ref[v];
    };
    auto child_ref = [&](uint32_t v) -> uint32_t {
        if (explen(v) <= L) return leaf_ref(v);
        if (ref[v] == 0xFFFFFFFFu) {
            if (order.size() >= kLeafBit - 1) throw std::runtime_error("exporter: too many grammar nodes");
            ref[v] = (uint32_t)order.size();
            order.push_back(v);
        }
        return ref[v];
    };
    img.hdr.root_ref = child_ref(g.start_var);
    img.nodes.clear();
    for (size_t i = 0; i < order.size(); ++i) {     // order grows while we walk it: breadth-first
        const uint32_t v = order[i];
        const uint32_t lv = g.left[v - A], rv = g.right[v - A];
        if (explen(lv) + explen(rv) != explen(v)) throw std::runtime_error("exporter: inconsistent expansion lengths");
        SlpNode nd;
        nd.left = child_ref(lv);
        nd.right = child_ref(rv);
        nd.left_len = explen(lv);
        img.nodes.push_back(nd);
    }
    for (int i = 0; i < 32; ++i) pool.push_back(0);   // slack so 16-byte block loads never run off the end
    img.hdr.n_nodes = img.nodes.size();
    img.hdr.pool_bytes = pool.size();
    // depth = nodes on the longest root-to-literal path.  A shared variable can be referenced from a
    // later node, so relax until stable (a couple of sweeps in practice).
    uint32_t depth_max = 0;
    {
        std::vector<uint16_t> depth(img.nodes.size(), 1);
        for (bool changed = true; changed;) {
            changed = false;
            for (size_t i = 0; i < img.nodes.size(); ++i) {
                const uint16_t d = (uint16_t)(depth[i] + 1);
                const SlpNode& nd = img.nodes[i];
                if (!(nd.left & kLeafBit) && depth[nd.left] < d) { depth[nd.left] = d; changed = changed || nd.left < i; }
                if (!(nd.right & kLeafBit) && depth[nd.right] < d) { depth[nd.right] = d; changed = changed || nd.right < i; }
            }
        }
        for (uint16_t d : depth) depth_max = std::max<uint32_t>(depth_max, d);
    }
    img.hdr.slp_depth_max = depth_max;
    if (opt.verbose)
        fprintf(stderr, "[moni-gpu] grammar table: text %llu, rules %llu -> %llu nodes (depth <= %u) + %.1f MB literals (leaf_len %u), %.2f s\n",
                (unsigned long long)g.text_len, (unsigned long long)NR, (unsigned long long)img.nodes.size(), depth_max,
                (double)pool.size() / 1e6, (unsigned)L, now_s() - t0);
}

}  // namespace moni
