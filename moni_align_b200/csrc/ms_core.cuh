// Per-read arithmetic of the two passes over the HBM image (device_image.hpp), written once as
// host+device inline functions: the CUDA kernels (kernels.cu) call them per thread, and the CPU
// test harness (tests/host_sim.cpp) compiles the very same functions with g++ so the logic can be
// checked against the oracle on a machine without a GPU.  The host build is TEST-ONLY: the shipped
// library has no CPU execution path.
//
// Reference behaviour being reproduced bit for bit:
//   pass 1  ms_pointers<>::_query                      include/ms/ms_pointers.hpp:455-511
//   pass 2  ms_c::matching_statistics length loop      src/matching_statistics.cpp:246-255
//           mems_c::maxrimal_exact_matches MEM rule    src/mems.cpp:247-259
//           seed_finder::find_mems seed rule           include/aligner/seed_finder.hpp:126-166
#pragma once
#include <cstdint>

#include "device_image.hpp"

#if defined(__CUDACC__)
#define MONI_HD __host__ __device__ __forceinline__
#else
#define MONI_HD inline
#endif

// -DMONI_BOUNDS_CHECK builds the kernels with an index check in front of every global access whose index is computed from
// input data (rows, read bytes, text words, flag words, restart lists, records): a violation prints its source line and traps,
// which the host sees as a CUDA error.  The pool's compute-sanitizer is closed, so this build (variants/libmoni_gpu_check.so,
// tests/test_gpu_parity.py::test_bounds_checked_build_runs_the_suite) is the memcheck of this tree.
#if defined(MONI_BOUNDS_CHECK) && defined(__CUDA_ARCH__)
#define MONI_ASSERT(cond)                                                          \
    do {                                                                           \
        if (!(cond)) {                                                             \
            printf("moni bounds check failed: %s (%s:%d)\n", #cond, __FILE__, __LINE__); \
            __trap();                                                              \
        }                                                                          \
    } while (0)
#else
#define MONI_ASSERT(cond) ((void)0)
#endif

namespace moni {

struct ImageView {
    const MoveRow* rows = nullptr;
    const uint64_t* row_pos = nullptr;
    const uint32_t* crun_row = nullptr;
    const GapSlot* slots = nullptr;
    const SlpNode* nodes = nullptr;
    const uint8_t* pool = nullptr;
    const SymbolTables* sym = nullptr;
    const PhiEntry* phi = nullptr;       // `moni align` indexes only
    const PhiEntry* phi_inv = nullptr;
    const uint32_t* phi_dir = nullptr;
    const uint32_t* phi_inv_dir = nullptr;
    ImageHeader hdr;
};

// Event counters of SURVEY.md 8(d) (B, J, S, D, C) plus the image-specific ones.
struct Counters {
    unsigned long long bases, jumps, restarts, restart_depth, chars, scan_rows, general_steps, descents;
};

// ---- a MoveRow held in registers ---------------------------------------------------------------
struct RowRegs {
    uint32_t w[8];
    MONI_HD uint32_t dest_row() const { return w[0]; }
    MONI_HD uint32_t dest_off() const { return w[1] & 0xFFFFu; }
    MONI_HD uint32_t len() const { return w[1] >> 16; }
    MONI_HD uint32_t head() const { return (w[3] >> 16) & 0xFFu; }
    MONI_HD uint32_t flags() const { return w[3] >> 24; }
    MONI_HD uint32_t t_rel(int s) const {
        const uint64_t t = (uint64_t)w[2] | ((uint64_t)(w[3] & 0xFFFFu) << 32);
        return (uint32_t)(t >> (16 * s)) & 0xFFFFu;
    }
    MONI_HD uint64_t s_start() const { return (uint64_t)w[4] | ((uint64_t)(w[5] & 0xFFu) << 32); }
    MONI_HD uint64_t s_last() const { return (uint64_t)(w[5] >> 8) | ((uint64_t)(w[6] & 0xFFFFu) << 24); }
    MONI_HD uint32_t dist_up(int s) const {
        const uint32_t d = (w[6] >> 16) | ((w[7] & 0xFFu) << 16);
        return (d >> (8 * s)) & 0xFFu;
    }
    MONI_HD uint32_t dist_dn(int s) const { return ((w[7] >> 8) >> (8 * s)) & 0xFFu; }
};

// One 32-byte row = one DRAM sector = one load instruction (sm_100 LDG.E.256).
MONI_HD RowRegs load_row(const MoveRow* rows, uint64_t i) {
    RowRegs r;
#if defined(__CUDA_ARCH__)
    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]), "=r"(r.w[4]), "=r"(r.w[5]), "=r"(r.w[6]), "=r"(r.w[7])
                 : "l"(rows + i));
#else
    const uint32_t* p = reinterpret_cast<const uint32_t*>(rows + i);
    for (int k = 0; k < 8; ++k) r.w[k] = p[k];
#endif
    return r;
}

template <typename T>
MONI_HD T ld(const T* p) {
#if defined(__CUDA_ARCH__)
    return __ldg(p);
#else
    return *p;
#endif
}

// ---- exact general path (any symbol, any position incl. pos == n) ------------------------------
// largest i in [0, n_rows] with row_pos[i] <= pos   (pos <= n; pos == n gives the sentinel row)
MONI_HD uint64_t locate_row(const ImageView& v, uint64_t pos) {
    uint64_t lo = 0, hi = v.hdr.n_rows;       // invariant: row_pos[lo] <= pos, answer in [lo, hi]
    while (lo < hi) {
        const uint64_t mid = lo + (hi - lo + 1) / 2;
        if (ld(v.row_pos + mid) <= pos) lo = mid; else hi = mid - 1;
    }
    return lo;
}

// number of entries of a[0..cnt) that are < key
MONI_HD uint64_t count_less(const uint32_t* a, uint64_t cnt, uint64_t key) {
    uint64_t lo = 0, hi = cnt;
    while (lo < hi) {
        const uint64_t mid = lo + (hi - lo) / 2;
        if ((uint64_t)ld(a + mid) < key) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// One iteration of ms_pointers.hpp:466-507 on an absolute BWT position.
MONI_HD void general_step(const ImageView& v, uint8_t c, uint64_t& pos, uint64_t& sample, bool& jumped) {
    const SymbolTables& S = *v.sym;
    jumped = false;
    const uint64_t n_c = ld(&S.count[c]);
    const uint64_t Fc = ld(&S.F[c]);
    if (n_c == 0) {                                  // :470-475  rank() of an absent letter is 0
        sample = 0;
        pos = Fc;
        return;
    }
    const uint64_t i = locate_row(v, pos);
    const uint32_t* cr = v.crun_row + ld(&S.crun_base[c]);
    const GapSlot* sl = v.slots + ld(&S.slot_base[c]);
    const uint64_t nruns = ld(&S.nruns[c]);
    if (pos < v.hdr.n && v.rows[i].head == c) {      // :476-481
        const uint64_t j = count_less(cr, nruns, i + 1) - 1;      // the c-run whose first piece is at or before row i
        const uint64_t first = ld(cr + j);
        const uint64_t rank = ld(&sl[j].cum) + (pos - ld(v.row_pos + first));
        sample = sample - 1;
        pos = Fc + rank;
        return;
    }
    jumped = true;                                   // :482-505
    const uint64_t jc = pos >= v.hdr.n ? nruns : count_less(cr, nruns, i);
    const uint64_t thr = ld(&sl[jc].thr);
    const uint64_t cum = ld(&sl[jc].cum);
    if (pos < thr) {                                 // rnk_c.first > thresholds.rank(pos + 1, c)
        sample = ld(&sl[jc].s_last_prev);
        pos = Fc + cum - 1;
    } else {
        sample = ld(&sl[jc].s_start);
        pos = Fc + cum;
    }
}

// ---- pass 1 -------------------------------------------------------------------------------------
struct ScanState {
    uint32_t cur;       // row
    uint32_t off;       // offset inside the row (may exceed the row length until scanned forward)
    uint64_t sample;
};

MONI_HD void scan_init(const ImageView& v, ScanState& s) {
    s.cur = v.hdr.init_row;
    s.off = v.hdr.init_off;
    s.sample = v.hdr.init_sample;
}

// Consumes one symbol; on return s.sample is the pointer to store.  `tcode` is the 256-byte symbol
// class table (shared memory on the device).
MONI_HD void scan_step(const ImageView& v, const uint8_t* tcode, uint8_t c, ScanState& s, Counters* cnt) {
    MONI_ASSERT(s.cur <= v.hdr.n_rows);
    RowRegs row = load_row(v.rows, s.cur);
    while (s.off >= row.len()) {                     // forward scan after the previous LF (bounded by balancing)
        s.off -= row.len();
        s.cur += 1;
        MONI_ASSERT(s.cur <= v.hdr.n_rows);
        row = load_row(v.rows, s.cur);
        if (cnt) cnt->scan_rows++;
    }
    const uint32_t tc = tcode[c];
    bool general = (tc == 0xFFu) || (row.flags() & kRowSlow);
    if (!general) {
        if (row.head() == c) {                       // match: rle_string::rank via the row's LF image
            s.sample -= 1;
            s.off += row.dest_off();
            s.cur = row.dest_row();
            return;
        }
        const uint32_t th = tcode[row.head()];
        const int slot = (int)(tc < th ? tc : tc - 1);
        const bool up = s.off < row.t_rel(slot);
        const uint32_t d = up ? row.dist_up(slot) : row.dist_dn(slot);
        if (d != kFar) {
            MONI_ASSERT(up ? s.cur >= d : (uint64_t)s.cur + d <= v.hdr.n_rows);
            const RowRegs q = load_row(v.rows, up ? (uint64_t)s.cur - d : (uint64_t)s.cur + d);
            s.sample = up ? q.s_last() : q.s_start();
            s.off = q.dest_off() + (up ? q.len() - 1 : 0);
            s.cur = q.dest_row();
            if (cnt) cnt->jumps++;
            return;
        }
        general = true;
    }
    uint64_t pos = ld(v.row_pos + s.cur) + s.off;
    bool jumped;
    general_step(v, c, pos, s.sample, jumped);
    const uint64_t i = locate_row(v, pos);
    s.cur = (uint32_t)i;
    s.off = (uint32_t)(pos - ld(v.row_pos + i));
    if (cnt) { cnt->general_steps++; cnt->jumps += jumped ? 1 : 0; }
}

// ---- pass 2 -------------------------------------------------------------------------------------
// Eight bytes from an arbitrary address (little endian).  May touch up to 15 bytes past p: every
// buffer read this way (device read batches, the literal pool) carries that much slack.
MONI_HD uint64_t load8(const uint8_t* p) {
#if defined(__CUDA_ARCH__)
    const unsigned long long a = (unsigned long long)p;
    const uint32_t sh = (uint32_t)(a & 7ull) * 8u;
    const uint64_t* q = reinterpret_cast<const uint64_t*>(a & ~7ull);
    const uint64_t lo = __ldg(q), hi = __ldg(q + 1);
    return sh ? (lo >> sh) | (hi << (64u - sh)) : lo;
#else
    uint64_t x = 0;
    for (int k = 7; k >= 0; --k) x = (x << 8) | p[k];
    return x;
#endif
}

MONI_HD uint32_t first_diff_byte(uint64_t x) {      // x != 0
#if defined(__CUDA_ARCH__)
    return (uint32_t)(__ffsll((long long)x) - 1) >> 3;
#else
    return (uint32_t)__builtin_ctzll(x) >> 3;
#endif
}

// ---- restart runs: what pass 1 hands to pass 2 ---------------------------------------------------
// The reference's length loop (src/matching_statistics.cpp:246-255) only does work at query offsets i where
// `i == 0 || pointers[i] != pointers[i-1] + 1` (u64 arithmetic): call them restarts.  Everywhere else the pointer is the
// previous one plus one and the length the previous one minus one (floor 0).  So pass 1 emits, per read,
//   flags : one bit per query offset, set at the restarts (word w of the read holds offsets 32 w .. 32 w + 31)
//   rptr  : pointers[i] of the restarts, in the order the backward scan meets them (descending i; the last entry is i = 0)
// and pass 2 walks the restarts only.  The full arrays are rebuilt from (flags, rptr, lengths at the restarts) where an
// interface wants them (expand_runs below; k_ms_expand on the device).
//
// Flag words of read r (r-th of its chunk, chunk-relative offset o) start at word (o >> 5) + r: disjoint for distinct reads
// (floor((o + m) / 32) - floor(o / 32) + 1 >= ceil(m / 32)), at most bases / 32 + n_reads + 1 words per chunk.
MONI_HD uint64_t flag_base(uint64_t o, uint64_t r) { return (o >> 5) + r; }

struct RunSink {
    uint32_t acc = 0;       // flag bits of the word being filled
    uint32_t k = 0;         // restarts written so far
};

// To be called after the backward scan consumed base j (from m - 1 down to 0): `old` is the sample before the step
// (= pointers[j + 1]), `now` the sample after it (= pointers[j]).  rptr / flags point at the read's own areas.
MONI_HD void run_note(RunSink& rs, uint32_t j, uint32_t m, uint64_t old, uint64_t now, uint64_t* rptr, uint32_t* flags) {
    const uint32_t i = j + 1;
    if (i < m) {
        if (old != now + 1) { rptr[rs.k++] = old; rs.acc |= 1u << (i & 31u); }
        if ((i & 31u) == 0) { flags[i >> 5] = rs.acc; rs.acc = 0; }          // offsets i .. i + 31 are decided
    }
    if (j == 0) { rptr[rs.k++] = now; flags[0] = rs.acc | 1u; }
}

MONI_HD uint32_t ctz32(uint32_t x) {     // x != 0
#if defined(__CUDA_ARCH__)
    return (uint32_t)(__ffs((int)x) - 1);
#else
    return (uint32_t)__builtin_ctz(x);
#endif
}

// Pass 2 as a per-read state machine over the restarts.  The reference's loop is
//     for i: pos = ptr[i]; while (i+l < m && pos+l < N && restart(i) && read[i+l] == charAt(pos+l)) ++l; len[i] = l; l -= (l != 0)
// with charAt = PlainSlp::charAt (thirdparty/ShapedSlp/PlainSlp.hpp:101-126), one root-to-leaf descent per character.  Here
// the offsets between two restarts are skipped arithmetically (l entering restart i = max(L' - (i - i'), 0) for the previous
// restart (i', L')), one descent reaches a literal block that is compared eight bytes at a time, and the work is cut into
// micro-steps -- NEXT restart, DESCEND to a literal block, COMPARE one word -- so that the 32 reads of a warp each take
// whatever micro-step they need in the same loop trip.  With the text in HBM a restart is normally one trip: NEXT falls through
// DESCEND (nothing to do) into COMPARE, whose first word already holds the mismatch.
struct RunState {
    enum : uint32_t { NEXT = 0, DESCEND = 1, COMPARE = 2, FINISHED = 3 };
    uint32_t mode = NEXT;
    uint32_t k = 0;                 // restarts finished
    uint32_t w = 0, bits = 0;       // flag cursor: bits = flags of word w not yet visited
    uint32_t ip = 0, Lp = 0;        // previous restart and its length
    uint32_t i = 0, l = 0, n_mems = 0;
    uint32_t n_Ns = 0;              // seeds only: length of the run of 'N' that ends the characters compared so far
    uint32_t ref = 0;               // DESCEND: current grammar reference
    uint64_t pos = 0;
    uint64_t nxt = 0, nxt2 = 0;     // pointers of the next two restarts, loaded ahead
    uint64_t span = 0, p = 0;       // DESCEND: expansion length of ref, offset inside it
    const uint8_t* t = nullptr;     // COMPARE: next text byte
    uint64_t rem = 0;               // COMPARE: text bytes left in the literal block
};

// The chunk buffers pass 2 works on (kernel-constant) and where one read lives in them (32-bit: a chunk holds fewer than 2^32
// bases and 2^28 reads).  Keeping offsets instead of six pointers per lane is worth eight registers in the kernel.
struct RunBufs {
    const uint8_t* seqs = nullptr;      // read bytes of the chunk
    const uint32_t* flags = nullptr;    // restart flags (flag_base)
    const uint64_t* rptr = nullptr;     // pointers at the restarts (run_base)
    uint64_t* rrec = nullptr;           // (i << 32) | len of every restart, same indexing as rptr; or null
    uint64_t* rec = nullptr;            // MEM / seed records at the read's own offset; or null
    uint64_t* rec_ref = nullptr;        // seeds: pointers[i] of every record; or null
    uint64_t seq_bytes = 0, flag_words = 0, run_entries = 0, rec_entries = 0;   // extents of the above (MONI_BOUNDS_CHECK builds test against them)
};
struct RunRead {
    uint32_t o = 0;         // chunk-relative offset of the read's first base (also where its records go)
    uint32_t m = 0;         // length, >= 1
    uint32_t fb = 0;        // first flag word
    uint32_t last = 0;      // index in rptr / rrec of the restart at query offset 0 (= first entry + cnt - 1)
    uint32_t cnt = 0;       // restarts
};

// One micro-step of a read.  The k-th restart in increasing query offset is rptr[last - k].  When a restart is finished, writes
// rrec[last - k] = (i << 32) | len (if rrec) and the MEM record(s) (packed (i << 32) | len, if rec): the reference's rule
// `i == 0 || len[i] >= len[i-1]` (src/mems.cpp:257-258) also fires at offsets between restarts whose predecessor has length 0
// (then len stays 0): those (i, 0) records are written on the way to the next restart.
// SEEDS: the record rule is seed_finder::find_mems' (include/aligner/seed_finder.hpp:148): l >= pl and n_Ns < l and
// l >= min_len, where n_Ns counts the consecutive 'N' at the end of the characters matched so far (:139-143; it is only
// touched when a character is compared and matches), which can only hold at a restart; rec_ref[k] receives pointers[i].
// With the text in HBM the root is itself a literal block: DESCEND falls straight through.  Two variations were measured on cfg3 and
// dropped (profiles/r02_c_launches.csv, r02_g_ab_variants.json, r02_h_ab_variants.json): requesting the text line of the NEXT
// restart with prefetch.global.L2 while this one is compared (DRAM reads doubled, 17.3 -> 38.6 GB per 2 M reads: the demand load
// fetches the line again; 5.0 -> 8.7 ms), and a text-only specialisation without the descent state (5.1 ms against 4.6 ms for this
// code at the same 64-register cap; at 40 registers and 1536 threads per SM 6.2 ms: the read bytes and restart lists of that many
// reads no longer fit L1).
// Call while st.mode != FINISHED.
template <bool SEEDS = false>
MONI_HD void runs_step(const ImageView& v, const RunBufs& B, const RunRead& R, RunState& st, Counters* ctr, uint32_t min_len = 0) {
    const uint64_t N = v.hdr.text_len;
    const uint32_t m = R.m;
    bool done = false;
    if (st.mode == RunState::NEXT) {
        MONI_ASSERT(R.cnt >= 1 && R.last >= R.cnt - 1 && R.last < B.run_entries && (uint64_t)R.fb + (m + 31) / 32 <= B.flag_words);
        MONI_ASSERT(!B.rec || (uint64_t)R.o + m <= B.rec_entries);
        if (st.k == R.cnt) {                                      // offsets after the last restart: lengths run down to 0, then (i, 0) records
            if (!SEEDS && B.rec) {
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
                for (uint32_t q = st.ip + st.Lp + 1; q < m; ++q) B.rec[(uint64_t)R.o + st.n_mems++] = (uint64_t)q << 32;
            }
            st.mode = RunState::FINISHED;
            return;
        }
        if (st.k == 0) {
            st.bits = ld(B.flags + R.fb);
            st.nxt = ld(B.rptr + R.last);
            if (R.cnt > 1) st.nxt2 = ld(B.rptr + (R.last - 1));
        }
        while (st.bits == 0) { MONI_ASSERT(st.w + 1 < (m + 31) / 32); st.bits = ld(B.flags + R.fb + (++st.w)); }
        const uint32_t i = 32u * st.w + ctz32(st.bits);
        MONI_ASSERT(i < m);
        st.bits &= st.bits - 1u;
        st.pos = st.nxt;
        st.nxt = st.nxt2;                                         // pointers are consumed two restarts after they were requested
        if (st.k + 2 < R.cnt) st.nxt2 = ld(B.rptr + (R.last - (st.k + 2)));
        if (st.k) {
            if (!SEEDS && B.rec) {
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
                for (uint32_t q = st.ip + st.Lp + 1; q < i; ++q) B.rec[(uint64_t)R.o + st.n_mems++] = (uint64_t)q << 32;
            }
            const uint32_t d = i - st.ip;                         // >= 1
            st.l = st.Lp > d ? st.Lp - d : 0u;                    // the l the reference carries into offset i
        } else {
            st.l = 0;
        }
        st.i = i;
        if (ctr) ctr->restarts++;
        done = true;
        if (i + st.l < m && st.pos + st.l < N) {                  // u64 add wraps exactly as in the reference
            done = false;
            st.mode = RunState::DESCEND;
            st.ref = v.hdr.root_ref;
            st.span = N;
            st.p = st.pos + st.l;
        }
    }
    if (st.mode == RunState::DESCEND) {
        // tight loop on purpose: a level is ~10 instructions, far less than one trip through this function
        while (!(st.ref & kLeafBit)) {
#if defined(__CUDA_ARCH__)
            const uint4 q = __ldg(reinterpret_cast<const uint4*>(v.nodes + st.ref));
            const uint32_t lc = q.x, rc = q.y;
            const uint64_t ll = (uint64_t)q.z | ((uint64_t)q.w << 32);
#else
            const uint32_t lc = v.nodes[st.ref].left, rc = v.nodes[st.ref].right;
            const uint64_t ll = v.nodes[st.ref].left_len;
#endif
            if (st.p < ll) { st.ref = lc; st.span = ll; } else { st.p -= ll; st.span -= ll; st.ref = rc; }
            if (ctr) ctr->restart_depth++;
        }
        st.t = v.pool + ((uint64_t)(st.ref & ~kLeafBit) << v.hdr.pool_shift) + st.p;
        st.rem = st.span - st.p;                                 // > 0, and the block lies inside the text: pos + l < N holds in it
        st.mode = RunState::COMPARE;
        if (ctr) ctr->descents++;
    }
    if (st.mode == RunState::COMPARE) {
        const uint32_t e = st.i + st.l;                          // < m
        uint32_t lim = m - e;
        if (lim > 8u) lim = 8u;
        if ((uint64_t)lim > st.rem) lim = (uint32_t)st.rem;
        MONI_ASSERT((uint64_t)R.o + e + 16 <= B.seq_bytes);
        MONI_ASSERT((uint64_t)(st.t - v.pool) + 16 <= v.hdr.pool_bytes);
        const uint64_t rd = load8(B.seqs + ((uint64_t)R.o + e));
        const uint64_t x = load8(st.t) ^ rd;
        uint32_t k = x ? first_diff_byte(x) : 8u;
        const bool mismatch = k < lim;
        if (!mismatch) k = lim;
        if (SEEDS && k) {
            // bit 7 of every byte of the read word that is NOT 'N', restricted to the k matched bytes
            const uint64_t z = rd ^ 0x4E4E4E4E4E4E4E4Eull;
            uint64_t nz = ((((z & 0x7F7F7F7F7F7F7F7Full) + 0x7F7F7F7F7F7F7F7Full) | z) & 0x8080808080808080ull);
            if (k < 8u) nz &= (1ull << (8u * k)) - 1ull;
            if (nz == 0) st.n_Ns += k;                               // all k are 'N': the run grows
            else {
#if defined(__CUDA_ARCH__)
                const uint32_t h = (uint32_t)(63 - __clzll((long long)nz)) >> 3;
#else
                const uint32_t h = (uint32_t)(63 - __builtin_clzll(nz)) >> 3;
#endif
                st.n_Ns = k - 1u - h;                                // 'N's after the last other character
            }
        }
        if (ctr) ctr->chars += k + (mismatch ? 1u : 0u);
        st.l += k;
        st.t += k;
        st.rem -= k;
        if (mismatch || e + k >= m) {
            done = true;
        } else if (st.rem == 0) {                                // block exhausted while still matching
            if (st.pos + st.l < N) {
                st.mode = RunState::DESCEND;
                st.ref = v.hdr.root_ref;
                st.span = N;
                st.p = st.pos + st.l;
            } else {
                done = true;
            }
        }
    }
    if (done) {
        const uint32_t i = st.i;
        if (B.rrec) B.rrec[R.last - st.k] = ((uint64_t)i << 32) | st.l;
        uint32_t prev_len = 0;                                   // len[i - 1]
        if (st.k) {
            const uint32_t d1 = i - st.ip - 1u;
            prev_len = st.Lp > d1 ? st.Lp - d1 : 0u;
        }
        const bool emit = SEEDS ? (st.l >= prev_len && st.n_Ns < st.l && st.l >= min_len)        // seed_finder.hpp:148 (pl starts at 0)
                                : (i == 0 || st.l >= prev_len);                                   // src/mems.cpp:257-258
        if (emit) {
            MONI_ASSERT(st.n_mems < m);
            if (B.rec) B.rec[(uint64_t)R.o + st.n_mems] = ((uint64_t)i << 32) | st.l;
            if (SEEDS && B.rec_ref) B.rec_ref[(uint64_t)R.o + st.n_mems] = st.pos;
            ++st.n_mems;
        }
        st.ip = i;
        st.Lp = st.l;
        ++st.k;
        st.mode = RunState::NEXT;
    }
}

// The whole length loop of one read (a single lane of the state machine).  Returns the number of MEM records.
template <bool SEEDS = false>
MONI_HD uint32_t runs_read(const ImageView& v, const RunBufs& B, const RunRead& R, Counters* ctr, uint32_t min_len = 0) {
    RunState st;
    while (st.mode != RunState::FINISHED) runs_step<SEEDS>(v, B, R, st, ctr, min_len);
    return st.n_mems;
}

// pointers[] / lengths[] of one read from its restarts (what the reference's vectors hold): between restarts the pointer
// grows by one and the length shrinks by one (floor 0), in u64 arithmetic.  Host-side helper of the run interfaces.
inline void expand_runs(uint64_t m, const uint32_t* flags, const uint64_t* rptr, const uint64_t* rrec, uint32_t cnt, uint64_t* ptr_out,
                        uint64_t* len_out) {
    uint64_t p0 = 0, l0 = 0, i0 = 0;
    uint32_t k = 0;
    for (uint64_t i = 0; i < m; ++i) {
        if ((flags[i >> 5] >> (i & 31)) & 1u) {
            p0 = rptr[cnt - 1 - k];
            l0 = rrec[cnt - 1 - k] & 0xFFFFFFFFull;
            i0 = i;
            ++k;
        }
        ptr_out[i] = p0 + (i - i0);
        len_out[i] = l0 > i - i0 ? l0 - (i - i0) : 0;
    }
}

// ---- occurrence enumeration (`moni align`, include/aligner/seed_finder.hpp:169-251, 346-368) ---------------------
// r_index::Phi (r_index.hpp:225-243) / ms_pointers::Phi_inv (moni.hpp:338-354) on a flattened table: jr = strict
// predecessor of i among the keys, circular (sparse_sd_vector.hpp:168-172); delta = key < i ? i - key : i + 1;
// result = (val + delta) % n.
// The directory narrows the search to the bucket of i (keys in [b << shift, (b + 1) << shift), about four of them).
MONI_HD uint64_t phi_step(const PhiEntry* tab, const uint32_t* dir, uint32_t shift, uint64_t cnt, uint64_t n, uint64_t i) {
    const uint64_t bk = i >> shift;                  // i < n, the directory has (n >> shift) + 2 words
    uint64_t lo = ld(dir + bk), hi = ld(dir + bk + 1);   // # keys < i lies in [lo, hi]
    while (lo < hi) {
        const uint64_t mid = lo + (hi - lo) / 2;
        if (ld(&tab[mid].key) < i) lo = mid + 1; else hi = mid;
    }
    const uint64_t jr = lo == 0 ? cnt - 1 : lo - 1;
    const uint64_t j = ld(&tab[jr].key);
    const uint64_t delta = j < i ? i - j : i + 1;
    return (ld(&tab[jr].val) + delta) % n;
}

// Is the longest common extension of text positions a and b at least len?  What the caller of lceToRBounded
// (thirdparty/ShapedSlp/Common.hpp:250-295) tests at seed_finder.hpp:178,208; requires the text in HBM (text mode) and
// a + len <= N, b + len <= N.
MONI_HD bool lce_at_least(const ImageView& v, uint64_t a, uint64_t b, uint64_t len) {
    const uint8_t* t = v.pool + ((uint64_t)(v.hdr.root_ref & ~kLeafBit) << v.hdr.pool_shift);
#if defined(__CUDA_ARCH__)
    // both sides as streams of aligned 8-byte words: one new load per side and step (load8 would issue two)
    const unsigned long long pa = (unsigned long long)(t + a), pb = (unsigned long long)(t + b);
    const uint32_t sa = (uint32_t)(pa & 7ull) * 8u, sb = (uint32_t)(pb & 7ull) * 8u;
    const uint64_t* qa = reinterpret_cast<const uint64_t*>(pa & ~7ull);
    const uint64_t* qb = reinterpret_cast<const uint64_t*>(pb & ~7ull);
    uint64_t wa = __ldg(qa), wb = __ldg(qb);
    for (uint64_t k = 0; k < len; k += 8) {
        const uint64_t na = __ldg(++qa), nb = __ldg(++qb);
        const uint64_t xa = sa ? (wa >> sa) | (na << (64u - sa)) : wa;
        const uint64_t xb = sb ? (wb >> sb) | (nb << (64u - sb)) : wb;
        uint64_t x = xa ^ xb;
        if (len - k < 8) x &= (1ull << (8 * (len - k))) - 1ull;
        if (x) return false;
        wa = na;
        wb = nb;
    }
    return true;
#else
    uint64_t k = 0;
    for (; k + 8 <= len; k += 8)
        if (load8(t + a + k) != load8(t + b + k)) return false;
    if (k < len) {
        const uint64_t x = load8(t + a + k) ^ load8(t + b + k);
        const uint64_t mask = (1ull << (8 * (len - k))) - 1ull;
        if (x & mask) return false;
    }
    return true;
#endif
}

// seed_finder::find_MEM_occs without seed filtering (:239-251): the seed position, then find_MEM_above (:169-197: Phi steps
// while the LCE with the previous suffix reaches len; get_prev_occ_with_lcp :359-370), then find_MEM_below (:201-229, Phi^-1,
// get_next_occ_with_lcp :346-357).  Generalised to what populate_seed (:258-309) asks of the left half of a split MEM: the
// list starts with `head`, the walk upwards starts at `up` and the walk downwards at `down` (a plain seed has head = up = down
// = its position).  Writes at most `cap` positions per direction to out (if not null; out[0] = head) and returns the number of
// positions (1 + above + below); *truncated is set when a direction stopped at the cap; *n_above (if not null) receives the
// number found above.
MONI_HD uint64_t occ_walk(const ImageView& v, uint64_t head, uint64_t up, uint64_t down, uint64_t len, uint64_t cap, uint64_t* out,
                          bool* truncated, uint32_t* n_above = nullptr) {
    const uint64_t N = v.hdr.text_len, n = v.hdr.n;
    uint64_t k = 0;
    if (out) out[k] = head;
    ++k;
    if (n_above) *n_above = 0;
    for (int dir = 0; dir < 2; ++dir) {
        const PhiEntry* tab = dir == 0 ? v.phi : v.phi_inv;
        const uint32_t* bdir = dir == 0 ? v.phi_dir : v.phi_inv_dir;
        const uint64_t stop_at = dir == 0 ? v.hdr.first_run_sample : v.hdr.init_sample;   // get_first / get_last_run_sample
        uint64_t curr = dir == 0 ? up : down, taken = 0;
        if (curr < n)                    // else not a text position (a wrapped pointer): the reference's Phi is undefined there
            for (;;) {
                if (curr == stop_at) break;                                   // the reference returns lcp 0 here
                const uint64_t next = phi_step(tab, bdir, v.hdr.phi_shift, v.hdr.phi_n, n, curr);
                if (!((N - curr) >= len && (N - next) >= len)) break;         // lcp stays 0 (u64 arithmetic as in the reference)
                if (!lce_at_least(v, curr, next, len)) break;
                if (taken == cap) { if (truncated) *truncated = true; break; }
                if (out) out[k] = next;
                ++k;
                ++taken;
                curr = next;
            }
        if (dir == 0 && n_above) *n_above = (uint32_t)taken;
    }
    return k;
}
MONI_HD uint64_t occ_walk(const ImageView& v, uint64_t pos, uint64_t len, uint64_t cap, uint64_t* out, bool* truncated) {
    return occ_walk(v, pos, pos, pos, len, cap, out, truncated, nullptr);
}

}  // namespace moni
