/* TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT.
 *
 * CPU restatement of the reference's `moni ms` / `moni mems` per-read computation, in plain C,
 * following the reference's control flow query by query (the succinct containers are replaced by
 * sorted arrays that answer the same rank/select questions).  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline leg may load this file's shared object, and only as the checker.
 *
 * Pinned against the UNMODIFIED reference binaries (oracle/_ref/bin/ms, mems, built from
 * /root/reference by oracle/build_ref.sh): tests/test_oracle_golden.py compares this restatement
 * with the committed reference outputs under tests/golden/ byte for byte.
 *
 * Reference lines followed (all under /root/reference):
 *   ms_pointers<>::_query (thr_bv specialisation)     include/ms/ms_pointers.hpp:455-511
 *   ms_pointers::LF                                    include/ms/ms_pointers.hpp:264-274
 *   rle_string::rank / run_of_position / operator[]    thirdparty/r-index/internal/rle_string.hpp:230-285,198-201
 *   ms_rle_string::run_and_head_rank / run_head_select include/ms/ms_rle_string.hpp:152-167
 *   ms_rle_string::number_of_letter                    include/ms/ms_rle_string.hpp:121-124
 *   thr_bv::rank                                       include/ms/thresholds_ds.hpp:494-497
 *   r_index::get_last_run_sample                       thirdparty/r-index/internal/r_index.hpp:470-472
 *   PlainSlp::charAt                                   thirdparty/ShapedSlp/PlainSlp.hpp:101-126
 *   ms_c::matching_statistics length loop              src/matching_statistics.cpp:246-255
 *   mems_c::maxrimal_exact_matches                     src/mems.cpp:247-259
 *   seed_finder::find_mems                             include/aligner/seed_finder.hpp:126-166  (pinned against
 *       oracle/_ref/bin/ref_seeds, the reference header driven by oracle/ref_seeds_harness.cpp: tests/golden/*.seeds*)
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef uint64_t u64;
typedef uint8_t u8;

typedef struct {
    /* run-length BWT */
    u64 n, R;
    const u64* F;              /* 256 */
    const u8* heads;           /* R */
    const u64* starts;         /* R + 1 */
    const u64* samples_start;  /* R */
    const u64* samples_last;   /* R */
    const u64* thr_vals;       /* stored thresholds, letters concatenated */
    const u64* thr_off;        /* 257 offsets into thr_vals */
    /* derived per-letter run lists (built by oracle_prepare) */
    u64 count[256];            /* runs_per_letter[c].size() */
    u64 nruns[256];
    u64* crun[256];            /* global run index of the j-th c-run */
    u64* cend[256];            /* # of c in the first j+1 c-runs  ( runs_per_letter[c].select(j) + 1 ) */
    /* PlainSlp */
    u64 text_len, num_rules, alph_size;
    uint32_t start_var;
    const u8* alph;
    const uint32_t* left;
    const uint32_t* right;
    const u64* explen;
    /* event counters of SURVEY.md 8(d) */
    u64 cnt_bases, cnt_jumps, cnt_restarts, cnt_restart_depth, cnt_chars;
} oracle_index;

oracle_index* oracle_new(void) { return (oracle_index*)calloc(1, sizeof(oracle_index)); }

void oracle_free(oracle_index* ix) {
    if (!ix) return;
    for (int c = 0; c < 256; ++c) { free(ix->crun[c]); free(ix->cend[c]); }
    free(ix);
}

void oracle_set_bwt(oracle_index* ix, u64 n, u64 R, const u64* F, const u8* heads, const u64* starts,
                    const u64* samples_start, const u64* samples_last, const u64* thr_vals, const u64* thr_off) {
    ix->n = n; ix->R = R; ix->F = F; ix->heads = heads; ix->starts = starts;
    ix->samples_start = samples_start; ix->samples_last = samples_last;
    ix->thr_vals = thr_vals; ix->thr_off = thr_off;
    for (int c = 0; c < 256; ++c) { free(ix->crun[c]); free(ix->cend[c]); ix->crun[c] = ix->cend[c] = NULL; ix->count[c] = ix->nruns[c] = 0; }
    for (u64 k = 0; k < R; ++k) ix->nruns[heads[k]]++;
    for (int c = 0; c < 256; ++c)
        if (ix->nruns[c]) {
            ix->crun[c] = (u64*)malloc(ix->nruns[c] * sizeof(u64));
            ix->cend[c] = (u64*)malloc(ix->nruns[c] * sizeof(u64));
            ix->nruns[c] = 0;
        }
    for (u64 k = 0; k < R; ++k) {
        const u8 c = heads[k];
        const u64 j = ix->nruns[c]++;
        ix->count[c] += starts[k + 1] - starts[k];
        ix->crun[c][j] = k;
        ix->cend[c][j] = ix->count[c];
    }
}

void oracle_set_slp(oracle_index* ix, u64 text_len, u64 num_rules, uint32_t start_var, u64 alph_size, const u8* alph,
                    const uint32_t* left, const uint32_t* right, const u64* explen) {
    ix->text_len = text_len; ix->num_rules = num_rules; ix->start_var = start_var; ix->alph_size = alph_size;
    ix->alph = alph; ix->left = left; ix->right = right; ix->explen = explen;
}

void oracle_counters(oracle_index* ix, u64* out5, int reset) {
    out5[0] = ix->cnt_bases; out5[1] = ix->cnt_jumps; out5[2] = ix->cnt_restarts; out5[3] = ix->cnt_restart_depth; out5[4] = ix->cnt_chars;
    if (reset) ix->cnt_bases = ix->cnt_jumps = ix->cnt_restarts = ix->cnt_restart_depth = ix->cnt_chars = 0;
}

/* # of elements of the sorted array a[0..m) that are < key */
static u64 lower_bound(const u64* a, u64 m, u64 key) {
    u64 lo = 0, hi = m;
    while (lo < hi) { u64 mid = lo + (hi - lo) / 2; if (a[mid] < key) lo = mid + 1; else hi = mid; }
    return lo;
}

/* rle_string::run_of_position (rle_string.hpp:265-285); i == n yields R as observed on the reference */
static u64 run_of_position(const oracle_index* ix, u64 i) {
    if (i >= ix->n) return ix->R;
    return lower_bound(ix->starts, ix->R + 1, i + 1) - 1;   /* last run with start <= i */
}

/* run_heads.rank(run, c): # c-runs among runs [0, run) */
static u64 head_rank(const oracle_index* ix, u64 run, u8 c) { return lower_bound(ix->crun[c], ix->nruns[c], run); }

/* rle_string::rank(i, c) (rle_string.hpp:230-260) */
static u64 bwt_rank(const oracle_index* ix, u64 i, u8 c) {
    if (ix->count[c] == 0) return 0;
    if (i == ix->n) return ix->count[c];
    const u64 run = run_of_position(ix, i);
    const u64 dist = i - ix->starts[run];
    const u64 rk = head_rank(ix, run, c);
    const u64 tail = (ix->heads[run] == c) ? dist : 0;
    if (rk == 0) return tail;
    return ix->cend[c][rk - 1] + tail;
}

/* thr_bv::rank(i, c): # stored thresholds of c that are < i (thresholds_ds.hpp:494-497) */
static u64 thr_rank(const oracle_index* ix, u64 i, u8 c) {
    return lower_bound(ix->thr_vals + ix->thr_off[c], ix->thr_off[c + 1] - ix->thr_off[c], i);
}

/* ms_pointers<>::_query, thr_bv specialisation (ms_pointers.hpp:455-511) */
void oracle_ms_pointers(oracle_index* ix, const u8* pattern, u64 m, u64* ms_pointers) {
    u64 pos = ix->n - 1;                                                /* bwt_size() - 1 */
    u64 sample = (ix->samples_last[ix->R - 1] + 1) % ix->n;             /* get_last_run_sample() */
    for (u64 i = 0; i < m; ++i) {
        const u8 c = pattern[m - i - 1];
        const u64 n_c = ix->count[c];                                   /* number_of_letter(c) */
        if (n_c == 0) {
            sample = 0;
            pos = ix->F[c] + bwt_rank(ix, pos, c);                      /* LF(pos, c), rank is 0 */
        } else if (pos < ix->n && ix->heads[run_of_position(ix, pos)] == c) {
            sample--;
            pos = ix->F[c] + bwt_rank(ix, pos, c);
        } else {
            const u64 run_of_pos = run_of_position(ix, pos);
            const u64 j = head_rank(ix, run_of_pos, c);                 /* run_and_head_rank */
            const u64 k = j < 1 ? j : ix->cend[c][j - 1];
            const u64 thr_c = thr_rank(ix, pos + 1, c);
            ix->cnt_jumps++;
            if (j > thr_c) {                                            /* jump up */
                const u64 run_of_j = ix->crun[c][j - 1];                /* run_head_select(j, c) */
                sample = ix->samples_last[run_of_j];
                pos = ix->F[c] + k - 1;
            } else {                                                    /* jump down */
                const u64 run_of_j = ix->crun[c][j];                    /* run_head_select(j + 1, c) */
                sample = ix->samples_start[run_of_j];
                pos = ix->F[c] + k;
            }
        }
        ms_pointers[m - i - 1] = sample;
    }
    ix->cnt_bases += m;
}

/* PlainSlp::charAt (PlainSlp.hpp:101-126), the recursion written as a loop; *depth counts rules visited */
static u8 char_at(const oracle_index* ix, u64 pos, u64* depth) {
    u64 var = ix->start_var;
    while (var >= ix->alph_size) {
        const u64 lv = ix->left[var - ix->alph_size];
        const u64 left_len = (lv < ix->alph_size) ? 1 : ix->explen[lv - ix->alph_size];
        if (pos < left_len) var = lv;
        else { pos -= left_len; var = ix->right[var - ix->alph_size]; }
        if (depth) (*depth)++;
    }
    return ix->alph[var];
}

u8 oracle_char_at(oracle_index* ix, u64 pos) { return char_at(ix, pos, NULL); }

/* Length loop shared by ms (matching_statistics.cpp:246-255) and mems (mems.cpp:247-259).
 * mems may be NULL; otherwise it receives (i, len) pairs and *n_mems their number. */
void oracle_lengths(oracle_index* ix, const u8* seq, u64 m, const u64* pointers, u64* lengths, u64* mems, u64* n_mems) {
    const u64 n = ix->text_len;                                         /* ra.getLen() */
    u64 l = 0, q = 0;
    for (u64 i = 0; i < m; ++i) {
        const u64 pos = pointers[i];
        const int restart = (i < 1 || pos != (pointers[i - 1] + 1));
        if (restart) {
            ix->cnt_restarts++;
            if ((i + l) < m && (pos + l) < n) { u64 d = 0; char_at(ix, pos + l, &d); ix->cnt_restart_depth += d; }
        }
        while ((i + l) < m && (pos + l) < n && restart) {
            ix->cnt_chars++;
            if (seq[i + l] != char_at(ix, pos + l, NULL)) break;
            ++l;
        }
        lengths[i] = l;
        l = (l == 0 ? 0 : (l - 1));
        if (mems && ((i == 0) || (lengths[i] >= lengths[i - 1]))) { mems[2 * q] = i; mems[2 * q + 1] = lengths[i]; ++q; }
    }
    if (n_mems) *n_mems = q;
}

/* Whole batches in the C-ABI layout: reads concatenated, offsets[n_reads + 1]. */
void oracle_ms_batch(oracle_index* ix, const u8* seqs, const u64* offsets, u64 n_reads, u64* pointers, u64* lengths) {
    for (u64 r = 0; r < n_reads; ++r) {
        const u64 o = offsets[r], m = offsets[r + 1] - o;
        oracle_ms_pointers(ix, seqs + o, m, pointers + o);
        oracle_lengths(ix, seqs + o, m, pointers + o, lengths + o, NULL, NULL);
    }
}

/* mem_pos/mem_len need offsets[n_reads] entries (upper bound); mem_offsets has n_reads + 1 entries. */
void oracle_mems_batch(oracle_index* ix, const u8* seqs, const u64* offsets, u64 n_reads, u64* mem_offsets, u64* mem_pos, u64* mem_len) {
    u64 total = 0, cap = 0;
    for (u64 r = 0; r < n_reads; ++r) { u64 m = offsets[r + 1] - offsets[r]; if (m > cap) cap = m; }
    u64* ptr = (u64*)malloc((cap + 1) * sizeof(u64));
    u64* len = (u64*)malloc((cap + 1) * sizeof(u64));
    u64* rec = (u64*)malloc((2 * cap + 2) * sizeof(u64));
    for (u64 r = 0; r < n_reads; ++r) {
        const u64 o = offsets[r], m = offsets[r + 1] - o;
        u64 q = 0;
        oracle_ms_pointers(ix, seqs + o, m, ptr);
        oracle_lengths(ix, seqs + o, m, ptr, len, rec, &q);
        mem_offsets[r] = total;
        for (u64 t = 0; t < q; ++t) { mem_pos[total + t] = rec[2 * t]; mem_len[total + t] = rec[2 * t + 1]; }
        total += q;
    }
    mem_offsets[n_reads] = total;
    free(ptr); free(len); free(rec);
}

/* seed_finder<slp_t, ms_t>::find_mems (include/aligner/seed_finder.hpp:126-166), statement by statement: the same
 * pointers (ms.query, include/ms/moni.hpp:568-626 == ms_pointers.hpp:455-511), the same extension loop with the
 * trailing-N counter (:139-143), and the record rule of :148.  rec receives (pointers[i], l, i) triples. */
void oracle_find_mems(oracle_index* ix, const u8* seq, u64 m, const u64* pointers, u64 min_len, u64* rec, u64* n_rec) {
    const u64 n = ix->text_len;                                         /* seed_finder::n = ra.getLen() (:91) */
    u64 l = 0, pl = 0, n_Ns = 0, q = 0;
    u64 prev_pos_plus_one = n + 1;
    for (u64 i = 0; i < m; ++i) {
        const u64 pos = pointers[i];
        while (pos != prev_pos_plus_one && (i + l) < m && (pos + l) < n && seq[i + l] == char_at(ix, pos + l, NULL)) {
            if (seq[i + l] == 'N') n_Ns++;
            else n_Ns = 0;
            ++l;
        }
        if (l >= pl && n_Ns < l && l >= min_len) { rec[3 * q] = pointers[i]; rec[3 * q + 1] = l; rec[3 * q + 2] = i; ++q; }
        pl = l;
        l = (l == 0 ? 0 : (l - 1));
        prev_pos_plus_one = pos + 1;
    }
    *n_rec = q;
}

/* seed_ref/seed_qpos/seed_len need offsets[n_reads] entries (upper bound); seed_offsets has n_reads + 1 entries. */
void oracle_seeds_batch(oracle_index* ix, const u8* seqs, const u64* offsets, u64 n_reads, u64 min_len, u64* seed_offsets,
                        u64* seed_ref, u64* seed_qpos, u64* seed_len) {
    u64 total = 0, cap = 0;
    for (u64 r = 0; r < n_reads; ++r) { u64 m = offsets[r + 1] - offsets[r]; if (m > cap) cap = m; }
    u64* ptr = (u64*)malloc((cap + 1) * sizeof(u64));
    u64* rec = (u64*)malloc((3 * cap + 3) * sizeof(u64));
    for (u64 r = 0; r < n_reads; ++r) {
        const u64 o = offsets[r], m = offsets[r + 1] - o;
        u64 q = 0;
        oracle_ms_pointers(ix, seqs + o, m, ptr);
        oracle_find_mems(ix, seqs + o, m, ptr, min_len, rec, &q);
        seed_offsets[r] = total;
        for (u64 t = 0; t < q; ++t) { seed_ref[total + t] = rec[3 * t]; seed_len[total + t] = rec[3 * t + 1]; seed_qpos[total + t] = rec[3 * t + 2]; }
        total += q;
    }
    seed_offsets[n_reads] = total;
    free(ptr); free(rec);
}

/* ---- occurrence enumeration of `moni align` (seed filtering off) -------------------------------------------------
 * r_index::Phi (thirdparty/r-index/internal/r_index.hpp:225-243), ms_pointers::Phi_inv (include/ms/moni.hpp:338-354),
 * sparse_sd_vector::predecessor_rank_circular (sparse_sd_vector.hpp:168-172), seed_finder::get_prev/next_occ_with_lcp
 * (include/aligner/seed_finder.hpp:346-370), find_MEM_above / find_MEM_below / find_MEM_occs (:169-251).
 * The bounded LCE (thirdparty/ShapedSlp/Common.hpp:250-295) is only ever tested for ">= len" by its callers, which is
 * what lce_reaches answers character by character.  Pinned against oracle/_ref/bin/ref_seeds ... occs (tests/golden/*.occs12). */
typedef struct {
    u64 R;
    const u64* pred; const u64* pred_to_run;               /* r_index::pred (positions of the set bits), pred_to_run */
    const u64* pred_start; const u64* pred_start_to_run;   /* ms_pointers::pred_start, pred_start_to_run */
} oracle_phi;

static oracle_phi g_phi;   /* one index at a time is enough for the tests */

void oracle_set_phi(u64 R, const u64* pred, const u64* pred_to_run, const u64* pred_start, const u64* pred_start_to_run) {
    g_phi.R = R; g_phi.pred = pred; g_phi.pred_to_run = pred_to_run; g_phi.pred_start = pred_start; g_phi.pred_start_to_run = pred_start_to_run;
}

static u64 pred_rank_circular(const u64* ones, u64 m, u64 i) {          /* rank(i) == 0 ? number_of_1() - 1 : rank(i) - 1 */
    const u64 rk = lower_bound(ones, m, i);
    return rk == 0 ? m - 1 : rk - 1;
}

static u64 phi(const oracle_index* ix, u64 i) {
    const u64 jr = pred_rank_circular(g_phi.pred, g_phi.R, i);
    const u64 j = g_phi.pred[jr];
    const u64 delta = j < i ? i - j : i + 1;
    const u64 prev_sample = ix->samples_last[g_phi.pred_to_run[jr] - 1];
    return (prev_sample + delta) % ix->n;
}

static u64 phi_inv(const oracle_index* ix, u64 i) {
    const u64 jr = pred_rank_circular(g_phi.pred_start, g_phi.R, i);
    const u64 j = g_phi.pred_start[jr];
    const u64 delta = j < i ? i - j : i + 1;
    const u64 prev_sample = ix->samples_start[g_phi.pred_start_to_run[jr] + 1];
    return (prev_sample + delta) % ix->n;
}

static int lce_reaches(const oracle_index* ix, u64 a, u64 b, u64 len) {
    for (u64 k = 0; k < len; ++k)
        if (char_at(ix, a + k, NULL) != char_at(ix, b + k, NULL)) return 0;
    return 1;
}

/* find_MEM_occs: occs = [pos] + above + below.  Returns the count; writes at most cap entries.
 * Generalised to what populate_seed asks of the left half of a split MEM (seed_finder.hpp:277-291): the list starts with `head`,
 * find_MEM_above starts at `up`, find_MEM_below at `down`; *n_above (if not NULL) receives the number found above. */
u64 oracle_find_mem_occs3(oracle_index* ix, u64 head, u64 up, u64 down, u64 len, u64* occs, u64 cap, u64* n_above) {
    const u64 n = ix->text_len;                                          /* seed_finder::n */
    const u64 first_run_sample = (ix->samples_start[0] + 1) % ix->n;
    const u64 last_run_sample = (ix->samples_last[ix->R - 1] + 1) % ix->n;
    u64 k = 0;
    if (k < cap) occs[k] = head;
    ++k;
    for (u64 curr = up;;) {                                              /* find_MEM_above */
        if (curr == first_run_sample) break;
        const u64 prev = phi(ix, curr);
        u64 ok = 0;
        if ((n - curr) >= len && (n - prev) >= len) ok = lce_reaches(ix, curr, prev, len);
        if (!ok) break;
        if (k < cap) occs[k] = prev;
        ++k;
        curr = prev;
    }
    if (n_above) *n_above = k - 1;
    for (u64 curr = down;;) {                                            /* find_MEM_below */
        if (curr == last_run_sample) break;
        const u64 next = phi_inv(ix, curr);
        u64 ok = 0;
        if ((n - curr) >= len && (n - next) >= len) ok = lce_reaches(ix, curr, next, len);
        if (!ok) break;
        if (k < cap) occs[k] = next;
        ++k;
        curr = next;
    }
    return k;
}

u64 oracle_find_mem_occs(oracle_index* ix, u64 pos, u64 len, u64* occs, u64 cap) {
    return oracle_find_mem_occs3(ix, pos, pos, pos, len, occs, cap, NULL);
}
