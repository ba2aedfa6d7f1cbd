/* moni_gpu.h -- C ABI of libmoni_gpu.so: the B200 (sm_100a) implementation of the computation
 * behind `moni ms` and `moni mems`.
 *
 * The reference (maxrossi91/moni-align) has no FFI for this path; the seam is a handful of C++
 * call sites inside its two executables.  Each entry point below names the reference interface
 * it replaces (paths relative to the reference checkout).  INTEGRATION.md shows the binding a
 * maintainer would add on the reference side.
 *
 * Conventions
 *   - every int-returning function returns 0 on success and a negative MONI_GPU_E* code on
 *     failure; moni_gpu_last_error() then returns a thread-local message.  The library never
 *     calls exit() (the reference's error() does: include/common/common.hpp:108-117).
 *   - all buffers are caller-owned host memory.  Pinned memory (moni_gpu_host_alloc) makes the
 *     PCIe copies asynchronous and ~2x faster; pageable memory works.
 *   - a batch of reads is passed as concatenated bytes `seqs` plus `offsets[n_reads + 1]`
 *     (read r = seqs[offsets[r] .. offsets[r+1])).  Read bytes are used verbatim (no case folding),
 *     exactly like kseq's read->seq.s in the reference.  Results are in input order for any GPU count.
 *   - all values are full 64-bit at the ABI, like the reference's size_t vectors.
 *   - there is NO CPU execution path: without a CUDA device every call fails with MONI_GPU_ECUDA.
 */
#ifndef MONI_GPU_H
#define MONI_GPU_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MONI_GPU_OK 0
#define MONI_GPU_EINVAL (-1)   /* bad argument */
#define MONI_GPU_EIO (-2)      /* index file missing / malformed / unsupported */
#define MONI_GPU_ECUDA (-3)    /* CUDA runtime failure (incl. no device) */
#define MONI_GPU_ENOMEM (-4)   /* caller-provided capacity too small, or allocation failure */

typedef struct moni_gpu_index moni_gpu_index; /* opaque: host image + one replica per GPU */
typedef struct moni_gpu_batch moni_gpu_batch; /* opaque: a read batch resident in one GPU's HBM */

/* Replaces ms_c<slp_t>::ms_c(std::string) / mems_c::mems_c  (src/matching_statistics.cpp:195-228,
 * src/mems.cpp:195-228): reads `<prefix>.thrbv.ms` (ms_pointers::load, include/ms/ms_pointers.hpp:307-320)
 * and `<prefix>.plain.slp` (PlainSlp::load, thirdparty/ShapedSlp/PlainSlp.hpp:305-319), flattens them
 * and uploads one replica to every listed CUDA device.  gpus == NULL: devices 0..n_gpus-1.
 * If `<prefix>.thrbv.ms` does not exist, the `moni align` index `<prefix>.thrbv.full.ms` (ms_pointers<>::load,
 * include/ms/moni.hpp:392-409) or `<prefix>.thrbv.full.lcp.ms` (moni_lcp<>::load, include/aligner/moni_lcp.hpp:208-225) is
 * read instead: same query structures, the Phi / LCP members are skipped.  The grammar is `<prefix>.plain.slp`; when only
 * `<prefix>.slp` exists, or after moni_gpu_set_option("grammar", "shaped"), the shaped grammar is read instead
 * (SelfShapedSlp::load, thirdparty/ShapedSlp/SelfShapedSlp.hpp:378-397; ms_c<shaped_slp_t>, src/matching_statistics.cpp:624-627). */
int moni_gpu_index_load(const char* prefix, const int* gpus, int n_gpus, moni_gpu_index** out);
void moni_gpu_index_free(moni_gpu_index* idx);

/* n = ms.bwt_size() (r_index.hpp:455-457), r = number of BWT runs, text_len = ra.getLen()
 * (src/matching_statistics.cpp:221), device_bytes = size of one replica.  Any pointer may be NULL. */
int moni_gpu_index_info(const moni_gpu_index* idx, uint64_t* n, uint64_t* r, uint64_t* text_len, uint64_t* device_bytes);

/* Replaces ms_c::matching_statistics(kseq_t*, FILE*) (src/matching_statistics.cpp:242-278) for a whole
 * batch: pointers = ms.query(seq, len) (include/ms/ms_pointers.hpp:241, 455-511), lengths = the
 * charAt loop (:246-255).  Both outputs have offsets[n_reads] entries, laid out like `seqs`. */
int moni_gpu_ms(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads,
                uint64_t* pointers, uint64_t* lengths);

/* Replaces mems_c::maxrimal_exact_matches(kseq_t*, FILE*) (src/mems.cpp:241-280): for read r the records
 * (mem_pos[k], mem_len[k]), k in [mem_offsets[r], mem_offsets[r+1]), are the reference's
 * std::vector<std::pair<size_t,size_t>> mems (query offset, length).  mem_offsets has n_reads + 1
 * entries.  If mem_capacity is too small the call fails with MONI_GPU_ENOMEM after filling
 * mem_offsets (so mem_offsets[n_reads] is the capacity needed); offsets[n_reads] always suffices. */
int moni_gpu_mems(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads,
                  uint64_t* mem_offsets, uint64_t* mem_pos, uint64_t* mem_len, uint64_t mem_capacity);

/* Same records with 32-bit fields: a query offset and a match length are both smaller than the read length
 * (< 2^32, enforced), so this halves the device->host traffic of the call, which is what bounds it. */
int moni_gpu_mems32(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads,
                    uint64_t* mem_offsets, uint32_t* mem_pos, uint32_t* mem_len, uint64_t mem_capacity);

/* Same records as one 32-bit word each, (query offset << 16) | length: for batches whose reads are all shorter than
 * 65536 bases (else MONI_GPU_EINVAL).  Halves the device->host bytes of moni_gpu_mems32 again (4 bytes per record). */
int moni_gpu_mems_packed(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads,
                         uint64_t* mem_offsets, uint32_t* mem_rec, uint64_t mem_capacity);

/* Same records as one 16-bit word each, (query offset << 8) | length: for batches whose reads are all shorter than 256 bases
 * (short-read sequencing; else MONI_GPU_EINVAL).  2 bytes per record leave the device. */
int moni_gpu_mems_packed16(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads,
                           uint64_t* mem_offsets, uint16_t* mem_rec, uint64_t mem_capacity);

/* Replaces seed_finder<slp_t, ms_t>::find_mems(const kseq_t*, std::vector<mem_t>&, r_offset, mate)
 * (include/aligner/seed_finder.hpp:126-166), the seed stage of `moni align`, for a whole batch: same two passes, with the
 * record rule of :148 -- l >= pl and n_Ns < l and l >= min_len (n_Ns = the run of 'N' ending the characters matched so
 * far, :139-143).  Record k of read r, k in [seed_offsets[r], seed_offsets[r+1]): mem_t.pos = seed_ref[k] (= pointers[i]),
 * mem_t.len = seed_len[k], mem_t.idx = seed_qpos[k]; mem_t.rpos = r_offset + idx + len - 1 and mem_t.mate are the caller's.
 * Works on any of the three index files (moni_gpu_index_load).  Capacity rule as for moni_gpu_mems. */
int moni_gpu_seeds(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads, uint64_t min_len,
                   uint64_t* seed_offsets, uint64_t* seed_ref, uint32_t* seed_qpos, uint32_t* seed_len, uint64_t seed_capacity);

/* Replaces seed_finder::find_MEM_occs(mem_t&) (include/aligner/seed_finder.hpp:239-251) with seed filtering off, for a batch
 * of seeds (e.g. the records moni_gpu_seeds returned): for seed s the text positions occs[k], k in [occ_offsets[s],
 * occ_offsets[s+1]), are mem_t::occs in the reference's order -- the seed's own position, then the suffixes above it
 * (find_MEM_above :169-197, r_index::Phi, thirdparty/r-index/internal/r_index.hpp:225-243) and below it (find_MEM_below
 * :201-229, ms_pointers::Phi_inv, include/ms/moni.hpp:338-354) whose common prefix with their neighbour reaches the seed
 * length (get_prev/next_occ_with_lcp :346-370; the bounded LCE of thirdparty/ShapedSlp/Common.hpp:250-295 is a direct
 * compare of the text in HBM).  Needs an index loaded from `<prefix>.thrbv.full.ms` (the Phi structures) with the text
 * resident.  max_per_direction bounds each walk (the reference's is unbounded); *truncated (may be NULL) is set to 1 if a
 * walk hit it (with max_per_direction <= 4096 and room for 2 * max + 1 staged positions per seed the lists come from a single
 * walk; otherwise a counting walk and a filling walk).  Runs on the index's first GPU.  Capacity rule as for moni_gpu_mems.  moni_gpu_timing afterwards: t[0] seeds
 * in, t[1] counting walk + scan, t[2] filling walk, t[4] positions out, t[5] whole call. */
int moni_gpu_seed_occs(moni_gpu_index* idx, const uint64_t* seed_ref, const uint32_t* seed_len, uint64_t n_seeds,
                       uint64_t max_per_direction, uint64_t* occ_offsets, uint64_t* occs, uint64_t occ_capacity, int* truncated);

/* Replaces seed_finder::populate_seeds(std::vector<mem_t>&) (include/aligner/seed_finder.hpp:311-319 -> populate_seed :258-309) with seed
 * filtering off, for a batch of seeds (the records moni_gpu_seeds returned; seed_qpos = mem_t.idx).  Every seed gets its find_MEM_occs
 * list (occ_offsets / occs, exactly what moni_gpu_seed_occs returns).  A seed with seed_len >= 2 * min_len additionally yields the two
 * seeds the reference appends to the read's vector (:277-306), stored as two consecutive entries h, h + 1 of the half_* arrays, in seed
 * order: the left half, mem_t(upper_suffix, len >> 1, idx), whose list is upper_suffix, the suffixes above it and the suffixes below
 * lower_suffix at the shorter length (upper / lower_suffix = the last position found above / below the seed, :272-275); and the right
 * half, mem_t(pos + (len >> 1), len - (len >> 1), idx + (len >> 1)), with its own find_MEM_occs list.  half_parent[h] = index of the
 * seed the half came from; half_ref / half_len / half_qpos = mem_t.pos / .len / .idx; lists in half_occ_offsets / half_occs.  The
 * reference's per-read vector is the read's seeds followed by their halves in seed order.  *n_halves receives the number of half
 * entries (capacity needed: 2 * n_seeds).  min_len >= 1.  Per-reference seed filtering (populate_dict, :331-) needs the lift-over
 * index and is not done.  Needs the `moni align` index with the text resident, like moni_gpu_seed_occs; runs on the index's first GPU. */
int moni_gpu_populate_seeds(moni_gpu_index* idx, const uint64_t* seed_ref, const uint32_t* seed_len, const uint32_t* seed_qpos, uint64_t n_seeds,
                            uint64_t min_len, uint64_t max_per_direction, uint64_t* occ_offsets, uint64_t* occs, uint64_t occ_capacity,
                            uint64_t* n_halves, uint64_t* half_parent, uint64_t* half_ref, uint32_t* half_len, uint32_t* half_qpos,
                            uint64_t* half_occ_offsets, uint64_t* half_occs, uint64_t half_occ_capacity, int* truncated);

/* moni_gpu_ms with 7 bytes of results per base instead of 16 (the call is bound by the device->host copy): the pointer's low
 * 40 bits as ptr_lo (u32) + ptr_hi (u8), the length as u16.  pointer = lo | hi << 32, and if that is >= 2^40 - 2^31 subtract
 * 2^40 in 64-bit arithmetic: such a value is one of the reference's wrapped pointers 2^64 - k (its `sample--` after an absent
 * symbol set sample to 0, ms_pointers.hpp:470-481).  Needs reads shorter than 65536 bases and n <= 2^40 - 2^31 (else EINVAL). */
int moni_gpu_ms_packed(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads,
                       uint32_t* ptr_lo, uint8_t* ptr_hi, uint16_t* lengths);

/* moni_gpu_ms as restart runs: what ms_c::matching_statistics computes (src/matching_statistics.cpp:242-278) without the
 * redundancy of its two vectors.  The reference's length loop (:246-255) only does work at query offsets i with
 * `i == 0 || pointers[i] != pointers[i-1] + 1` (restarts); at every other offset pointers[i] = pointers[i-1] + 1 and
 * lengths[i] = lengths[i-1] - 1 (floor 0).  For read r the runs k in [run_offsets[r], run_offsets[r+1]) are its restarts in
 * increasing query offset: run_qpos[k] = i, run_ptr[k] = pointers[i], run_len[k] = lengths[i]; the first run of a non-empty
 * read is at i = 0.  moni_gpu_ms_runs_expand rebuilds the two vectors on the host (format conversion, no device work), so
 * expand(runs) == moni_gpu_ms bit for bit, wrapped pointers included.  About 1 + #mismatches runs per read instead of two
 * u64 per base leave the device.  Capacity rule as for moni_gpu_mems (offsets[n_reads] always suffices). */
int moni_gpu_ms_runs(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads, uint64_t* run_offsets,
                     uint32_t* run_qpos, uint32_t* run_len, uint64_t* run_ptr, uint64_t run_capacity);
/* Same runs in 9 bytes each: run_rec = (query offset << 16) | length (reads shorter than 65536 bases, else MONI_GPU_EINVAL) and the
 * pointer's low 40 bits as ptr_lo + ptr_hi (decoded like moni_gpu_ms_packed's; needs n <= 2^40 - 2^31). */
int moni_gpu_ms_runs_packed(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads, uint64_t* run_offsets,
                            uint32_t* run_rec, uint32_t* ptr_lo, uint8_t* ptr_hi, uint64_t run_capacity);
int moni_gpu_ms_runs_expand(const uint64_t* offsets, uint64_t n_reads, const uint64_t* run_offsets, const uint32_t* run_qpos,
                            const uint32_t* run_len, const uint64_t* run_ptr, uint64_t* pointers, uint64_t* lengths);

/* Event counters accumulated over all calls since the last reset, when counting is enabled
 * (moni_gpu_set_option("count", "1"); off by default, it costs a few percent):
 * c[0]=bases, c[1]=threshold jumps (ms_pointers.hpp:482-505 taken), c[2]=LCE restarts,
 * c[3]=grammar nodes visited by descents, c[4]=characters compared (charAt evaluations,
 * src/matching_statistics.cpp:250), c[5]=rows scanned forward after LF, c[6]=steps through the
 * general path, c[7]=descents.  SURVEY.md 8(d) defines the roofline numerator from c[0..4]. */
int moni_gpu_counters(moni_gpu_index* idx, uint64_t c[8], int reset);

/* Device time of the last moni_gpu_ms / moni_gpu_mems call on this index, in milliseconds, measured
 * with CUDA events on the launching stream and maxed over the index's GPUs:
 * t[0]=host->device copies, t[1]=pass 1 kernel, t[2]=pass 2 kernel, t[3]=MEM compaction,
 * t[4]=device->host copies, t[5]=whole call (first copy to last copy). */
int moni_gpu_timing(const moni_gpu_index* idx, double t[6]);

/* Options.  "leaf_len", "text_budget", "max_scan", "grammar", "chunk_bases", "chunk_min_reads", "lanes", "dynamic", "l2_fetch" apply to
 * indexes loaded afterwards; the others apply at once.
 *   index image
 *     "leaf_len"      grammar literal block length; default 0 = auto: the whole text as one literal block if it fits "text_budget"
 *                     bytes (default 64 GiB), else 128
 *     "grammar"       "plain" = <prefix>.plain.slp, falling back to <prefix>.slp when only that exists; "shaped" = <prefix>.slp, the
 *                     file `-q` names (SelfShapedSlp::load, thirdparty/ShapedSlp/SelfShapedSlp.hpp:378-397); default "plain"
 *     "max_scan"      move-table balancing target, default 8
 *   batching
 *     "chunk_bases"   bases per device chunk, default 2^24
 *     "chunk_min_reads"  a chunk is not closed before it holds this many reads, default 98304 (less when the batch has too few reads to
 *                     give every lane two chunks); hard cap 2^30 bases, less when free HBM is short
 *     "lanes"         chunks in flight per GPU, 1..8; default 0 = 4
 *     "guided", "ramp"   chunk-size schedules for batches of long reads, both off by default (measured no better)
 *   kernels
 *     "dynamic"       0 = one read per lane, 1 = persistent kernels whose lanes pull reads, 2 = pull for ragged chunks; default 2
 *     "dyn_reads_per_lane"  reads per thread the persistent kernels are sized for, default 2: fewer threads than reads, so that lanes
 *                     that finish a short read take another
 *     "l2_fetch"      cudaLimitMaxL2FetchGranularity hint; 0 = device default
 *   results
 *     "widen_threads" moni_gpu_mems and moni_gpu_ms: host threads that widen the results -- which then cross the link in 2 / 4 / 8 bytes
 *                     per record resp. 7 bytes per base -- into the u64 columns.  0 = off (the u64 columns cross the link); -1 = auto,
 *                     the default: two thirds of the CPUs the process may run on, at most 16, off below 4 CPUs.  moni_gpu_mems widens
 *                     with 2 pool threads or more, moni_gpu_ms with 6 or more
 *   diagnostics
 *     "verbose" (0/1), "count" (0/1: event counters, moni_gpu_counters) */
int moni_gpu_set_option(const char* key, const char* value);
/* Environment: MONI_GPU_VERBOSE=1 prints the exporter's and the index load's phase times on stderr (as option "verbose");
 * MONI_GPU_TRACE=1 prints a per-chunk timeline of every call. */

/* Pinned host memory for batch buffers (cudaHostAlloc / cudaFreeHost). */
void* moni_gpu_host_alloc(uint64_t bytes);
void moni_gpu_host_free(void* p);
/* Same, placed on the NUMA node of CUDA device `gpu` (sysfs numa_node of its PCI function; plain moni_gpu_host_alloc when the
 * node is unknown or the policy call is refused), so copies to and from that GPU stay off the socket interconnect: what a
 * caller feeding several GPUs of a two-socket host wants for its batch buffers.  moni_gpu_numa_node returns that node or -1. */
void* moni_gpu_host_alloc_near(int gpu, uint64_t bytes);
int moni_gpu_numa_node(int gpu);

/* ---- resident batches: inputs stay in HBM across runs (pipelines, kernel-only measurement) ----
 * gpu_slot indexes the `gpus` list given to moni_gpu_index_load. */
int moni_gpu_batch_create(moni_gpu_index* idx, int gpu_slot, const char* seqs, const uint64_t* offsets,
                          uint64_t n_reads, moni_gpu_batch** out);
/* what: 1 = ms (pass 1 + pass 2), 2 = mems (pass 1 + pass 2 + compaction).  Kernels only; returns the
 * CUDA-event time of each stage in milliseconds (kernel_ms[0]=pass 1, [1]=pass 2, [2]=compaction). */
int moni_gpu_batch_run(moni_gpu_batch* b, int what, float kernel_ms[3]);
int moni_gpu_batch_fetch_ms(moni_gpu_batch* b, uint64_t* pointers, uint64_t* lengths);
int moni_gpu_batch_fetch_mems(moni_gpu_batch* b, uint64_t* mem_offsets, uint64_t* mem_pos, uint64_t* mem_len,
                              uint64_t mem_capacity);
void moni_gpu_batch_free(moni_gpu_batch* b);

/* ---- allocation-free call ----
 * No counterpart in the reference: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  Computes exactly what moni_gpu_ms computes
 * -- pointers[] and lengths[] of every base, laid out like seqs -- but
 *   - seqs, offsets (n_reads + 1 entries, offsets[0] = 0, offsets[n_reads] = n_bases), pointers and lengths (n_bases entries
 *     each) are DEVICE pointers on the GPU of `gpu_slot` (index into the `gpus` list of moni_gpu_index_load), owned by the caller;
 *   - everything the kernels exchange (staged read bytes, restart flags, restart pointers / records, restarts per read) lives in
 *     one fixed array in the library's device image, about 21 MiB, which every GPU that loads the library holds until the process
 *     exits.  Hence the limits: at most MONI_GPU_NOALLOC_MAX_READS reads and MONI_GPU_NOALLOC_MAX_BASES bases per call; anything
 *     larger is refused with MONI_GPU_EINVAL and a message that says so (split the batch, or use moni_gpu_ms);
 *   - the call itself performs no cudaMalloc / cudaFree / cudaHostAlloc and calls no allocating library routine (no cub, no
 *     torch): one memset of the status word, one check kernel over offsets (bad offsets -> MONI_GPU_EINVAL), a device-to-device
 *     copy of the reads into the array, the three ms kernels (one read per lane; no length sort), on `stream` (a cudaStream_t, NULL
 *     = the default stream).  It returns after `stream` has drained, results in place.
 * All calls share the one array per GPU and it keeps the last call's contents (the call clears what has to start at zero: the
 * status word and the 16 bytes behind the reads).  Calls through this entry point are serialised inside the process; work from
 * ANOTHER process's copy of the library has its own array.  Two calls that overlap on one GPU would race on it -- there is no
 * way to do that through this synchronous entry point short of bypassing its lock, and none should be added. */
#define MONI_GPU_NOALLOC_MAX_READS 65536u
#define MONI_GPU_NOALLOC_MAX_BASES 1048576u
int moni_gpu_ms_noalloc(moni_gpu_index* idx, int gpu_slot, const char* d_seqs, const uint64_t* d_offsets, uint64_t n_reads,
                        uint64_t n_bases, uint64_t* d_pointers, uint64_t* d_lengths, void* stream);

/* Diagnostics: the random-access roofline of CUDA device `gpu` ("measured random-sector bandwidth of the box,
 * obtained by a gather micro-benchmark").  Dependent chains over a `table_bytes` table (next granule = hash of the words just
 * loaded: the access pattern of an LF step over the move table, include/ms/ms_pointers.hpp:476-481 + rle_string::rank); a miss
 * costs one 128-byte DRAM line on B200, so the result is random lines per second.  shape: 0 = one 32-byte load per access into
 * registers (what pass 1 does), 1 = one 16-byte load into registers, 2 = one 16-byte cp.async into shared memory;
 * chains_per_thread x (blocks_per_sm x 256 threads) chains are resident per SM; only the combinations the library was compiled
 * with are accepted (else MONI_GPU_EINVAL).  Not on the product path. */
int moni_gpu_probe_random_lines(int gpu, uint64_t table_bytes, int shape, int chains_per_thread, int blocks_per_sm, int iters,
                                double* lines_per_s, int* threads_per_sm);
void moni_gpu_probe_release(void);

/* Number of kernel launches issued by this process so far (all indexes). */
uint64_t moni_gpu_launch_count(void);

const char* moni_gpu_last_error(void);
const char* moni_gpu_version(void);

#ifdef __cplusplus
}
#endif
#endif /* MONI_GPU_H */
