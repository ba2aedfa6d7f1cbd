// libmoni_gpu.so: CUDA kernels (sm_100a) and the C ABI of include/moni_gpu.h.
//
// Kernel mapping (DESIGN.md "Kernels"): one read per thread, 32 reads per warp.  Every backward step
// of pass 1 is one dependent random 32-byte sector (a MoveRow, fetched with one LDG.256), so
// throughput comes from the ~10^5 independent reads in flight per GPU, not from cooperation inside
// a read.  What the warp does share is the traffic around the walk: the read bytes of its 32 reads
// are staged through shared memory 32 steps at a time with coalesced loads, and the pointers are
// stored warp-transposed (step-major, lane-minor) so that every store instruction writes one
// contiguous 256-byte line.  The per-read arithmetic lives in ms_core.cuh.
// No tensor cores, no CPU fallback: without a device every entry point fails.
#include <cuda_runtime.h>
#include <sys/syscall.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cctype>
#include <chrono>
#include <cstring>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "../../include/moni_gpu.h"
#include "device_image.hpp"
#include "host_widen.hpp"
#include "index_files.hpp"
#include "ms_core.cuh"

using namespace moni;

// ================================================================================ kernels
namespace {

#ifndef MONI_P1_MINBLOCKS
#define MONI_P1_MINBLOCKS 12         // resident CTAs per SM the pass-1 kernels are compiled for (register cap 65536 / (128 * 12) -> 40, no spills)
#endif
#ifndef MONI_P2_MINBLOCKS
#define MONI_P2_MINBLOCKS 8          // pass 2: 64 registers, 1024 threads per SM -- faster than 40 registers x 1536 threads (4.6 vs 6.2 ms
#endif                               // per 2 M cfg3 reads, profiles/r02_h_ab_variants.json): more reads than that thrash L1
constexpr int kBlock = 128;          // 4 warps; warps never synchronise with each other after the prologue
constexpr int kTile = 32;            // steps staged per refill
constexpr int kTileStride = 36;      // bytes per read in the tile: 9 words, odd -> conflict-free column reads

__device__ __forceinline__ void flush_counters(const Counters& c, Counters* g) {
    atomicAdd(&g->bases, c.bases);
    atomicAdd(&g->jumps, c.jumps);
    atomicAdd(&g->restarts, c.restarts);
    atomicAdd(&g->restart_depth, c.restart_depth);
    atomicAdd(&g->chars, c.chars);
    atomicAdd(&g->scan_rows, c.scan_rows);
    atomicAdd(&g->general_steps, c.general_steps);
    atomicAdd(&g->descents, c.descents);
}

// How the reads of a chunk map to threads and where a read keeps what the two passes exchange.
//   slot  = global thread index; read = perm ? perm[slot] : slot   (perm sorts ragged chunks by length, longest first, so
//           the 32 reads of a warp retire together)
//   read r at chunk-relative offset o: flag words at flags[flag_base(o, r)], restart pointers (pass 1) and restart records
//           (pass 2) at rptr / rrec[run_base(o, r)], filled from index 0 in the order of the backward scan.
struct ChunkMap {
    const uint64_t* offs;      // n_reads + 1 absolute offsets into the caller's seqs
    const uint32_t* perm;      // or null
    uint64_t base;             // offs[0]: seqs on the device start here
    uint32_t n_reads;
    uint64_t seq_bytes, flag_words, run_entries;      // extents of the chunk's buffers (MONI_BOUNDS_CHECK builds test against them)
};

// First restart entry of read r: a multiple of 4 entries (one 32-byte sector of pointers), at least m entries before the
// next read's ((o + m + 4 (r + 1)) & ~3) - ((o + 4 r) & ~3) >= m + 1.  At most bases + 4 n_reads + 4 entries per chunk.
__host__ __device__ __forceinline__ uint64_t run_base(uint64_t o, uint64_t r) { return (o + 4 * r) & ~3ull; }

__device__ __forceinline__ uint32_t warp_max(uint32_t x) {
#pragma unroll
    for (int o = 16; o; o >>= 1) x = max(x, __shfl_xor_sync(0xFFFFFFFFu, x, o));
    return x;
}

// ---- chunk preparation (ragged chunks only) -----------------------------------------------------
__global__ void k_read_len(const uint64_t* __restrict__ offs, uint32_t n, uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    keys[r] = (uint32_t)(offs[r + 1] - offs[r]);
    vals[r] = r;
}

// ---- pass 1: MS pointers, kept at the restarts ------------------------------------------------------
// One read per thread.  Output per read (ms_core.cuh "restart runs"): the restart flags, the pointers at the restarts in scan
// order, and their number.  A lane's pointers are staged four at a time in shared memory and leave as one aligned 32-byte
// sector, so DRAM sees whole-sector writes although every lane appends to its own list at its own pace.
constexpr int kStage = 4;

template <bool COUNT>
__global__ void __launch_bounds__(kBlock, MONI_P1_MINBLOCKS) k_ms_pointers(ImageView v, const uint8_t* __restrict__ seqs, ChunkMap cm,
                                                        uint64_t* __restrict__ rptr, uint32_t* __restrict__ flags,
                                                        uint32_t* __restrict__ rcnt, Counters* g) {
    __shared__ uint8_t tcode[256];
    __shared__ __align__(16) uint8_t tiles[kBlock / 32][32 * kTileStride];
    __shared__ uint64_t stage[kStage][kBlock];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) tcode[i] = v.sym->tcode[i];
    __syncthreads();
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31, w = slot >> 5;
    if ((w << 5) >= cm.n_reads) return;                       // warp-uniform
    uint8_t* tile = tiles[threadIdx.x >> 5];
    uint32_t o = 0, m = 0, r = 0;                             // chunk-relative offset (< 2^32), length, read
    if (slot < cm.n_reads) {
        r = cm.perm ? __ldg(cm.perm + slot) : slot;
        const uint64_t a = __ldg(cm.offs + r);
        o = (uint32_t)(a - cm.base);
        m = (uint32_t)(__ldg(cm.offs + r + 1) - a);
    }
    const uint32_t mw = warp_max(m);
    uint64_t* const out = rptr + run_base(o, r);
    uint32_t* const fl = flags + flag_base(o, r);
    ScanState s;
    scan_init(v, s);
    Counters c = {0, 0, 0, 0, 0, 0, 0, 0};
    uint32_t acc = 0, k = 0;
    auto push = [&](uint64_t x) {
        stage[k & (kStage - 1)][threadIdx.x] = x;
        ++k;
        MONI_ASSERT(k <= m && run_base(o, r) + k <= cm.run_entries);
        if ((k & (kStage - 1)) == 0) {
            ulonglong2* dst = reinterpret_cast<ulonglong2*>(out + (k - kStage));
            dst[0] = make_ulonglong2(stage[0][threadIdx.x], stage[1][threadIdx.x]);
            dst[1] = make_ulonglong2(stage[2][threadIdx.x], stage[3][threadIdx.x]);
        }
    };
    MONI_ASSERT(flag_base(o, r) + (m + 31) / 32 <= cm.flag_words && (uint64_t)o + m + 16 <= cm.seq_bytes);
    for (uint32_t t0 = 0; t0 < mw; t0 += kTile) {
        __syncwarp();
        // refill: lane L fetches the byte of step t0 + L for each of the warp's reads in turn (one
        // 32-byte window per read: 1-2 sectors per load instruction instead of 32)
#pragma unroll 4
        for (int rr = 0; rr < 32; ++rr) {
            const uint32_t o_r = __shfl_sync(0xFFFFFFFFu, o, rr), m_r = __shfl_sync(0xFFFFFFFFu, m, rr);
            const uint32_t step = t0 + lane;
            if (step < m_r) tile[rr * kTileStride + lane] = __ldg(seqs + (uint64_t)o_r + (m_r - 1 - step));
        }
        __syncwarp();
        const uint32_t lim = min((uint32_t)kTile, mw - t0);
        for (uint32_t t = 0; t < lim; ++t) {
            const uint32_t step = t0 + t;
            if (step < m) {
                const uint64_t old = s.sample;                // pointers[j + 1]
                scan_step(v, tcode, tile[lane * kTileStride + t], s, COUNT ? &c : nullptr);
                // ms_core.cuh run_note, with the pointer list staged
                const uint32_t i = m - step;                  // j + 1
                if (i < m) {
                    if (old != s.sample + 1) { push(old); acc |= 1u << (i & 31u); }
                    if ((i & 31u) == 0) { fl[i >> 5] = acc; acc = 0; }
                }
                if (i == 1) { push(s.sample); fl[0] = acc | 1u; }
            }
        }
    }
    for (uint32_t q = k & ~(uint32_t)(kStage - 1); q < k; ++q) out[q] = stage[q & (kStage - 1)][threadIdx.x];
    if (slot < cm.n_reads) rcnt[r] = k;
    if (COUNT) { c.bases = m; flush_counters(c, g); }
}

// ---- dynamic variants: every lane pulls its next read from a counter ------------------------------------
// A read is a chain of dependent steps (a 10 kbp read is 10^4 of them, each a DRAM round trip), so a kernel whose lanes each
// own one read lasts as long as its longest read while the short ones leave their lanes idle.  Here the kernel is persistent
// (one resident wave) and a lane that finishes a read takes the next one of the chunk (ragged chunks hand them out longest
// first): the memory system sees the same number of chains in flight from the first read to the last.
__device__ __forceinline__ uint32_t pull_slot(uint32_t* ctr, bool want, uint32_t lane) {
    const uint32_t mask = __ballot_sync(0xFFFFFFFFu, want);
    if (!mask) return 0xFFFFFFFFu;
    const uint32_t leader = (uint32_t)__ffs((int)mask) - 1u;
    uint32_t first = 0;
    if (lane == leader) first = atomicAdd(ctr, (uint32_t)__popc(mask));
    first = __shfl_sync(0xFFFFFFFFu, first, leader);
    return want ? first + (uint32_t)__popc(mask & ((1u << lane) - 1u)) : 0xFFFFFFFFu;
}

__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void* gptr) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(gptr) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Pass 1, dynamic.  The bases of a lane's read arrive in shared memory as aligned 16-byte blocks fetched with cp.async one
// block ahead of the scan (two half-windows per lane), so the read bytes cost no registers and no exposed latency.
template <bool COUNT>
__global__ void __launch_bounds__(kBlock, MONI_P1_MINBLOCKS) k_ms_pointers_dyn(ImageView v, const uint8_t* __restrict__ seqs, ChunkMap cm,
                                                            uint64_t* __restrict__ rptr, uint32_t* __restrict__ flags,
                                                            uint32_t* __restrict__ rcnt, uint32_t* __restrict__ next, Counters* g) {
    __shared__ uint8_t tcode[256];
    __shared__ __align__(16) uint8_t win[2][kBlock][16];
    __shared__ uint64_t stage[kStage][kBlock];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) tcode[i] = v.sym->tcode[i];
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t wsm = (uint32_t)__cvta_generic_to_shared(&win[0][threadIdx.x][0]);
    constexpr uint32_t kHalf = kBlock * 16;
    bool have = false, dry = false;
    uint32_t o = 0, m = 0, r = 0, j = 0, acc = 0, k = 0;
    uint64_t* out = rptr;
    uint32_t* fl = flags;
    ScanState s;
    scan_init(v, s);
    Counters c = {0, 0, 0, 0, 0, 0, 0, 0};
    auto push = [&](uint64_t x) {
        stage[k & (kStage - 1)][threadIdx.x] = x;
        ++k;
        if ((k & (kStage - 1)) == 0) {
            ulonglong2* dst = reinterpret_cast<ulonglong2*>(out + (k - kStage));
            dst[0] = make_ulonglong2(stage[0][threadIdx.x], stage[1][threadIdx.x]);
            dst[1] = make_ulonglong2(stage[2][threadIdx.x], stage[3][threadIdx.x]);
        }
    };
    for (;;) {
        const bool want = !have && !dry;
        const uint32_t slot = pull_slot(next, want, lane);
        if (want) {
            if (slot < cm.n_reads) {
                r = cm.perm ? __ldg(cm.perm + slot) : slot;
                const uint64_t a = __ldg(cm.offs + r);
                o = (uint32_t)(a - cm.base);
                m = (uint32_t)(__ldg(cm.offs + r + 1) - a);
                if (m == 0) {
                    rcnt[r] = 0;
                } else {
                    out = rptr + run_base(o, r);
                    fl = flags + flag_base(o, r);
                    MONI_ASSERT(flag_base(o, r) + (m + 31) / 32 <= cm.flag_words && (uint64_t)o + m + 16 <= cm.seq_bytes);
                    MONI_ASSERT(run_base(o, r) + m <= cm.run_entries);
                    scan_init(v, s);
                    acc = 0;
                    k = 0;
                    j = m - 1;
                    have = true;
                    const uint32_t blk = (o + j) & ~15u;                     // block of the read's last base, and the one below it
                    MONI_ASSERT((uint64_t)blk + 16 <= cm.seq_bytes);
                    cp_async16(wsm + ((blk >> 4) & 1u) * kHalf, seqs + blk);
                    if (blk >= 16u && blk > o) cp_async16(wsm + (((blk >> 4) & 1u) ^ 1u) * kHalf, seqs + (blk - 16u));
                    cp_async_commit();
                    cp_async_wait_all();
                    if (COUNT) c.bases += m;
                }
            } else {
                dry = true;
            }
        }
        if (!__any_sync(0xFFFFFFFFu, have)) break;
        if (have) {
            const uint32_t at = o + j;
            const uint8_t ch = win[(at >> 4) & 1u][threadIdx.x][at & 15u];
            const uint64_t old = s.sample;                                    // pointers[j + 1]
            scan_step(v, tcode, ch, s, COUNT ? &c : nullptr);
            const uint32_t i = j + 1;                                         // ms_core.cuh run_note, with the pointer list staged
            if (i < m) {
                if (old != s.sample + 1) { push(old); acc |= 1u << (i & 31u); }
                if ((i & 31u) == 0) { fl[i >> 5] = acc; acc = 0; }
            }
            if (j == 0) {
                push(s.sample);
                fl[0] = acc | 1u;
                for (uint32_t q = k & ~(uint32_t)(kStage - 1); q < k; ++q) out[q] = stage[q & (kStage - 1)][threadIdx.x];
                rcnt[r] = k;
                have = false;
            } else {
                --j;
                const uint32_t nat = o + j;
                if ((nat & 15u) == 15u) {                                     // entering the block below: requested 16 steps ago
                    cp_async_wait_all();
                    const uint32_t blk = nat & ~15u;
                    if (blk >= 16u && blk > o) {                              // the read reaches into the block below that one
                        cp_async16(wsm + (((blk >> 4) & 1u) ^ 1u) * kHalf, seqs + (blk - 16u));
                        cp_async_commit();
                    }
                }
            }
        }
    }
    if (COUNT) flush_counters(c, g);
}

// Pass 2, dynamic: a lane walks the restarts of its read (ms_core.cuh runs_step) and then takes the next read.
template <int MODE, bool COUNT>
__global__ void __launch_bounds__(kBlock, MONI_P2_MINBLOCKS) k_ms_lengths_dyn(ImageView v, RunBufs B, ChunkMap cm, const uint32_t* __restrict__ rcnt,
                                                                              uint32_t* __restrict__ counts, uint32_t* __restrict__ next, Counters* g,
                                                                              uint32_t min_len) {
    const uint32_t lane = threadIdx.x & 31;
    bool have = false, dry = false;
    uint32_t r = 0;
    RunRead R;
    RunState st;
    Counters c = {0, 0, 0, 0, 0, 0, 0, 0};
    for (;;) {
        const bool want = !have && !dry;
        const uint32_t slot = pull_slot(next, want, lane);
        if (want) {
            if (slot < cm.n_reads) {
                r = cm.perm ? __ldg(cm.perm + slot) : slot;
                const uint64_t a = __ldg(cm.offs + r);
                R.o = (uint32_t)(a - cm.base);
                R.m = (uint32_t)(__ldg(cm.offs + r + 1) - a);
                if (R.m == 0) {
                    if (MODE) counts[r] = 0;
                } else {
                    R.cnt = __ldg(rcnt + r);
                    R.fb = (uint32_t)flag_base(R.o, r);
                    R.last = (uint32_t)run_base(R.o, r) + R.cnt - 1u;
                    st = RunState();
                    have = true;
                }
            } else {
                dry = true;
            }
        }
        if (!__any_sync(0xFFFFFFFFu, have)) break;
        if (have) {
            runs_step<MODE == 2>(v, B, R, st, COUNT ? &c : nullptr, min_len);
            if (st.mode == RunState::FINISHED) {
                if (MODE) counts[r] = st.n_mems;
                have = false;
            }
        }
    }
    if (COUNT) flush_counters(c, g);
}

// ---- pass 2: MS lengths at the restarts / MEM records / seeds ----------------------------------------
// MODE 0 (ms, runs): rrec[] receives (i << 32) | len of every restart, next to its pointer.  MODE 1 (mems): the read's MEM
// records, packed (i << 32) | len, go to rec[] at the read's own offset (at most one per base) and counts[read] receives their
// number.  MODE 2 (seeds): same with seed_finder::find_mems' rule (minimum length, N filter) and rec_ref[] = pointers[i].
template <int MODE, bool COUNT>
__global__ void __launch_bounds__(kBlock, MONI_P2_MINBLOCKS) k_ms_lengths(ImageView v, RunBufs B, ChunkMap cm, const uint32_t* __restrict__ rcnt,
                                                                          uint32_t* __restrict__ counts, Counters* g, uint32_t min_len) {
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= cm.n_reads) return;
    const uint32_t r = cm.perm ? __ldg(cm.perm + slot) : slot;
    const uint64_t a = __ldg(cm.offs + r);
    RunRead R;
    R.o = (uint32_t)(a - cm.base);
    R.m = (uint32_t)(__ldg(cm.offs + r + 1) - a);
    if (R.m == 0) { if (MODE) counts[r] = 0; return; }
    R.cnt = __ldg(rcnt + r);
    R.fb = (uint32_t)flag_base(R.o, r);
    R.last = (uint32_t)run_base(R.o, r) + R.cnt - 1u;
    Counters c = {0, 0, 0, 0, 0, 0, 0, 0};
    const uint32_t k = runs_read<MODE == 2>(v, B, R, COUNT ? &c : nullptr, min_len);
    if (MODE) counts[r] = k;
    if (COUNT) flush_counters(c, g);
}

// ms: restart runs -> the caller's layout (pointers[] / lengths[] laid out like seqs).  One warp per read, 32 query offsets per
// trip: every lane finds the restart at or before its offset in the flag word (or the carry from earlier words), reads that
// restart's pointer and length, and adds / subtracts the distance -- what the reference's loop does one offset at a time
// (src/matching_statistics.cpp:246-255).  Stores are 256 bytes along the read.
// PACKED: 7 bytes per base instead of 16 -- the low 40 bits of the pointer as u32 + u8 (a pointer is a text position < n,
// or 2^64 - k with k < 2^31 after the reference's `sample--` wrapped: both survive `& (2^40 - 1)` when n <= 2^40 - 2^31)
// and the length as u16 (reads shorter than 65536 bases).
template <bool PACKED>
__global__ void __launch_bounds__(kBlock) k_ms_expand(ChunkMap cm, const uint64_t* __restrict__ rptr, const uint64_t* __restrict__ rrec,
                                                      const uint32_t* __restrict__ flags, const uint32_t* __restrict__ rcnt,
                                                      uint64_t* __restrict__ ptr_out, uint64_t* __restrict__ len_out,
                                                      uint32_t* __restrict__ lo_out, uint8_t* __restrict__ hi_out,
                                                      uint16_t* __restrict__ len16_out) {
    const uint32_t r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (r >= cm.n_reads) return;
    const uint64_t a = __ldg(cm.offs + r);
    const uint64_t o = a - cm.base, m = __ldg(cm.offs + r + 1) - a;
    if (m == 0) return;
    const uint32_t cnt = __ldg(rcnt + r);
    const uint64_t* const rp = rptr + run_base(o, r) + (cnt - 1);     // restart k (in increasing query offset) is rp[-k]
    const uint64_t* const rr = rrec + run_base(o, r) + (cnt - 1);
    const uint32_t* const fl = flags + flag_base(o, r);
    uint32_t seen = 0;                                                // restarts in earlier words
    uint64_t carry = 0;                                               // query offset of the last of them
    for (uint64_t i0 = 0; i0 < m; i0 += 32) {
        const uint32_t word = __ldg(fl + (i0 >> 5));
        const uint32_t upto = word & (0xFFFFFFFFu >> (31u - lane));    // restarts at offsets i0 .. i0 + lane
        const uint64_t at = upto ? i0 + (31u - (uint32_t)__clz((int)upto)) : carry;
        const uint32_t kk = seen + (uint32_t)__popc(upto) - 1u;         // offset 0 is a restart, so kk >= 0 from the first word on
        const uint64_t i = i0 + lane;
        if (i < m) {
            MONI_ASSERT(kk < cnt && at <= i);
            const uint64_t d = i - at;
            const uint64_t pv = __ldg(rp - (int64_t)kk) + d;
            const uint64_t l0 = __ldg(rr - (int64_t)kk) & 0xFFFFFFFFull;
            const uint64_t lv = l0 > d ? l0 - d : 0;
            if (PACKED) {
                lo_out[o + i] = (uint32_t)pv;
                hi_out[o + i] = (uint8_t)(pv >> 32);
                len16_out[o + i] = (uint16_t)lv;
            } else {
                ptr_out[o + i] = pv;
                len_out[o + i] = lv;
            }
        }
        if (word) { carry = i0 + (31u - (uint32_t)__clz((int)word)); seen += (uint32_t)__popc(word); }
    }
}

// runs: the restarts of every read, compacted in increasing query offset.  WIDE: (query offset u32, length u32, pointer u64);
// else one word (query offset << 16) | length for reads shorter than 65536 bases and the pointer's low 40 bits as u32 + u8.
template <bool WIDE>
__global__ void __launch_bounds__(kBlock) k_runs_gather(const uint64_t* __restrict__ rptr, const uint64_t* __restrict__ rrec,
                                                        const uint64_t* __restrict__ offs, uint64_t base, uint32_t n_reads,
                                                        const uint64_t* __restrict__ moff, uint32_t* __restrict__ c0,
                                                        uint32_t* __restrict__ c1, uint64_t* __restrict__ c2_wide, uint8_t* __restrict__ c2_hi) {
    const uint32_t r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (r >= n_reads) return;
    const uint64_t o = offs[r] - base, a = moff[r], cnt = moff[r + 1] - a;
    if (cnt == 0) return;
    const uint64_t last = run_base(o, r) + (cnt - 1);
    for (uint64_t k = lane; k < cnt; k += 32) {
        const uint64_t x = rrec[last - k], pv = rptr[last - k];
        if (WIDE) {
            c0[a + k] = (uint32_t)(x >> 32);
            c1[a + k] = (uint32_t)x;
            c2_wide[a + k] = pv;
        } else {
            c0[a + k] = ((uint32_t)(x >> 32) << 16) | ((uint32_t)x & 0xFFFFu);
            c1[a + k] = (uint32_t)pv;
            c2_hi[a + k] = (uint8_t)(pv >> 32);
        }
    }
}

// MEM compaction: one warp per read copies its packed records to their final place.
template <typename OutT>
__global__ void __launch_bounds__(kBlock) k_mem_gather(const uint64_t* __restrict__ rec, const uint64_t* __restrict__ offs,
                                                       uint64_t base, uint32_t n_reads, const uint64_t* __restrict__ moff,
                                                       OutT* __restrict__ mem_pos, OutT* __restrict__ mem_len) {
    const uint32_t r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (r >= n_reads) return;
    const uint64_t o = offs[r] - base, a = moff[r], cnt = moff[r + 1] - a;
    for (uint64_t k = lane; k < cnt; k += 32) {
        const uint64_t x = rec[o + k];
        mem_pos[a + k] = (OutT)(x >> 32);
        mem_len[a + k] = (OutT)(x & 0xFFFFFFFFull);
    }
}

// Seeds: (query offset, length) as u32 and the text position as u64.
__global__ void __launch_bounds__(kBlock) k_seed_gather(const uint64_t* __restrict__ rec, const uint64_t* __restrict__ rec_ref,
                                                        const uint64_t* __restrict__ offs, uint64_t base, uint32_t n_reads,
                                                        const uint64_t* __restrict__ moff, uint32_t* __restrict__ qpos,
                                                        uint32_t* __restrict__ len, uint64_t* __restrict__ ref) {
    const uint32_t r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (r >= n_reads) return;
    const uint64_t o = offs[r] - base, a = moff[r], cnt = moff[r + 1] - a;
    for (uint64_t k = lane; k < cnt; k += 32) {
        const uint64_t x = rec[o + k];
        qpos[a + k] = (uint32_t)(x >> 32);
        len[a + k] = (uint32_t)(x & 0xFFFFFFFFull);
        ref[a + k] = rec_ref[o + k];
    }
}

// Same, one word per record: (query offset << SHIFT) | length -- 32-bit words with SHIFT 16 for chunks whose reads are shorter
// than 2^16, 16-bit words with SHIFT 8 for reads shorter than 2^8 (what bounds the call is the bytes that leave the device).
template <typename RecT, int SHIFT>
__global__ void __launch_bounds__(kBlock) k_mem_gather_packed(const uint64_t* __restrict__ rec, const uint64_t* __restrict__ offs,
                                                              uint64_t base, uint32_t n_reads, const uint64_t* __restrict__ moff,
                                                              RecT* __restrict__ mem_rec) {
    const uint32_t r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (r >= n_reads) return;
    const uint64_t o = offs[r] - base, a = moff[r], cnt = moff[r + 1] - a;
    for (uint64_t k = lane; k < cnt; k += 32) {
        const uint64_t x = rec[o + k];
        mem_rec[a + k] = (RecT)(((uint32_t)(x >> 32) << SHIFT) | (uint32_t)(x & ((1u << SHIFT) - 1)));
    }
}

// ---- occurrence enumeration: one seed per thread, Phi / Phi^-1 walks with a bounded LCE on the text in HBM -------------
// head / up / down: first list entry, start of the walk upwards, start of the walk downwards (ms_core.cuh occ_walk; one array
// three times for plain seeds); above (may be null) receives the number of positions found above.
__global__ void __launch_bounds__(kBlock) k_occ_count(ImageView v, const uint64_t* __restrict__ head, const uint64_t* __restrict__ up,
                                                      const uint64_t* __restrict__ down, const uint32_t* __restrict__ len, uint64_t n_seeds,
                                                      uint64_t cap, uint64_t* __restrict__ counts, uint32_t* __restrict__ above,
                                                      uint32_t* __restrict__ truncated) {
    const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (i > n_seeds) return;
    if (i == n_seeds) { counts[i] = 0; return; }
    bool tr = false;
    uint32_t na = 0;
    counts[i] = occ_walk(v, head[i], up[i], down[i], len[i], cap, nullptr, &tr, &na);
    if (above) above[i] = na;
    if (tr) atomicOr(truncated, 1u);
}

__global__ void __launch_bounds__(kBlock) k_occ_fill(ImageView v, const uint64_t* __restrict__ head, const uint64_t* __restrict__ up,
                                                     const uint64_t* __restrict__ down, const uint32_t* __restrict__ len, uint64_t n_seeds,
                                                     uint64_t cap, const uint64_t* __restrict__ off, uint64_t* __restrict__ occs) {
    const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (i >= n_seeds) return;
    occ_walk(v, head[i], up[i], down[i], len[i], cap, occs + off[i], nullptr, nullptr);
}

// One walk instead of two when the per-direction cap bounds a seed's list: positions go to a fixed-stride staging area
// (stride = 2 cap + 1; HBM has the room), and a streaming pass packs them once the offsets are known.
__global__ void __launch_bounds__(kBlock) k_occ_walk_staged(ImageView v, const uint64_t* __restrict__ head, const uint64_t* __restrict__ up,
                                                            const uint64_t* __restrict__ down, const uint32_t* __restrict__ len, uint64_t n_seeds,
                                                            uint64_t cap, uint64_t stride, uint64_t* __restrict__ counts,
                                                            uint64_t* __restrict__ staging, uint32_t* __restrict__ above,
                                                            uint32_t* __restrict__ truncated) {
    const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (i > n_seeds) return;
    if (i == n_seeds) { counts[i] = 0; return; }
    bool tr = false;
    uint32_t na = 0;
    counts[i] = occ_walk(v, head[i], up[i], down[i], len[i], cap, staging + i * stride, &tr, &na);
    if (above) above[i] = na;
    if (tr) atomicOr(truncated, 1u);
}

// populate_seed's two halves of every seed of at least 2 * min_len bases (include/aligner/seed_finder.hpp:277-306), from the seed's
// own list: upper_suffix = the last position found above (the seed's position if none), lower_suffix = the last position of the
// whole list.  slot[j] = index of the seed's left half among the halves (exclusive scan of 2 * qualifies), right half at slot + 1.
__global__ void __launch_bounds__(kBlock) k_seed_qualifies(const uint32_t* __restrict__ len, uint64_t n_seeds, uint64_t min_len, uint64_t* __restrict__ slot) {
    const uint64_t j = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (j > n_seeds) return;
    slot[j] = (j < n_seeds && (uint64_t)len[j] >= (min_len << 1)) ? 2 : 0;
}

__global__ void __launch_bounds__(kBlock) k_make_halves(const uint64_t* __restrict__ ref, const uint32_t* __restrict__ len, const uint32_t* __restrict__ qpos,
                                                        uint64_t n_seeds, uint64_t min_len, const uint64_t* __restrict__ slot,
                                                        const uint64_t* __restrict__ off, const uint64_t* __restrict__ occs,
                                                        const uint32_t* __restrict__ above, uint64_t* __restrict__ h_head, uint64_t* __restrict__ h_up,
                                                        uint64_t* __restrict__ h_down, uint32_t* __restrict__ h_len, uint32_t* __restrict__ h_qpos,
                                                        uint64_t* __restrict__ h_parent) {
    const uint64_t j = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (j >= n_seeds) return;
    const uint64_t l = len[j];
    if (l < (min_len << 1)) return;
    const uint64_t a = off[j], cnt = off[j + 1] - a;
    const uint64_t upper = occs[a + above[j]], lower = occs[a + cnt - 1];     // occs[a] is the seed itself, then the positions above
    const uint64_t ll = l >> 1, h = slot[j];
    h_head[h] = upper; h_up[h] = upper; h_down[h] = lower; h_len[h] = (uint32_t)ll; h_qpos[h] = qpos[j]; h_parent[h] = j;
    const uint64_t rp = ref[j] + ll;
    h_head[h + 1] = rp; h_up[h + 1] = rp; h_down[h + 1] = rp; h_len[h + 1] = (uint32_t)(l - ll); h_qpos[h + 1] = qpos[j] + (uint32_t)ll; h_parent[h + 1] = j;
}

__global__ void __launch_bounds__(kBlock) k_occ_pack(const uint64_t* __restrict__ staging, uint64_t stride, const uint64_t* __restrict__ off,
                                                     uint64_t n_seeds, uint64_t* __restrict__ occs) {
    const uint64_t i = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;      // a warp per seed
    const uint32_t lane = threadIdx.x & 31;
    if (i >= n_seeds) return;
    const uint64_t a = off[i], cnt = off[i + 1] - a;
    const uint64_t* src = staging + i * stride;
    for (uint64_t k = lane; k < cnt; k += 32) occs[a + k] = src[k];
}

// Offsets of a chunk whose reads all have the same length: computed on the device instead of crossing the host link (8 bytes per
// read are 5 % of what a batch of 150-base reads sends).
__global__ void k_uniform_offsets(uint64_t* __restrict__ offs, uint64_t base, uint32_t len, uint32_t n1) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n1) offs[i] = base + (uint64_t)i * len;
}

__global__ void k_widen_counts(const uint32_t* __restrict__ c, uint64_t* __restrict__ out, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = c[i];
    if (i == n) out[i] = 0;
}

}  // namespace

// ================================================================================ host side
namespace {

thread_local std::string g_err;
std::atomic<uint64_t> g_launches{0};
std::atomic<bool> g_widen_ready{false};   // the widening pool has the size option "widen_threads" asks for
std::atomic<int> g_dyn_rpt{2};            // option "dyn_reads_per_lane": reads per thread the persistent kernels are sized for
std::atomic<int> g_ramp{0};               // option "ramp": 0 off (default: measured no better on 1 M ONT-like reads), 1 on, 2 = for batches of long reads
std::atomic<int> g_guided{0};             // option "guided": chunks shrink towards the end of a batch of long reads (default off: measured
                                          // 5.8 against 9.6 G bases/s on 1 M ONT-like reads -- small chunks make small, latency-bound kernels)
std::atomic<bool> g_count{false};   // "count" option: read at every launch, so it can be toggled on a loaded index

struct Options {
    int grammar = 0;                          // 0 = <prefix>.plain.slp, else <prefix>.slp if that is the only one; 1 = <prefix>.slp (`-q`)
    uint64_t leaf_len = 0;                    // 0 = auto: the whole text as one literal block when it fits text_budget, else 128
    uint64_t text_budget = 64ull << 30;       // bytes of HBM the plain text may take per GPU
    uint32_t max_scan = 8;
    bool verbose = false, count = false;
    uint64_t chunk_bases = 1ull << 24;
    uint64_t chunk_min_reads = 96 * 1024;
    uint32_t l2_fetch = 0;          // 0 = leave the device default
    uint32_t lanes = 0;             // chunks in flight per GPU; 0 = auto (4, or 6 for batches of long reads)
    int dynamic = 2;                // 0 = one read per lane, 1 = persistent kernels whose lanes pull reads, 2 = auto (pull when reads are ragged)
    int widen_threads = -1;         // host threads that widen packed MEM records into the caller's u64 columns; 0 = off (u64 over the link), -1 = auto
};
Options g_opt;
inline double wall_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
std::mutex g_opt_mu;

struct CudaError : std::runtime_error {
    using std::runtime_error::runtime_error;
};
#define CK(call)                                                                                              \
    do {                                                                                                      \
        cudaError_t e_ = (call);                                                                              \
        if (e_ != cudaSuccess)                                                                                \
            throw CudaError(std::string(#call) + ": " + cudaGetErrorString(e_) + " (" __FILE__ ":" + std::to_string(__LINE__) + ")"); \
    } while (0)

struct DeviceOom : std::runtime_error {
    using std::runtime_error::runtime_error;
};

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    void reserve(size_t bytes) {
        if (bytes <= cap) return;
        if (p) CK(cudaFree(p));
        p = nullptr;
        cap = 0;
        const size_t want = bytes + bytes / 8 + 256;
        const cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaErrorMemoryAllocation) {
            cudaGetLastError();
            p = nullptr;
            size_t free_b = 0, total_b = 0;
            cudaMemGetInfo(&free_b, &total_b);
            throw DeviceOom("out of device memory: a scratch buffer of " + std::to_string(want) + " bytes is needed, " + std::to_string(free_b) +
                            " bytes are free (smaller batches or fewer lanes need less)");
        }
        CK(e);
        cap = want;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <typename T>
    T* as() const { return (T*)p; }
};

// What the host knows about a chunk before it is enqueued.
struct ChunkPlan {
    uint64_t first = 0, n = 0;          // reads [first, first + n) of the caller's batch
    uint64_t base = 0, bases = 0;       // offsets[first], total bases
    uint32_t uniform_len = 0;           // common read length, or 0 for a ragged chunk
    uint64_t max_len = 0;
    bool ragged() const { return uniform_len == 0 && max_len != 0; }
    uint64_t run_entries() const { return bases + 4 * n + 8; }        // restart pointers / records: run_base()
    uint64_t flag_words() const { return bases / 32 + n + 2; }        // flag_base()
};

// What a call wants from a chunk: decides which scratch buffers exist and how many bytes of HBM a base costs.
enum Mode { kModeMs = 0, kModeMems = 1, kModeSeeds = 2, kModeRuns = 3 };

struct Scratch {   // device buffers of one chunk
    DevBuf seqs, offs, keys_in, keys_out, vals_in, perm, cub, ctr;
    DevBuf flags, rptr, rcnt;    // pass 1 -> pass 2: restart flags, pointers at the restarts, restarts per read
    DevBuf rrec;                 // ms / runs: (query offset, length) of every restart
    DevBuf out_ptr, out_len;     // ms: results in the caller's layout (u64 each; packed: u32 pointer halves, u16 lengths)
    DevBuf out_hi;               // ms, packed outputs: bits 32..39 of the pointers
    DevBuf rec, counts, moff, mem_pos, mem_len;      // mems / seeds / runs: records per read, their compaction
    DevBuf rec_ref, mem_ref;     // seeds: text positions; runs: pointers (mem_ref)
    void release() {
        for (DevBuf* b : {&seqs, &offs, &keys_in, &keys_out, &vals_in, &perm, &cub, &ctr, &flags, &rptr, &rcnt, &rrec, &out_ptr, &out_len, &out_hi, &rec,
                          &counts, &moff, &mem_pos, &mem_len, &rec_ref, &mem_ref})
            b->release();
    }
    size_t held() const {
        size_t t = 0;
        for (const DevBuf* b : {&seqs, &offs, &keys_in, &keys_out, &vals_in, &perm, &cub, &ctr, &flags, &rptr, &rcnt, &rrec, &out_ptr, &out_len, &out_hi, &rec,
                                &counts, &moff, &mem_pos, &mem_len, &rec_ref, &mem_ref})
            t += b->cap;
        return t;
    }
    // HBM bytes per base of a chunk in this mode (plus a few dozen bytes per read): what bounds the chunk size on a device
    // whose memory is mostly taken (large indexes, several replicas or processes per GPU)
    // bytes per record of the two compacted columns for a launch_compact format: records of one word need one narrow column
    static void column_bytes(int fmt, size_t& pos_b, size_t& len_b) {
        switch (fmt) {
            case 1: pos_b = 2; len_b = 0; break;          // u16 (offset:8, length:8)
            case 2: pos_b = 4; len_b = 0; break;          // u32 (offset:16, length:16)
            case 8: pos_b = 8; len_b = 8; break;          // two u64 columns
            default: pos_b = 4; len_b = 4; break;         // 4, 12 (seeds), 16 / 9 (runs): two u32 columns
        }
    }
    static double bytes_per_base(int mode, int fmt = 8) {
        const double common = 1.0 + 8.0 + 1.0 / 8.0;                  // seqs, rptr, flags
        size_t pb = 8, lb = 8;
        column_bytes(fmt, pb, lb);
        switch (mode) {
            case kModeMs: return (common + 8.0 + 16.0 + 1.0) * 1.125;     // rrec, out_ptr + out_len, out_hi
            case kModeMems: return (common + 8.0 + (double)(pb + lb)) * 1.125;         // rec, mem_pos + mem_len
            case kModeSeeds: return (common + 8.0 + 8.0 + 16.0) * 1.125;  // rec, mem_pos + mem_len (u32), rec_ref, mem_ref
            default: return (common + 8.0 + 8.0 + 8.0 + 1.0) * 1.125;     // runs: rrec, two u32 columns, pointers, out_hi
        }
    }
    // fmt: the launch_compact format the chunk's records will take (ignored for ms)
    void reserve(const ChunkPlan& p, int mode, int fmt = 8) {
        size_t pb = 8, lb = 8;
        column_bytes(fmt, pb, lb);
        seqs.reserve(p.bases + 16);
        offs.reserve((p.n + 1) * 8);
        flags.reserve(p.flag_words() * 4);
        rptr.reserve(p.run_entries() * 8);
        rcnt.reserve((p.n + 1) * 4);
        ctr.reserve(16);
        if (p.ragged()) { keys_in.reserve(p.n * 4); keys_out.reserve(p.n * 4); vals_in.reserve(p.n * 4); perm.reserve(p.n * 4); }
        if (mode == kModeMs || mode == kModeRuns) rrec.reserve(p.run_entries() * 8);
        if (mode == kModeMs) { out_ptr.reserve(p.bases * 8 + 16); out_len.reserve(p.bases * 8 + 16); out_hi.reserve(p.bases + 16); }
        if (mode == kModeMems || mode == kModeSeeds) {
            rec.reserve(p.bases * 8 + 16);
            counts.reserve((p.n + 1) * 4);
        }
        if (mode != kModeMs) {
            moff.reserve((p.n + 1) * 16);
            mem_pos.reserve(p.bases * pb + 16);
            if (lb) mem_len.reserve(p.bases * lb + 16);
        }
        if (mode == kModeSeeds) { rec_ref.reserve(p.bases * 8 + 16); mem_ref.reserve(p.bases * 8 + 16); }
        if (mode == kModeRuns) { mem_ref.reserve(p.bases * 8 + 16); out_hi.reserve(p.bases + 16); }
    }
};

constexpr int kMaxLanes = 8;     // chunks in flight per GPU (option "lanes"): copies of one overlap the kernels of the others
constexpr int kEv = 6;           // events per chunk: start, h2d done, pass 1 done, pass 2 done, compaction done, d2h done

// Pinned host staging of a lane: packed records land here and host threads widen them into the caller's u64 columns.
struct HostBuf {
    void* p = nullptr;
    size_t cap = 0;
    void reserve(size_t bytes) {
        if (bytes <= cap) return;
        release();
        const size_t want = bytes + bytes / 4 + 4096;
        CK(cudaHostAlloc(&p, want, cudaHostAllocDefault));
        cap = want;
    }
    void release() {
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
    }
};

struct Lane {
    cudaStream_t stream = nullptr;
    Scratch scratch;
    HostBuf stage;
    std::vector<cudaEvent_t> events;     // kEv per chunk processed on this lane in the current call
    size_t used = 0;
    cudaEvent_t* take() {
        if (used + kEv > events.size()) {
            const size_t old = events.size();
            events.resize(old + kEv, nullptr);
            for (size_t i = old; i < events.size(); ++i) CK(cudaEventCreate(&events[i]));
        }
        cudaEvent_t* e = events.data() + used;     // callers finish with the pointer before the next take()
        used += kEv;
        return e;
    }
};

struct Replica {
    int dev = 0;
    int sms = 0;
    int dynamic = 2;
    std::vector<Lane> lanes;
    DevBuf rows, row_pos, crun_row, slots, nodes, pool, sym, counters;
    DevBuf phi, phi_inv, phi_dir, phi_inv_dir;            // `moni align` indexes only
    DevBuf occ_ref, occ_len, occ_off, occ_out, occ_flag, occ_cub, occ_stage;   // moni_gpu_seed_occs scratch
    DevBuf pop_qpos, pop_above, pop_slot, pop_head, pop_up, pop_down, pop_parent, pop_len, pop_hq, pop_hoff, pop_hout;   // moni_gpu_populate_seeds scratch
    ImageView view;
    std::mutex mu;
    double t[6] = {0, 0, 0, 0, 0, 0};
    cudaStream_t stream() const { return lanes[0].stream; }
    ~Replica() {
        cudaSetDevice(dev);
        rows.release(); row_pos.release(); crun_row.release(); slots.release(); nodes.release(); pool.release();
        sym.release(); counters.release();
        phi.release(); phi_inv.release(); phi_dir.release(); phi_inv_dir.release();
        occ_ref.release(); occ_len.release(); occ_off.release(); occ_out.release(); occ_flag.release(); occ_cub.release(); occ_stage.release();
        for (DevBuf* b : {&pop_qpos, &pop_above, &pop_slot, &pop_head, &pop_up, &pop_down, &pop_parent, &pop_len, &pop_hq, &pop_hoff, &pop_hout}) b->release();
        for (Lane& L : lanes) {
            L.scratch.release();
            L.stage.release();
            for (auto& e : L.events) if (e) cudaEventDestroy(e);
            if (L.stream) cudaStreamDestroy(L.stream);
        }
    }
};

// One table of the image into a replica: from the host for the first replica, from the first replica's HBM for the others
// (device to device over NVLink / NVSwitch where the GPUs are peers; cudaMemcpyPeerAsync stages through the host otherwise),
// so the host link carries the image once however many GPUs hold it.
template <typename T>
void upload(DevBuf& b, const std::vector<T>& v, cudaStream_t s, const DevBuf* from = nullptr, int from_dev = 0, int to_dev = 0) {
    b.reserve(std::max<size_t>(v.size() * sizeof(T), 16));
    if (v.empty()) return;
    if (from) CK(cudaMemcpyPeerAsync(b.p, to_dev, from->p, from_dev, v.size() * sizeof(T), s));
    else CK(cudaMemcpyAsync(b.p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, s));
}

}  // namespace

struct moni_gpu_index {
    ImageHeader hdr;
    uint64_t device_bytes = 0;
    std::vector<std::unique_ptr<Replica>> reps;
    Options opt;
    double t[6] = {0, 0, 0, 0, 0, 0};
};

struct moni_gpu_batch {
    moni_gpu_index* idx = nullptr;
    Replica* rep = nullptr;
    Scratch s;
    ChunkPlan plan;
    uint64_t n_mems = 0;
    int last = 0;
    bool prepared = false;
};

namespace {

// ---- one chunk on one stream (device pointers already filled) ------------------------------------
ChunkMap chunk_map(const Scratch& S, const ChunkPlan& p) {
    ChunkMap cm;
    cm.offs = S.offs.as<uint64_t>();
    cm.perm = p.ragged() ? S.perm.as<uint32_t>() : nullptr;
    cm.base = p.base;
    cm.n_reads = (uint32_t)p.n;
    cm.seq_bytes = p.bases + 16;
    cm.flag_words = p.flag_words();
    cm.run_entries = p.run_entries();
    return cm;
}

// Ragged chunks: sort the reads by length (longest first) so that the 32 reads of a warp finish together.
void launch_prepare(Scratch& S, const ChunkPlan& p, cudaStream_t st) {
    if (!p.ragged()) return;
    const uint32_t n = (uint32_t)p.n;
    k_read_len<<<(n + 255) / 256, 256, 0, st>>>(S.offs.as<uint64_t>(), n, S.keys_in.as<uint32_t>(), S.vals_in.as<uint32_t>());
    g_launches++;
    int bits = 1;
    while (bits < 32 && (p.max_len >> bits)) ++bits;
    size_t tmp = 0;
    CK(cub::DeviceRadixSort::SortPairsDescending(nullptr, tmp, S.keys_in.as<uint32_t>(), S.keys_out.as<uint32_t>(), S.vals_in.as<uint32_t>(),
                                                 S.perm.as<uint32_t>(), (int)n, 0, bits, st));
    S.cub.reserve(tmp);
    CK(cub::DeviceRadixSort::SortPairsDescending(S.cub.p, tmp, S.keys_in.as<uint32_t>(), S.keys_out.as<uint32_t>(), S.vals_in.as<uint32_t>(),
                                                 S.perm.as<uint32_t>(), (int)n, 0, bits, st));
    g_launches += 3;
    CK(cudaGetLastError());
}

// Blocks of a persistent kernel: at most one resident wave, and few enough that a thread takes several reads of the chunk ("dyn_reads_per_lane",
// default 2).  A read is a chain of dependent steps, so a kernel with as many threads as reads lasts as long as its longest read while the
// threads of the short ones idle (ONT-like reads: longest 24.8 kbp, mean 10 kbp -> 40 % of the thread-steps used); with the reads handed out
// longest first and fewer threads than reads, a thread that finishes early takes another read and the threads end together.  The chains in
// flight that the memory system needs come from the lanes' kernels running side by side.
template <typename K>
unsigned persistent_grid(const Replica& R, K kernel, uint64_t n_reads) {
    int per_sm = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kBlock, 0));
    const uint64_t wave = (uint64_t)std::max(per_sm, 1) * (uint64_t)std::max(R.sms, 1);
    const uint64_t per_block = (uint64_t)kBlock * (uint64_t)std::max(g_dyn_rpt.load(), 1);
    return (unsigned)std::max<uint64_t>(1, std::min<uint64_t>(wave, (n_reads + per_block - 1) / per_block));
}

bool use_dynamic(const Replica& R, const ChunkPlan& p) { return R.dynamic == 1 || (R.dynamic == 2 && p.ragged()); }

void launch_pass1(Replica& R, Scratch& S, const ChunkPlan& p, cudaStream_t st) {
    if (!p.n) return;
    const ChunkMap cm = chunk_map(S, p);
    const bool count = g_count.load();
    Counters* g = count ? R.counters.as<Counters>() : nullptr;
    if (use_dynamic(R, p)) {
        CK(cudaMemsetAsync(S.ctr.p, 0, 16, st));
        if (count)
            k_ms_pointers_dyn<true><<<persistent_grid(R, k_ms_pointers_dyn<true>, p.n), kBlock, 0, st>>>(
                R.view, S.seqs.as<uint8_t>(), cm, S.rptr.as<uint64_t>(), S.flags.as<uint32_t>(), S.rcnt.as<uint32_t>(), S.ctr.as<uint32_t>(), g);
        else
            k_ms_pointers_dyn<false><<<persistent_grid(R, k_ms_pointers_dyn<false>, p.n), kBlock, 0, st>>>(
                R.view, S.seqs.as<uint8_t>(), cm, S.rptr.as<uint64_t>(), S.flags.as<uint32_t>(), S.rcnt.as<uint32_t>(), S.ctr.as<uint32_t>(), nullptr);
    } else {
        const dim3 grid((unsigned)((p.n + kBlock - 1) / kBlock));
        if (count)
            k_ms_pointers<true><<<grid, kBlock, 0, st>>>(R.view, S.seqs.as<uint8_t>(), cm, S.rptr.as<uint64_t>(), S.flags.as<uint32_t>(),
                                                          S.rcnt.as<uint32_t>(), g);
        else
            k_ms_pointers<false><<<grid, kBlock, 0, st>>>(R.view, S.seqs.as<uint8_t>(), cm, S.rptr.as<uint64_t>(), S.flags.as<uint32_t>(),
                                                           S.rcnt.as<uint32_t>(), nullptr);
    }
    g_launches++;
    CK(cudaGetLastError());
}

template <int MODE>
void launch_pass2_mode(Replica& R, Scratch& S, const ChunkPlan& p, cudaStream_t st, uint32_t min_len) {
    const ChunkMap cm = chunk_map(S, p);
    Counters* g = g_count.load() ? R.counters.as<Counters>() : nullptr;
    RunBufs B;
    B.seqs = S.seqs.as<uint8_t>();
    B.flags = S.flags.as<uint32_t>();
    B.rptr = S.rptr.as<uint64_t>();
    B.rrec = MODE == 0 ? S.rrec.as<uint64_t>() : nullptr;
    B.rec = MODE ? S.rec.as<uint64_t>() : nullptr;
    B.rec_ref = MODE == 2 ? S.rec_ref.as<uint64_t>() : nullptr;
    B.seq_bytes = p.bases + 16;
    B.flag_words = p.flag_words();
    B.run_entries = p.run_entries();
    B.rec_entries = p.bases + 2;
    const uint32_t* rc = S.rcnt.as<uint32_t>();
    uint32_t* cnt = MODE ? S.counts.as<uint32_t>() : nullptr;
    if (use_dynamic(R, p)) {
        uint32_t* next = S.ctr.as<uint32_t>() + 1;        // zeroed with pass 1's counter
        if (g) k_ms_lengths_dyn<MODE, true><<<persistent_grid(R, k_ms_lengths_dyn<MODE, true>, p.n), kBlock, 0, st>>>(R.view, B, cm, rc, cnt, next, g, min_len);
        else k_ms_lengths_dyn<MODE, false><<<persistent_grid(R, k_ms_lengths_dyn<MODE, false>, p.n), kBlock, 0, st>>>(R.view, B, cm, rc, cnt, next, nullptr, min_len);
    } else {
        const dim3 grid((unsigned)((p.n + kBlock - 1) / kBlock));
        if (g) k_ms_lengths<MODE, true><<<grid, kBlock, 0, st>>>(R.view, B, cm, rc, cnt, g, min_len);
        else k_ms_lengths<MODE, false><<<grid, kBlock, 0, st>>>(R.view, B, cm, rc, cnt, nullptr, min_len);
    }
}

void launch_pass2(Replica& R, Scratch& S, const ChunkPlan& p, int mode, cudaStream_t st, uint32_t min_len = 0) {
    if (!p.n) return;
    if (mode == kModeSeeds) launch_pass2_mode<2>(R, S, p, st, min_len);
    else if (mode == kModeMems) launch_pass2_mode<1>(R, S, p, st, 0);
    else launch_pass2_mode<0>(R, S, p, st, 0);
    g_launches++;
    CK(cudaGetLastError());
}

// ms: pointers and lengths from the restart runs to the caller's layout
void launch_expand(Scratch& S, const ChunkPlan& p, cudaStream_t st, bool packed = false) {
    if (!p.n) return;
    const dim3 grid((unsigned)((p.n * 32 + kBlock - 1) / kBlock));
    if (packed)
        k_ms_expand<true><<<grid, kBlock, 0, st>>>(chunk_map(S, p), S.rptr.as<uint64_t>(), S.rrec.as<uint64_t>(), S.flags.as<uint32_t>(),
                                                   S.rcnt.as<uint32_t>(), nullptr, nullptr, S.out_ptr.as<uint32_t>(), S.out_hi.as<uint8_t>(),
                                                   S.out_len.as<uint16_t>());
    else
        k_ms_expand<false><<<grid, kBlock, 0, st>>>(chunk_map(S, p), S.rptr.as<uint64_t>(), S.rrec.as<uint64_t>(), S.flags.as<uint32_t>(),
                                                    S.rcnt.as<uint32_t>(), S.out_ptr.as<uint64_t>(), S.out_len.as<uint64_t>(), nullptr, nullptr, nullptr);
    g_launches++;
    CK(cudaGetLastError());
}

// counts -> exclusive offsets (n + 1 entries) -> gather.  mem_pos/mem_len hold `bases` entries.
// rec_bytes: 8 = two u64 arrays, 4 = two u32 arrays, 2 = one u32 array of packed (offset:16, length:16) records, 1 = one u16 array of
// packed (offset:8, length:8) records, 12 = seeds (u32 query offset, u32 length, u64 text position),
// 16 = runs (u32 query offset, u32 length, u64 pointer), 9 = packed runs (u32 (offset:16, length:16), u32 + u8 pointer)
void launch_compact(Scratch& S, const ChunkPlan& p, int rec_bytes, cudaStream_t st) {
    if (!p.n) return;
    const uint32_t n = (uint32_t)p.n;
    const bool runs = rec_bytes == 16 || rec_bytes == 9;
    uint64_t* wide = S.moff.as<uint64_t>() + (n + 1);          // second half of moff: widened counts
    k_widen_counts<<<(n + 1 + 255) / 256, 256, 0, st>>>(runs ? S.rcnt.as<uint32_t>() : S.counts.as<uint32_t>(), wide, n);
    g_launches++;
    size_t tmp = 0;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp, wide, S.moff.as<uint64_t>(), (int)(n + 1), st));
    S.cub.reserve(tmp);
    CK(cub::DeviceScan::ExclusiveSum(S.cub.p, tmp, wide, S.moff.as<uint64_t>(), (int)(n + 1), st));
    g_launches += 2;
    const uint64_t threads = (uint64_t)n * 32;
    const unsigned blocks = (unsigned)((threads + kBlock - 1) / kBlock);
    if (rec_bytes == 16)
        k_runs_gather<true><<<blocks, kBlock, 0, st>>>(S.rptr.as<uint64_t>(), S.rrec.as<uint64_t>(), S.offs.as<uint64_t>(), p.base, n, S.moff.as<uint64_t>(),
                                                       S.mem_pos.as<uint32_t>(), S.mem_len.as<uint32_t>(), S.mem_ref.as<uint64_t>(), nullptr);
    else if (rec_bytes == 9)
        k_runs_gather<false><<<blocks, kBlock, 0, st>>>(S.rptr.as<uint64_t>(), S.rrec.as<uint64_t>(), S.offs.as<uint64_t>(), p.base, n, S.moff.as<uint64_t>(),
                                                        S.mem_pos.as<uint32_t>(), S.mem_len.as<uint32_t>(), nullptr, S.out_hi.as<uint8_t>());
    else if (rec_bytes == 12)
        k_seed_gather<<<blocks, kBlock, 0, st>>>(S.rec.as<uint64_t>(), S.rec_ref.as<uint64_t>(), S.offs.as<uint64_t>(), p.base, n, S.moff.as<uint64_t>(),
                                                 S.mem_pos.as<uint32_t>(), S.mem_len.as<uint32_t>(), S.mem_ref.as<uint64_t>());
    else if (rec_bytes == 2)
        k_mem_gather_packed<uint32_t, 16><<<blocks, kBlock, 0, st>>>(S.rec.as<uint64_t>(), S.offs.as<uint64_t>(), p.base, n, S.moff.as<uint64_t>(),
                                                                     S.mem_pos.as<uint32_t>());
    else if (rec_bytes == 1)
        k_mem_gather_packed<uint16_t, 8><<<blocks, kBlock, 0, st>>>(S.rec.as<uint64_t>(), S.offs.as<uint64_t>(), p.base, n, S.moff.as<uint64_t>(),
                                                                    S.mem_pos.as<uint16_t>());
    else if (rec_bytes == 4)
        k_mem_gather<uint32_t><<<blocks, kBlock, 0, st>>>(S.rec.as<uint64_t>(), S.offs.as<uint64_t>(), p.base, n, S.moff.as<uint64_t>(),
                                                          S.mem_pos.as<uint32_t>(), S.mem_len.as<uint32_t>());
    else
        k_mem_gather<uint64_t><<<blocks, kBlock, 0, st>>>(S.rec.as<uint64_t>(), S.offs.as<uint64_t>(), p.base, n, S.moff.as<uint64_t>(),
                                                          S.mem_pos.as<uint64_t>(), S.mem_len.as<uint64_t>());
    g_launches++;
    CK(cudaGetLastError());
}

float ev_ms(cudaEvent_t a, cudaEvent_t b) {
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, a, b));
    return ms;
}

struct BadBatch : std::runtime_error {
    using std::runtime_error::runtime_error;
};

// Next chunk of reads [a, r1).  A chunk is closed once it holds chunk_bases bases AND min_reads reads (a kernel
// needs ~10^5 independent reads in flight whatever their length), or when it reaches hard_cap bases (device scratch
// is ~24 bytes per base per lane; chunk-relative offsets are 32-bit).  Always at least one read.  Validates the
// offsets it walks over (the only pass the host makes over them).
constexpr uint64_t kChunkHardCap = 1ull << 30;

ChunkPlan plan_chunk(const uint64_t* offsets, uint64_t a, uint64_t r1, uint64_t chunk_bases, uint64_t min_reads,
                     uint64_t hard_cap = kChunkHardCap) {
    ChunkPlan p;
    p.first = a;
    p.base = offsets[a];
    uint64_t b = a, lo = ~0ull, hi = 0;
    while (b < r1 && b - a < (1ull << 28)) {          // run_base / flag_base stay below 2^32
        const uint64_t s = offsets[b], e = offsets[b + 1];
        if (e < s) throw BadBatch("offsets must be non-decreasing");
        if (e - s >= 0x7FFFFFFFull) throw BadBatch("reads of 2^31 bases or more are not supported");
        if (b != a) {
            const uint64_t with = e - p.base;
            if (with > hard_cap || (with > chunk_bases && b - a >= min_reads)) break;
        }
        lo = std::min(lo, e - s);
        hi = std::max(hi, e - s);
        ++b;
    }
    p.n = b - a;
    p.bases = offsets[b] - p.base;
    p.max_len = hi;
    p.uniform_len = (lo == hi) ? (uint32_t)hi : 0;
    return p;
}

__global__ void k_shift(uint64_t* __restrict__ a, uint64_t n, uint64_t by) {
    const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (i < n) a[i] += by;
}

// One host batch, cut into chunks that the GPUs of an index take in turn (in input order, whoever has a free lane next).  For the
// record-producing calls a chunk's records belong at the position given by the record counts of all chunks before it: every
// chunk publishes its count as soon as its kernels are done and waits only for the counts of earlier chunks -- which started
// earlier, on whatever GPU -- so results stream home in input order while later chunks are still being computed, with one GPU
// or eight, and no host-side stitching.
struct ChunkFeed {
    std::mutex mu;
    std::condition_variable cv;
    const uint64_t* offsets = nullptr;
    uint64_t next_read = 0, n_reads = 0;
    uint64_t chunk_bases = 0, min_reads = 0, hard_cap = kChunkHardCap;
    int workers = 1, arrived = 0;
    std::vector<int64_t> count;          // records of chunk c; -1 = not yet known
    std::vector<uint64_t> prefix{0};     // prefix[c] = records of chunks 0..c-1, valid for c <= known
    size_t known = 0;
    bool failed = false;
    int lanes = 4;                       // chunks in flight per GPU in this call
    bool ramp = false;                   // shorter first chunks (see next())
    bool guided = true;                  // shrinking last chunks (see next())
    bool widen = false;                  // moni_gpu_mems: records cross the link packed, host threads widen them (host_widen.hpp)

    // every worker reports how many bases of scratch its GPU can take per lane; chunks are planned once all have (smallest wins)
    void arrive(uint64_t fit) {
        std::unique_lock<std::mutex> lock(mu);
        hard_cap = std::max<uint64_t>(1u << 16, std::min(hard_cap, fit));
        ++arrived;
        cv.notify_all();
        cv.wait(lock, [&] { return failed || arrived >= workers; });
        if (failed) throw std::runtime_error("another GPU's part of the batch failed");
    }
    // next chunk in input order; false when the batch is exhausted
    bool next(ChunkPlan& p, size_t& c) {
        std::lock_guard<std::mutex> lock(mu);
        if (failed) throw std::runtime_error("another GPU's part of the batch failed");
        if (next_read >= n_reads) return false;
        // ramp: the first chunk of every lane is cut shorter (1/L, 2/L, ... of the nominal size), so that the lanes -- which all start at
        // once -- do not copy at the same time and compute at the same time for the rest of the batch, but spread over the period
        uint64_t cb = std::min(chunk_bases, hard_cap), mr = min_reads, hc = hard_cap;
        // guided: where the read-count rule decides the chunk size (long reads), chunks shrink with what is left of the batch, so that
        // the batch does not end with one or two large chunks on a GPU whose other lanes have run dry
        if (guided) mr = std::max<uint64_t>(std::min<uint64_t>(1024, min_reads), std::min(min_reads, (n_reads - next_read) / (2 * (uint64_t)std::max(lanes, 1) * (uint64_t)std::max(workers, 1))));
        const uint64_t c_now = count.size();
        if (ramp && lanes > 1 && c_now < (uint64_t)lanes) {
            const uint64_t num = c_now + 1, den = (uint64_t)lanes;
            cb = std::max<uint64_t>(1, cb * num / den);
            mr = std::max<uint64_t>(1, mr * num / den);
            hc = std::max<uint64_t>(1u << 16, hc * num / den);
        }
        p = plan_chunk(offsets, next_read, n_reads, cb, mr, hc);
        next_read += p.n;
        c = count.size();
        count.push_back(-1);
        prefix.push_back(0);
        return true;
    }
    void publish(size_t c, uint64_t k) {
        {
            std::lock_guard<std::mutex> lock(mu);
            count[c] = (int64_t)k;
            while (known < count.size() && count[known] >= 0) {
                prefix[known + 1] = prefix[known] + (uint64_t)count[known];
                ++known;
            }
        }
        cv.notify_all();
    }
    uint64_t wait_prefix(size_t c) {
        std::unique_lock<std::mutex> lock(mu);
        cv.wait(lock, [&] { return failed || known >= c; });
        if (failed) throw std::runtime_error("another GPU's part of the batch failed");
        return prefix[c];
    }
    void fail() {
        { std::lock_guard<std::mutex> lock(mu); failed = true; }
        cv.notify_all();
    }
    uint64_t total() {
        std::lock_guard<std::mutex> lock(mu);
        return prefix[known];
    }
};

// Where a mems / seeds call wants its records: up to three host columns of `width` bytes per record.
struct RecOut {
    uint64_t* off = nullptr;                     // per-read record offsets
    void* col[3] = {nullptr, nullptr, nullptr};
    int width[3] = {0, 0, 0};
    int ncol = 0;
    uint64_t cap = 0;                            // records the columns can hold
    int fmt = 8;                                 // launch_compact format
    uint32_t min_len = 0;                        // seeds
};

// Size the process-wide widening pool as option "widen_threads" asks (auto: from the CPUs this process may run on).
int widen_pool_threads() {
    moni::WidenPool& P = moni::WidenPool::get();
    if (!g_widen_ready.load()) {
        std::lock_guard<std::mutex> lock(g_opt_mu);
        if (!g_widen_ready.load()) {
            const int want = g_opt.widen_threads < 0 ? moni::WidenPool::default_threads() : g_opt.widen_threads;
            if (want != P.threads()) P.resize(want);
            g_widen_ready.store(true);
        }
    }
    return P.threads();
}

// One per GPU worker of a moni_gpu_mems call: takes (event, staged packed records, destination) jobs in chunk order, waits for the
// chunk's device->host copy and has the pool widen the records into the caller's u64 columns.  A lane's staging buffer is free again
// when its job is done (wait_lane).
class Widener {
  public:
    struct Job {
        cudaEvent_t ev;
        int cfmt;
        const void *src, *src2, *src3;      // cfmt 7 (ms): pointer low words, pointer bits 32..39, u16 lengths
        uint64_t k;
        uint64_t *dpos, *dlen;
        size_t lane;
    };
    Widener(int dev, size_t lanes) : dev_(dev), pending_(lanes, 0) {}
    ~Widener() { finish(); }
    void start() { if (!th_.joinable()) th_ = std::thread([this] { loop(); }); }
    void push(const Job& j) {
        { std::lock_guard<std::mutex> lock(mu_); jobs_.push_back(j); ++pending_[j.lane]; }
        cv_.notify_all();
    }
    void wait_lane(size_t lane) {
        std::unique_lock<std::mutex> lock(mu_);
        cv_.wait(lock, [&] { return pending_[lane] == 0; });
    }
    // all jobs done, thread joined; returns the first CUDA error text seen (empty if none)
    std::string finish() {
        if (th_.joinable()) {
            { std::lock_guard<std::mutex> lock(mu_); stop_ = true; }
            cv_.notify_all();
            th_.join();
        }
        return err_;
    }
    double busy_ms = 0;      // time the pool spent widening
    uint64_t bytes = 0;      // bytes written to the caller's columns

  private:
    void loop() {
        cudaSetDevice(dev_);
        for (;;) {
            Job j;
            {
                std::unique_lock<std::mutex> lock(mu_);
                cv_.wait(lock, [&] { return stop_ || !jobs_.empty(); });
                if (jobs_.empty()) return;
                j = jobs_.front();
                jobs_.erase(jobs_.begin());
            }
            const cudaError_t e = cudaEventSynchronize(j.ev);
            if (e != cudaSuccess) {
                if (err_.empty()) err_ = std::string("cudaEventSynchronize (widening): ") + cudaGetErrorString(e);
            } else if (err_.empty()) {
                const double t0 = wall_s();
                if (j.cfmt == 7) moni::WidenPool::get().widen_ms((const uint32_t*)j.src, (const uint8_t*)j.src2, (const uint16_t*)j.src3, j.k, j.dpos, j.dlen);
                else moni::WidenPool::get().widen(j.cfmt, j.src, j.src2, j.k, j.dpos, j.dlen);
                busy_ms += (wall_s() - t0) * 1e3;
                bytes += j.k * 16;
            }
            { std::lock_guard<std::mutex> lock(mu_); --pending_[j.lane]; }
            cv_.notify_all();
        }
    }
    int dev_;
    std::mutex mu_;
    std::condition_variable cv_;
    std::vector<Job> jobs_;
    std::vector<int> pending_;
    bool stop_ = false;
    std::string err_;
    std::thread th_;
};

void collect_times(moni_gpu_index* idx) {
    for (int k = 0; k < 6; ++k) {
        idx->t[k] = 0;
        for (auto& r : idx->reps) idx->t[k] = std::max(idx->t[k], r->t[k]);
    }
}

// One GPU's share of a host batch: takes chunks from the feed while it has a free lane (stream + scratch each), so the copies of
// one chunk overlap the kernels of the others.  mode: kModeMs, kModeMems, kModeSeeds, kModeRuns.  All but ms: records go to
// out.col[] (contiguous in input order, capacity out.cap) and per-read record offsets to out.off[0 .. n_reads].
void run_worker(moni_gpu_index* idx, Replica& R, int mode, const char* seqs, const uint64_t* offsets, uint64_t n_reads, ChunkFeed& feed,
                uint64_t* pointers, uint64_t* lengths, const RecOut& out) {
    const bool mems = mode != kModeMs;
    uint64_t* const rec_off = out.off;
    const uint64_t rec_cap = out.cap;
    struct InFlight { ChunkPlan p; size_t c, lane, ev_at; int cfmt; };
    const bool widen = feed.widen && mems && out.fmt == 8;
    // moni_gpu_ms (u64 vectors): 7 bytes per base cross the link (the moni_gpu_ms_packed form) and host threads widen them, chunk by chunk
    const bool widen_ms = feed.widen && mode == kModeMs && out.fmt != 7 && idx->hdr.n <= (1ull << 40) - (1ull << 31);
    Widener W(R.dev, R.lanes.size());
    if (widen || widen_ms) W.start();
    std::vector<InFlight> done;
    std::vector<InFlight> q;                 // chunks enqueued and not yet finished, oldest first (at most one per lane)
    const size_t kLanes_ = std::min<size_t>(R.lanes.size(), std::max(feed.lanes, 1));
    size_t issued = 0;

    auto enqueue = [&](const ChunkPlan& p, size_t c) {
        Lane& L = R.lanes[issued % kLanes_];
        Scratch& S = L.scratch;
        // format the chunk's records take on the device: the caller's, or -- when host threads widen them -- the narrowest that holds them
        const int cfmt = !widen ? out.fmt : (p.max_len <= 0xFFull ? 1 : p.max_len <= 0xFFFFull ? 2 : 4);
        S.reserve(p, mode, cfmt);
        InFlight f{p, c, issued % kLanes_, L.used, cfmt};
        ++issued;
        L.take();
        cudaEvent_t* e = L.events.data() + f.ev_at;
        CK(cudaEventRecord(e[0], L.stream));
        if (p.bases) CK(cudaMemcpyAsync(S.seqs.p, seqs + p.base, p.bases, cudaMemcpyHostToDevice, L.stream));
        if (p.uniform_len) {
            k_uniform_offsets<<<(unsigned)((p.n + 1 + 255) / 256), 256, 0, L.stream>>>(S.offs.as<uint64_t>(), p.base, p.uniform_len, (uint32_t)(p.n + 1));
            g_launches++;
        } else {
            CK(cudaMemcpyAsync(S.offs.p, offsets + p.first, (p.n + 1) * 8, cudaMemcpyHostToDevice, L.stream));
        }
        CK(cudaEventRecord(e[1], L.stream));
        launch_prepare(S, p, L.stream);
        launch_pass1(R, S, p, L.stream);
        CK(cudaEventRecord(e[2], L.stream));
        if (out.fmt == 2 && mems && p.max_len > 0xFFFFull) throw BadBatch("packed MEM records need reads shorter than 65536 bases");
        if (out.fmt == 1 && mems && p.max_len > 0xFFull) throw BadBatch("16-bit MEM records need reads shorter than 256 bases");
        if (out.fmt == 9 && p.max_len > 0xFFFFull) throw BadBatch("packed restart runs need reads shorter than 65536 bases");
        launch_pass2(R, S, p, mode, L.stream, out.min_len);
        CK(cudaEventRecord(e[3], L.stream));
        if (mems) {
            launch_compact(S, p, cfmt, L.stream);
            CK(cudaEventRecord(e[4], L.stream));
        } else {
            const bool staged = widen_ms && p.max_len <= 0xFFFFull;      // this chunk leaves packed and is widened on the host
            const bool packed = out.fmt == 7 || staged;      // moni_gpu_ms_packed: col[] = u32 pointer halves, u8 pointer bits 32..39, u16 lengths
            if (out.fmt == 7 && p.max_len > 0xFFFFull) throw BadBatch("packed matching statistics need reads shorter than 65536 bases");
            launch_expand(S, p, L.stream, packed);
            CK(cudaEventRecord(e[4], L.stream));
            if (p.bases && staged) {
                W.wait_lane(f.lane);                         // the lane's previous chunk has left the staging buffer
                const size_t a4 = (p.bases * 4 + 63) & ~(size_t)63, a1 = (p.bases + 63) & ~(size_t)63, a2 = (p.bases * 2 + 63) & ~(size_t)63;
                L.stage.reserve(a4 + a1 + a2);
                char* const st0 = (char*)L.stage.p;
                CK(cudaMemcpyAsync(st0, S.out_ptr.p, p.bases * 4, cudaMemcpyDeviceToHost, L.stream));
                CK(cudaMemcpyAsync(st0 + a4, S.out_hi.p, p.bases, cudaMemcpyDeviceToHost, L.stream));
                CK(cudaMemcpyAsync(st0 + a4 + a1, S.out_len.p, p.bases * 2, cudaMemcpyDeviceToHost, L.stream));
                CK(cudaEventRecord(e[5], L.stream));
                W.push({e[5], 7, st0, st0 + a4, st0 + a4 + a1, p.bases, pointers + p.base, lengths + p.base, f.lane});
                q.push_back(f);
                return;
            }
            if (p.bases && packed) {
                DevBuf* const src[3] = {&S.out_ptr, &S.out_hi, &S.out_len};
                for (int k = 0; k < 3; ++k)
                    CK(cudaMemcpyAsync((char*)out.col[k] + p.base * out.width[k], src[k]->p, p.bases * out.width[k], cudaMemcpyDeviceToHost, L.stream));
            } else if (p.bases) {
                CK(cudaMemcpyAsync(pointers + p.base, S.out_ptr.p, p.bases * 8, cudaMemcpyDeviceToHost, L.stream));
                CK(cudaMemcpyAsync(lengths + p.base, S.out_len.p, p.bases * 8, cudaMemcpyDeviceToHost, L.stream));
            }
            CK(cudaEventRecord(e[5], L.stream));
        }
        q.push_back(f);
    };
    // oldest chunk in flight: ms -- wait until its results are home (the lane is free again); records -- learn its record count,
    // publish it, wait for the counts of all earlier chunks, then send offsets (shifted to batch-relative) and records home
    auto finish = [&](const InFlight& f) {
        const ChunkPlan& p = f.p;
        Lane& L = R.lanes[f.lane];
        Scratch& S = L.scratch;
        cudaEvent_t* e = L.events.data() + f.ev_at;
        if (!mems) {
            CK(cudaEventSynchronize(e[5]));
            return;
        }
        uint64_t k = 0;
        CK(cudaMemcpyAsync(&k, S.moff.as<uint64_t>() + p.n, 8, cudaMemcpyDeviceToHost, L.stream));
        CK(cudaStreamSynchronize(L.stream));
        feed.publish(f.c, k);
        const uint64_t before = feed.wait_prefix(f.c);
        if (before) {
            k_shift<<<(unsigned)((p.n + 1 + 255) / 256), 256, 0, L.stream>>>(S.moff.as<uint64_t>(), p.n + 1, before);
            g_launches++;
        }
        const bool last = p.first + p.n == n_reads;
        DevBuf* const chunk_col[3] = {&S.mem_pos, &S.mem_len, out.fmt == 9 ? &S.out_hi : &S.mem_ref};
        CK(cudaMemcpyAsync(rec_off + p.first, S.moff.p, (p.n + (last ? 1 : 0)) * 8, cudaMemcpyDeviceToHost, L.stream));
        if (widen) {
            if (k && before + k <= rec_cap) {
                W.wait_lane(f.lane);                      // the lane's previous chunk has left the staging buffer
                const size_t w = f.cfmt == 1 ? 2 : 4, col_bytes = (k * w + 63) & ~(size_t)63;
                L.stage.reserve(f.cfmt == 4 ? 2 * col_bytes : col_bytes);
                CK(cudaMemcpyAsync(L.stage.p, S.mem_pos.p, k * w, cudaMemcpyDeviceToHost, L.stream));
                if (f.cfmt == 4) CK(cudaMemcpyAsync((char*)L.stage.p + col_bytes, S.mem_len.p, k * w, cudaMemcpyDeviceToHost, L.stream));
                CK(cudaEventRecord(e[5], L.stream));
                W.push({e[5], f.cfmt, L.stage.p, f.cfmt == 4 ? (char*)L.stage.p + col_bytes : nullptr, nullptr, k, (uint64_t*)out.col[0] + before,
                        (uint64_t*)out.col[1] + before, f.lane});
            } else {
                CK(cudaEventRecord(e[5], L.stream));
            }
            return;
        }
        for (int col = 0; col < out.ncol && k && before + k <= rec_cap; ++col)
            CK(cudaMemcpyAsync((char*)out.col[col] + before * out.width[col], chunk_col[col]->p, k * out.width[col], cudaMemcpyDeviceToHost, L.stream));
        CK(cudaEventRecord(e[5], L.stream));
    };

    try {
        CK(cudaSetDevice(R.dev));
        for (double& x : R.t) x = 0;
        for (Lane& L : R.lanes) L.used = 0;
        // A chunk may not need more scratch than a lane's share of the HBM that is free (or already held by the lanes): the index
        // image, other replicas on the same device and other processes have taken theirs.  A single read larger than that share
        // still forms a chunk of its own and then fails with MONI_GPU_ENOMEM and the size needed.
        {
            size_t free_b = 0, total_b = 0, held = 0;
            CK(cudaMemGetInfo(&free_b, &total_b));
            for (Lane& L : R.lanes) held += L.scratch.held();
            const double share = 0.85 * (double)(free_b + held) / (double)kLanes_;
            feed.arrive((uint64_t)(share / Scratch::bytes_per_base(mode, widen ? 4 : out.fmt)));
        }
        for (;;) {
            ChunkPlan p;
            size_t c = 0;
            while (q.size() < kLanes_ && feed.next(p, c)) enqueue(p, c);
            if (q.empty()) break;
            finish(q.front());
            done.push_back(q.front());
            q.erase(q.begin());
        }
        for (Lane& L : R.lanes) CK(cudaStreamSynchronize(L.stream));
        const std::string werr = W.finish();
        if (!werr.empty()) throw CudaError(werr);
        if ((widen || widen_ms) && getenv("MONI_GPU_TRACE"))
            fprintf(stderr, "[moni-gpu trace] gpu %d widened %.1f MB of records on %d host threads in %.2f ms (%.1f GB/s written)\n", R.dev,
                    W.bytes / 1e6, moni::WidenPool::get().threads(), W.busy_ms, W.busy_ms > 0 ? W.bytes / W.busy_ms / 1e6 : 0.0);
    } catch (...) {
        // a bad offset found mid-batch (or a CUDA failure): the other GPUs stop, and nothing may still be writing the caller's
        // buffers when we return
        feed.fail();
        for (Lane& L : R.lanes) cudaStreamSynchronize(L.stream);
        W.finish();
        throw;
    }
    if (!done.empty() && getenv("MONI_GPU_TRACE")) {      // per-chunk timeline (ms since the first event): where the pipeline idles
        cudaEvent_t* e0 = R.lanes[done[0].lane].events.data() + done[0].ev_at;
        for (const InFlight& f : done) {
            cudaEvent_t* e = R.lanes[f.lane].events.data() + f.ev_at;
            fprintf(stderr, "[moni-gpu trace] gpu %d chunk %zu lane %zu reads %llu:", R.dev, f.c, f.lane, (unsigned long long)f.p.n);
            for (int k = 0; k < kEv; ++k) fprintf(stderr, " %.3f", ev_ms(e0[0], e[k]));
            fprintf(stderr, "   (start h2d pass1 pass2 compact d2h)\n");
        }
    }
    for (const InFlight& f : done) {
        cudaEvent_t* e = R.lanes[f.lane].events.data() + f.ev_at;
        for (int k = 0; k < 5; ++k) R.t[k] += ev_ms(e[k], e[k + 1]);
        cudaEvent_t* e0 = R.lanes[done[0].lane].events.data() + done[0].ev_at;
        R.t[5] = std::max<double>(R.t[5], ev_ms(e0[0], e[5]));
    }
}

// Runs one batch over all replicas of the index; returns the number of records (0 for ms).
uint64_t run_batch(moni_gpu_index* idx, int mode, const char* seqs, const uint64_t* offsets, uint64_t n_reads, uint64_t* pointers,
                   uint64_t* lengths, const RecOut& out) {
    const int G = (int)idx->reps.size();
    // The workers of one call wait for each other (chunk counts), so a call owns ALL replicas while it runs: the locks are taken
    // here, in replica order, not by the workers -- two concurrent calls on one index can then never hold one replica each.
    std::vector<std::unique_lock<std::mutex>> locks;
    for (auto& r : idx->reps) locks.emplace_back(r->mu);
    ChunkFeed feed;
    feed.offsets = offsets;
    feed.n_reads = n_reads;
    feed.chunk_bases = idx->opt.chunk_bases;
    feed.workers = G;
    // measured on the bench boxes: a widening thread writes 9-15 GB/s; two of them beat 16-byte records on a shared link, moni_gpu_ms
    // (4.8 GB written per 2 M reads against 2.7 GB less on the link) needs six
    const int wt = ((mode == kModeMems && out.fmt == 8) || (mode == kModeMs && out.fmt != 7)) ? widen_pool_threads() : 0;
    feed.widen = mode == kModeMems ? wt >= 2 : wt >= 6;
    // A chunk is kept open until it holds chunk_min_reads reads so that its kernels fill the GPU; with long reads that rule alone
    // would cut a 1 G-base batch into two chunks and leave lanes (and the copy/compute overlap) unused, so the minimum shrinks
    // until the batch makes at least two chunks per lane.
    const uint64_t mean_len = n_reads ? (offsets[n_reads] - offsets[0]) / n_reads : 0;
    feed.lanes = idx->opt.lanes ? (int)idx->opt.lanes : 4;      // 6 and 8 lanes measured the same on 1 M ONT-like reads (profiles/r02_r_*)
    feed.guided = g_guided.load() != 0;
    feed.ramp = g_ramp.load() == 1 || (g_ramp.load() == 2 && mean_len >= 4096 && G == 1);
    feed.min_reads = std::min<uint64_t>(idx->opt.chunk_min_reads, std::max<uint64_t>(1024, n_reads / (2 * (uint64_t)feed.lanes * (uint64_t)G)));
    if (G == 1) {
        run_worker(idx, *idx->reps[0], mode, seqs, offsets, n_reads, feed, pointers, lengths, out);
    } else {
        std::vector<std::exception_ptr> errs(G);
        std::vector<std::thread> th;
        for (int g = 0; g < G; ++g)
            th.emplace_back([&, g]() {
                try {
                    run_worker(idx, *idx->reps[g], mode, seqs, offsets, n_reads, feed, pointers, lengths, out);
                } catch (...) {
                    errs[g] = std::current_exception();
                }
            });
        for (auto& t : th) t.join();
        // the failure that started it, not the "another GPU failed" of those that were stopped by it
        std::exception_ptr first, other;
        for (int g = 0; g < G; ++g) {
            if (!errs[g]) continue;
            try { std::rethrow_exception(errs[g]); }
            catch (const CudaError&) { if (!first) first = errs[g]; }
            catch (const DeviceOom&) { if (!first) first = errs[g]; }
            catch (const BadBatch&) { if (!first) first = errs[g]; }
            catch (...) { if (!other) other = errs[g]; }
        }
        if (first) std::rethrow_exception(first);
        if (other) std::rethrow_exception(other);
    }
    collect_times(idx);
    return feed.total();
}

template <typename F>
int guarded(F&& f) {
    try {
        return f();
    } catch (const CudaError& e) {
        g_err = e.what();
        return MONI_GPU_ECUDA;
    } catch (const DeviceOom& e) {
        g_err = e.what();
        return MONI_GPU_ENOMEM;
    } catch (const BadBatch& e) {
        g_err = e.what();
        return MONI_GPU_EINVAL;
    } catch (const std::bad_alloc&) {
        g_err = "out of host memory";
        return MONI_GPU_ENOMEM;
    } catch (const std::exception& e) {
        g_err = e.what();
        return MONI_GPU_EIO;
    }
}


}  // namespace

// ================================================================================ C ABI
extern "C" {

const char* moni_gpu_last_error(void) { return g_err.c_str(); }
const char* moni_gpu_version(void) { return "moni-gpu 0.3 (sm_100a)"; }
uint64_t moni_gpu_launch_count(void) { return g_launches.load(); }

int moni_gpu_set_option(const char* key, const char* value) {
    if (!key || !value) { g_err = "null option"; return MONI_GPU_EINVAL; }
    std::lock_guard<std::mutex> lock(g_opt_mu);
    const std::string k(key);
    const long long v = atoll(value);
    if (k == "leaf_len" && v >= 0 && v <= (1ll << 40)) g_opt.leaf_len = (uint64_t)v;
    else if (k == "text_budget" && v >= 0) g_opt.text_budget = (uint64_t)v;
    else if (k == "max_scan" && v >= 2 && v <= 4096) g_opt.max_scan = (uint32_t)v;
    else if (k == "verbose") g_opt.verbose = v != 0;
    else if (k == "grammar" && (std::string(value) == "plain" || std::string(value) == "shaped")) g_opt.grammar = value[0] == 's';
    else if (k == "count") { g_opt.count = v != 0; g_count.store(v != 0); }
    else if (k == "chunk_bases" && v >= 1 && v <= (1ll << 31)) g_opt.chunk_bases = (uint64_t)v;
    else if (k == "chunk_min_reads" && v >= 0) g_opt.chunk_min_reads = (uint64_t)v;
    else if (k == "lanes" && v >= 0 && v <= kMaxLanes) g_opt.lanes = (uint32_t)v;
    else if (k == "dynamic" && v >= 0 && v <= 2) g_opt.dynamic = (int)v;
    else if (k == "dyn_reads_per_lane" && v >= 1 && v <= 64) g_dyn_rpt.store((int)v);
    else if (k == "ramp" && v >= 0 && v <= 2) g_ramp.store((int)v);
    else if (k == "guided" && v >= 0 && v <= 1) g_guided.store((int)v);
    else if (k == "widen_threads" && v >= -1 && v <= 256) { g_opt.widen_threads = (int)v; g_widen_ready.store(false); }
    else if (k == "l2_fetch" && (v == 0 || v == 32 || v == 64 || v == 128)) g_opt.l2_fetch = (uint32_t)v;
    else { g_err = "unknown option or value out of range: " + k; return MONI_GPU_EINVAL; }
    return 0;
}

void* moni_gpu_host_alloc(uint64_t bytes) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) {
        g_err = "cudaHostAlloc failed";
        return nullptr;
    }
    return p;
}
// NUMA node of a CUDA device from sysfs (-1 if unknown): /sys/bus/pci/devices/<domain:bus:dev.fn>/numa_node
static int gpu_numa_node(int dev) {
    char bdf[32] = {0};
    if (cudaDeviceGetPCIBusId(bdf, sizeof(bdf), dev) != cudaSuccess) return -1;
    for (char* c = bdf; *c; ++c) *c = (char)tolower(*c);
    const std::string path = std::string("/sys/bus/pci/devices/") + bdf + "/numa_node";
    FILE* f = fopen(path.c_str(), "r");
    if (!f) return -1;
    int node = -1;
    if (fscanf(f, "%d", &node) != 1) node = -1;
    fclose(f);
    return node;
}

void* moni_gpu_host_alloc_near(int gpu, uint64_t bytes) {
    // Page-locked memory is faulted in by cudaHostAlloc itself, so a "preferred node" memory policy held across the call
    // places the buffer on the GPU's NUMA node: the PCIe copies then do not cross the socket interconnect.
    const int node = gpu_numa_node(gpu);
    bool policy_set = false;
    int old_mode = 0;
    unsigned long old_mask[16] = {0};
    if (node >= 0 && node < 1024) {
        // the caller's own policy is put back afterwards (get_mempolicy with flags 0 returns the thread's policy)
        const bool have_old = syscall(SYS_get_mempolicy, &old_mode, old_mask, (unsigned long)(sizeof(old_mask) * 8), nullptr, 0ul) == 0;
        unsigned long mask[16] = {0};
        mask[node / (8 * sizeof(unsigned long))] = 1ul << (node % (8 * sizeof(unsigned long)));
        policy_set = have_old && syscall(SYS_set_mempolicy, 1 /* MPOL_PREFERRED */, mask, (unsigned long)(sizeof(mask) * 8)) == 0;
    }
    void* p = moni_gpu_host_alloc(bytes);
    if (policy_set) {
        bool any = false;
        for (unsigned long w : old_mask) any = any || w != 0;
        syscall(SYS_set_mempolicy, old_mode, any ? old_mask : nullptr, any ? (unsigned long)(sizeof(old_mask) * 8) : 0ul);
    }
    return p;
}

int moni_gpu_numa_node(int gpu) { return gpu_numa_node(gpu); }

void moni_gpu_host_free(void* p) {
    if (p) cudaFreeHost(p);
}

int moni_gpu_index_load(const char* prefix, const int* gpus, int n_gpus, moni_gpu_index** out) {
    if (!prefix || !out || n_gpus < 1) { g_err = "bad arguments"; return MONI_GPU_EINVAL; }
    *out = nullptr;
    return guarded([&]() -> int {
        std::vector<int> devs(n_gpus);
        for (int g = 0; g < n_gpus; ++g) {
            devs[g] = gpus ? gpus[g] : g;
            if (devs[g] < 0) { g_err = "GPU ordinal out of range"; return MONI_GPU_EINVAL; }
        }
        std::unique_ptr<moni_gpu_index> idx(new moni_gpu_index);
        { std::lock_guard<std::mutex> lock(g_opt_mu); idx->opt = g_opt; }
        if (getenv("MONI_GPU_VERBOSE")) idx->opt.verbose = true;
        const double t_load0 = wall_s();
        // the driver (seconds in a fresh process, more with many GPUs visible) and the CUDA contexts come up while the files
        // are decoded; without a usable device the call still fails with MONI_GPU_ECUDA, whatever else went wrong
        int ndev = 0;
        cudaError_t init_err = cudaSuccess;
        double t_init = 0;
        std::thread ctx_thread([&]() {
            init_err = cudaGetDeviceCount(&ndev);
            if (init_err == cudaSuccess)
                for (int d : devs) { if (d < ndev && cudaSetDevice(d) == cudaSuccess) cudaFree(nullptr); }
            t_init = wall_s() - t_load0;
        });
        struct Joiner { std::thread& t; ~Joiner() { if (t.joinable()) t.join(); } } ctx_join{ctx_thread};
        auto require_devices = [&]() {
            if (ctx_thread.joinable()) ctx_thread.join();
            if (init_err != cudaSuccess || ndev < 1) {
                cudaGetLastError();
                throw CudaError(std::string("no CUDA device: libmoni_gpu has no CPU path") +
                                (init_err != cudaSuccess ? std::string(" (") + cudaGetErrorString(init_err) + ")" : std::string()));
            }
        };
        HostImage img;
        // the two files are decoded and exported side by side (they fill disjoint members of the image)
        std::exception_ptr slp_err;
        std::thread slp_thread([&]() {
            try {
                ExportOptions eo;
                eo.verbose = idx->opt.verbose;
                FlatSlp slp;
                // the grammar `ms` / `mems` load by default is the plain one; `-q` names the shaped one
                // (src/matching_statistics.cpp:178-188).  Both decode to the same rule table; results do not depend on which.
                const std::string plain = std::string(prefix) + ".plain.slp", shaped = std::string(prefix) + ".slp";
                const bool use_shaped = idx->opt.grammar == 1 || (access(plain.c_str(), R_OK) != 0 && access(shaped.c_str(), R_OK) == 0);
                if (use_shaped) read_shaped_slp(shaped, slp);
                else read_plain_slp(plain, slp);
                // B200 has the HBM to keep the text itself next to the move table (6 GB for a 6 Gbp pangenome): then an LCE
                // restart is one random line instead of a root-to-leaf walk.  Larger texts keep the grammar with 128-byte leaves.
                eo.leaf_len = idx->opt.leaf_len ? idx->opt.leaf_len : (slp.text_len <= idx->opt.text_budget ? std::max<uint64_t>(slp.text_len, 1) : 128);
                build_grammar_table(slp, eo, img);
            } catch (...) {
                slp_err = std::current_exception();
            }
        });
        try {
            ExportOptions eo;
            eo.max_scan = idx->opt.max_scan;
            eo.verbose = idx->opt.verbose;
            FlatRlbwt bwt;
            // `ms`/`mems` index first; else the `moni align` index (same query structures + Phi, include/ms/moni.hpp:384-387,
            // include/aligner/moni_lcp.hpp:200-203)
            const char* exts[3] = {".thrbv.ms", ".thrbv.full.ms", ".thrbv.full.lcp.ms"};
            int which = 0;
            for (int k = 0; k < 3; ++k)
                if (access((std::string(prefix) + exts[k]).c_str(), R_OK) == 0) { which = k; break; }
            read_thrbv_ms(std::string(prefix) + exts[which], bwt, which != 0);
            build_move_table(bwt, eo, img);
            build_phi_tables(bwt, img);
        } catch (...) {
            slp_thread.join();
            require_devices();
            throw;
        }
        slp_thread.join();
        require_devices();
        if (slp_err) std::rethrow_exception(slp_err);
        for (int d : devs)
            if (d >= ndev) { g_err = "GPU ordinal out of range"; return MONI_GPU_EINVAL; }
        idx->hdr = img.hdr;
        idx->device_bytes = img.bytes() + sizeof(SymbolTables);
        const double t_load1 = wall_s();
        for (int g = 0; g < n_gpus; ++g) {
            std::unique_ptr<Replica> R(new Replica);
            R->dev = devs[g];
            R->dynamic = idx->opt.dynamic;
            CK(cudaSetDevice(R->dev));
            CK(cudaDeviceGetAttribute(&R->sms, cudaDevAttrMultiProcessorCount, R->dev));
            if (idx->opt.l2_fetch) CK(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, idx->opt.l2_fetch));
            R->lanes.resize(idx->opt.lanes ? idx->opt.lanes : 4);
            for (Lane& L : R->lanes) CK(cudaStreamCreateWithFlags(&L.stream, cudaStreamNonBlocking));
            cudaStream_t st = R->stream();
            const Replica* src = g ? idx->reps[0].get() : nullptr;      // replica 0 is complete (synchronized) by now
            if (src && src->dev != R->dev) {                             // peer access lets the copy go GPU to GPU; failure is fine
                int can = 0;
                if (cudaDeviceCanAccessPeer(&can, R->dev, src->dev) == cudaSuccess && can) {
                    const cudaError_t pe = cudaDeviceEnablePeerAccess(src->dev, 0);
                    if (pe != cudaSuccess) cudaGetLastError();           // already enabled, or refused: the copy still works
                }
            }
#define MONI_UP(member, table) upload(R->member, img.table, st, src ? &src->member : nullptr, src ? src->dev : 0, R->dev)
            MONI_UP(rows, rows);
            MONI_UP(row_pos, row_pos);
            MONI_UP(crun_row, crun_row);
            MONI_UP(slots, slots);
            MONI_UP(nodes, nodes);
            MONI_UP(pool, pool);
            MONI_UP(phi, phi);
            MONI_UP(phi_inv, phi_inv);
            MONI_UP(phi_dir, phi_dir);
            MONI_UP(phi_inv_dir, phi_inv_dir);
#undef MONI_UP
            R->sym.reserve(sizeof(SymbolTables));
            CK(cudaMemcpyAsync(R->sym.p, &img.sym, sizeof(SymbolTables), cudaMemcpyHostToDevice, st));
            R->counters.reserve(sizeof(Counters));
            CK(cudaMemsetAsync(R->counters.p, 0, sizeof(Counters), st));
            CK(cudaStreamSynchronize(st));
            ImageView& v = R->view;
            v.rows = R->rows.as<MoveRow>();
            v.row_pos = R->row_pos.as<uint64_t>();
            v.crun_row = R->crun_row.as<uint32_t>();
            v.slots = R->slots.as<GapSlot>();
            v.nodes = R->nodes.as<SlpNode>();
            v.pool = R->pool.as<uint8_t>();
            v.sym = R->sym.as<SymbolTables>();
            v.phi = R->phi.as<PhiEntry>();
            v.phi_inv = R->phi_inv.as<PhiEntry>();
            v.phi_dir = R->phi_dir.as<uint32_t>();
            v.phi_inv_dir = R->phi_inv_dir.as<uint32_t>();
            v.hdr = img.hdr;
            idx->reps.push_back(std::move(R));
        }
        if (idx->opt.verbose)
            fprintf(stderr, "[moni-gpu] index load: driver + contexts %.2f s beside decode + export %.2f s, upload of %.2f GB to %d GPU(s) %.2f s\n", t_init,
                    t_load1 - t_load0, (double)idx->device_bytes / 1e9, n_gpus, wall_s() - t_load1);
        *out = idx.release();
        return 0;
    });
}

void moni_gpu_index_free(moni_gpu_index* idx) { delete idx; }

int moni_gpu_index_info(const moni_gpu_index* idx, uint64_t* n, uint64_t* r, uint64_t* text_len, uint64_t* device_bytes) {
    if (!idx) { g_err = "null index"; return MONI_GPU_EINVAL; }
    if (n) *n = idx->hdr.n;
    if (r) *r = idx->hdr.R;
    if (text_len) *text_len = idx->hdr.text_len;
    if (device_bytes) *device_bytes = idx->device_bytes;
    return 0;
}

static int ms_impl(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads, uint64_t* pointers, uint64_t* lengths,
                   const RecOut& out) {
    if (!idx || !offsets || (!seqs && n_reads && offsets[n_reads] != offsets[0])) { g_err = "bad arguments"; return MONI_GPU_EINVAL; }
    if (n_reads == 0) return 0;
    if (out.fmt == 7 ? (!out.col[0] || !out.col[1] || !out.col[2]) : (!pointers || !lengths)) { g_err = "null output"; return MONI_GPU_EINVAL; }
    if (out.fmt == 7 && idx->hdr.n > (1ull << 40) - (1ull << 31)) { g_err = "packed pointers need n <= 2^40 - 2^31"; return MONI_GPU_EINVAL; }
    return guarded([&]() -> int {
        run_batch(idx, kModeMs, seqs, offsets, n_reads, pointers, lengths, out);
        return 0;
    });
}

int moni_gpu_ms(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads, uint64_t* pointers, uint64_t* lengths) {
    return ms_impl(idx, seqs, offsets, n_reads, pointers, lengths, RecOut());
}

int moni_gpu_ms_packed(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads, uint32_t* ptr_lo, uint8_t* ptr_hi,
                       uint16_t* lengths) {
    RecOut o;
    o.col[0] = ptr_lo; o.width[0] = 4;
    o.col[1] = ptr_hi; o.width[1] = 1;
    o.col[2] = lengths; o.width[2] = 2;
    o.ncol = 3;
    o.fmt = 7;
    return ms_impl(idx, seqs, offsets, n_reads, nullptr, nullptr, o);
}

static int mems_impl(moni_gpu_index* idx, int mode, const char* seqs, const uint64_t* offsets, uint64_t n_reads, RecOut out) {
    uint64_t* const mem_offsets = out.off;
    if (!idx || !offsets || !mem_offsets) { g_err = "bad arguments"; return MONI_GPU_EINVAL; }
    if (n_reads == 0) { mem_offsets[0] = 0; return 0; }
    if (!seqs && offsets[n_reads] != offsets[0]) { g_err = "null seqs"; return MONI_GPU_EINVAL; }
    for (int q = 0; q < out.ncol; ++q)
        if (!out.col[q] && out.cap) { g_err = "null output"; return MONI_GPU_EINVAL; }
    return guarded([&]() -> int {
        const uint64_t needed = run_batch(idx, mode, seqs, offsets, n_reads, nullptr, nullptr, out);
        if (needed > out.cap) {
            g_err = "mem_capacity too small: " + std::to_string(needed) + " records needed";
            return MONI_GPU_ENOMEM;
        }
        return 0;
    });
}

static RecOut rec_out(uint64_t* off, void* a, int wa, void* b, int wb, void* c, int wc, uint64_t cap, int fmt, uint32_t min_len = 0) {
    RecOut o;
    o.off = off;
    o.col[0] = a; o.width[0] = wa;
    o.col[1] = b; o.width[1] = wb;
    o.col[2] = c; o.width[2] = wc;
    o.ncol = wc ? 3 : (wb ? 2 : 1);
    o.cap = cap;
    o.fmt = fmt;
    o.min_len = min_len;
    return o;
}

int moni_gpu_mems(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads, uint64_t* mem_offsets,
                  uint64_t* mem_pos, uint64_t* mem_len, uint64_t mem_capacity) {
    return mems_impl(idx, kModeMems, seqs, offsets, n_reads, rec_out(mem_offsets, mem_pos, 8, mem_len, 8, nullptr, 0, mem_capacity, 8));
}

int moni_gpu_mems32(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads, uint64_t* mem_offsets,
                    uint32_t* mem_pos, uint32_t* mem_len, uint64_t mem_capacity) {
    return mems_impl(idx, kModeMems, seqs, offsets, n_reads, rec_out(mem_offsets, mem_pos, 4, mem_len, 4, nullptr, 0, mem_capacity, 4));
}

int moni_gpu_mems_packed(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads, uint64_t* mem_offsets,
                         uint32_t* mem_rec, uint64_t mem_capacity) {
    return mems_impl(idx, kModeMems, seqs, offsets, n_reads, rec_out(mem_offsets, mem_rec, 4, nullptr, 0, nullptr, 0, mem_capacity, 2));
}

int moni_gpu_mems_packed16(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads, uint64_t* mem_offsets,
                           uint16_t* mem_rec, uint64_t mem_capacity) {
    return mems_impl(idx, kModeMems, seqs, offsets, n_reads, rec_out(mem_offsets, mem_rec, 2, nullptr, 0, nullptr, 0, mem_capacity, 1));
}

int moni_gpu_ms_runs(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads, uint64_t* run_offsets, uint32_t* run_qpos,
                     uint32_t* run_len, uint64_t* run_ptr, uint64_t run_capacity) {
    return mems_impl(idx, kModeRuns, seqs, offsets, n_reads, rec_out(run_offsets, run_qpos, 4, run_len, 4, run_ptr, 8, run_capacity, 16));
}

int moni_gpu_ms_runs_packed(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads, uint64_t* run_offsets,
                            uint32_t* run_rec, uint32_t* ptr_lo, uint8_t* ptr_hi, uint64_t run_capacity) {
    if (idx && idx->hdr.n > (1ull << 40) - (1ull << 31)) { g_err = "packed pointers need n <= 2^40 - 2^31"; return MONI_GPU_EINVAL; }
    return mems_impl(idx, kModeRuns, seqs, offsets, n_reads, rec_out(run_offsets, run_rec, 4, ptr_lo, 4, ptr_hi, 1, run_capacity, 9));
}

// Host-side expansion of restart runs into the reference's two vectors (pure format conversion: between restarts the pointer grows by
// one and the length shrinks by one, floor 0).  No device work.
int moni_gpu_ms_runs_expand(const uint64_t* offsets, uint64_t n_reads, const uint64_t* run_offsets, const uint32_t* run_qpos,
                            const uint32_t* run_len, const uint64_t* run_ptr, uint64_t* pointers, uint64_t* lengths) {
    if (!offsets || !run_offsets || (n_reads && (!run_qpos || !run_len || !run_ptr || !pointers || !lengths))) { g_err = "bad arguments"; return MONI_GPU_EINVAL; }
    for (uint64_t r = 0; r < n_reads; ++r) {
        const uint64_t o = offsets[r], m = offsets[r + 1] - o;
        uint64_t k = run_offsets[r];
        const uint64_t ke = run_offsets[r + 1];
        if (m && (k == ke || run_qpos[k] != 0)) { g_err = "read " + std::to_string(r) + ": runs must start at query offset 0"; return MONI_GPU_EINVAL; }
        for (; k < ke; ++k) {
            const uint64_t i0 = run_qpos[k], i1 = k + 1 < ke ? run_qpos[k + 1] : m;
            if (i1 <= i0 || i1 > m) { g_err = "read " + std::to_string(r) + ": run offsets must increase inside the read"; return MONI_GPU_EINVAL; }
            const uint64_t p0 = run_ptr[k], l0 = run_len[k];
            for (uint64_t i = i0; i < i1; ++i) {
                pointers[o + i] = p0 + (i - i0);
                lengths[o + i] = l0 > i - i0 ? l0 - (i - i0) : 0;
            }
        }
    }
    return 0;
}

int moni_gpu_seeds(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads, uint64_t min_len,
                   uint64_t* seed_offsets, uint64_t* seed_ref, uint32_t* seed_qpos, uint32_t* seed_len, uint64_t seed_capacity) {
    const uint32_t ml = min_len > 0x7FFFFFFFull ? 0x7FFFFFFFu : (uint32_t)min_len;     // reads are shorter than 2^31
    return mems_impl(idx, kModeSeeds, seqs, offsets, n_reads, rec_out(seed_offsets, seed_qpos, 4, seed_len, 4, seed_ref, 8, seed_capacity, 12, ml));
}

namespace {

// One enumeration job on a replica's first stream: lists for n seeds given on the device.  phase 1 walks (counting, or staging when
// the cap bounds the lists), scans the counts into offsets, sends the offsets home and returns the total; phase 2 produces the
// packed lists in `lists` (reserved by the caller once the total is known to fit its capacity).
struct OccJob {
    const uint64_t *head = nullptr, *up = nullptr, *down = nullptr;
    const uint32_t* len = nullptr;
    uint64_t n = 0, cap = 0, stride = 0;
    DevBuf *off = nullptr, *stage = nullptr, *cub = nullptr, *flag = nullptr, *above = nullptr, *lists = nullptr;
};

uint64_t occ_phase1(Replica& R, cudaStream_t st, OccJob& J, uint64_t* h_off, bool* truncated) {
    J.off->reserve((J.n + 1) * 8);
    J.flag->reserve(16);
    if (J.above) J.above->reserve((J.n + 1) * 4);
    CK(cudaMemsetAsync(J.flag->p, 0, 4, st));
    const unsigned blocks = (unsigned)((J.n + 1 + kBlock - 1) / kBlock);
    // staged single walk when 2 cap + 1 positions per seed fit a bounded share of the free HBM; else count, then fill
    J.stride = 0;
    if (J.cap <= 4096) {
        const uint64_t want = (2 * J.cap + 1) * 8 * J.n;
        size_t free_b = 0, total_b = 0;
        CK(cudaMemGetInfo(&free_b, &total_b));
        if (want <= J.stage->cap || want <= std::min<uint64_t>((uint64_t)free_b / 4, 8ull << 30)) J.stride = 2 * J.cap + 1;
    }
    uint32_t* ab = J.above ? J.above->as<uint32_t>() : nullptr;
    if (J.stride) {
        J.stage->reserve(J.stride * 8 * J.n);
        k_occ_walk_staged<<<blocks, kBlock, 0, st>>>(R.view, J.head, J.up, J.down, J.len, J.n, J.cap, J.stride, J.off->as<uint64_t>(),
                                                     J.stage->as<uint64_t>(), ab, J.flag->as<uint32_t>());
    } else {
        k_occ_count<<<blocks, kBlock, 0, st>>>(R.view, J.head, J.up, J.down, J.len, J.n, J.cap, J.off->as<uint64_t>(), ab, J.flag->as<uint32_t>());
    }
    g_launches++;
    size_t tmp = 0;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp, J.off->as<uint64_t>(), J.off->as<uint64_t>(), (int)(J.n + 1), st));
    J.cub->reserve(tmp);
    CK(cub::DeviceScan::ExclusiveSum(J.cub->p, tmp, J.off->as<uint64_t>(), J.off->as<uint64_t>(), (int)(J.n + 1), st));
    g_launches += 2;
    uint32_t flag = 0;
    CK(cudaMemcpyAsync(h_off, J.off->p, (J.n + 1) * 8, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(&flag, J.flag->p, 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (truncated && flag) *truncated = true;
    return h_off[J.n];
}

void occ_phase2(Replica& R, cudaStream_t st, OccJob& J, uint64_t total) {
    J.lists->reserve(total * 8 + 16);
    if (J.stride)
        k_occ_pack<<<(unsigned)((J.n * 32 + kBlock - 1) / kBlock), kBlock, 0, st>>>(J.stage->as<uint64_t>(), J.stride, J.off->as<uint64_t>(), J.n,
                                                                                    J.lists->as<uint64_t>());
    else
        k_occ_fill<<<(unsigned)((J.n + kBlock - 1) / kBlock), kBlock, 0, st>>>(R.view, J.head, J.up, J.down, J.len, J.n, J.cap, J.off->as<uint64_t>(),
                                                                               J.lists->as<uint64_t>());
    g_launches++;
    CK(cudaGetLastError());
}

int check_occ_args(moni_gpu_index* idx, const uint64_t* seed_ref, const uint32_t* seed_len, uint64_t n_seeds, uint64_t max_per_direction) {
    if (n_seeds >= 0x7FFFFFF0ull) { g_err = "too many seeds"; return MONI_GPU_EINVAL; }
    // the reference asserts len > 0 in find_MEM_above / find_MEM_below (include/aligner/seed_finder.hpp:171,203): a zero-length seed
    // would "match" every suffix and walk the whole suffix array; a cap beyond 2^32 positions per direction cannot be stored anyway
    if (max_per_direction > (1ull << 32)) { g_err = "max_per_direction must be at most 2^32"; return MONI_GPU_EINVAL; }
    for (uint64_t i = 0; i < n_seeds; ++i)
        if (seed_len[i] == 0) { g_err = "seed " + std::to_string(i) + " has length 0 (the reference requires len > 0)"; return MONI_GPU_EINVAL; }
    if (idx->hdr.phi_n == 0) { g_err = "this index has no Phi structures: load <prefix>.thrbv.full.ms (the `moni align` index)"; return MONI_GPU_EINVAL; }
    if (!(idx->hdr.root_ref & kLeafBit)) { g_err = "occurrence enumeration needs the text in HBM (leaf_len = 0 / text_budget)"; return MONI_GPU_EINVAL; }
    (void)seed_ref;
    return 0;
}

}  // namespace

int moni_gpu_seed_occs(moni_gpu_index* idx, const uint64_t* seed_ref, const uint32_t* seed_len, uint64_t n_seeds,
                       uint64_t max_per_direction, uint64_t* occ_offsets, uint64_t* occs, uint64_t occ_capacity, int* truncated) {
    if (!idx || !occ_offsets || (n_seeds && (!seed_ref || !seed_len))) { g_err = "bad arguments"; return MONI_GPU_EINVAL; }
    if (truncated) *truncated = 0;
    if (n_seeds == 0) { occ_offsets[0] = 0; return 0; }
    if (!occs && occ_capacity) { g_err = "null output"; return MONI_GPU_EINVAL; }
    if (const int rc = check_occ_args(idx, seed_ref, seed_len, n_seeds, max_per_direction)) return rc;
    return guarded([&]() -> int {
        Replica& R = *idx->reps[0];
        std::lock_guard<std::mutex> lock(R.mu);
        CK(cudaSetDevice(R.dev));
        cudaStream_t st = R.stream();
        Lane& L0 = R.lanes[0];
        L0.used = 0;
        L0.take();
        cudaEvent_t* e = L0.events.data();      // 0 start, 1 seeds on the device, 2 counted + scanned, 3 filled, 4 home
        CK(cudaEventRecord(e[0], st));
        R.occ_ref.reserve(n_seeds * 8);
        R.occ_len.reserve(n_seeds * 4);
        CK(cudaMemcpyAsync(R.occ_ref.p, seed_ref, n_seeds * 8, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(R.occ_len.p, seed_len, n_seeds * 4, cudaMemcpyHostToDevice, st));
        CK(cudaEventRecord(e[1], st));
        OccJob J;
        J.head = J.up = J.down = R.occ_ref.as<uint64_t>();
        J.len = R.occ_len.as<uint32_t>();
        J.n = n_seeds;
        J.cap = max_per_direction;
        J.off = &R.occ_off; J.stage = &R.occ_stage; J.cub = &R.occ_cub; J.flag = &R.occ_flag; J.lists = &R.occ_out;
        bool tr = false;
        const uint64_t total = occ_phase1(R, st, J, occ_offsets, &tr);
        CK(cudaEventRecord(e[2], st));
        if (truncated) *truncated = tr ? 1 : 0;
        if (total > occ_capacity) {
            g_err = "occ_capacity too small: " + std::to_string(total) + " positions needed";
            return MONI_GPU_ENOMEM;
        }
        occ_phase2(R, st, J, total);
        CK(cudaEventRecord(e[3], st));
        CK(cudaMemcpyAsync(occs, R.occ_out.p, total * 8, cudaMemcpyDeviceToHost, st));
        CK(cudaEventRecord(e[4], st));
        CK(cudaStreamSynchronize(st));
        // moni_gpu_timing: t[0] seeds in, t[1] count walk + scan, t[2] fill walk (after the offsets went home), t[4] positions out
        R.t[0] = ev_ms(e[0], e[1]); R.t[1] = ev_ms(e[1], e[2]); R.t[2] = ev_ms(e[2], e[3]); R.t[3] = 0; R.t[4] = ev_ms(e[3], e[4]);
        R.t[5] = ev_ms(e[0], e[4]);
        for (int k = 0; k < 6; ++k) idx->t[k] = R.t[k];
        return 0;
    });
}

int moni_gpu_populate_seeds(moni_gpu_index* idx, const uint64_t* seed_ref, const uint32_t* seed_len, const uint32_t* seed_qpos, uint64_t n_seeds,
                            uint64_t min_len, uint64_t max_per_direction, uint64_t* occ_offsets, uint64_t* occs, uint64_t occ_capacity,
                            uint64_t* n_halves, uint64_t* half_parent, uint64_t* half_ref, uint32_t* half_len, uint32_t* half_qpos,
                            uint64_t* half_occ_offsets, uint64_t* half_occs, uint64_t half_occ_capacity, int* truncated) {
    if (!idx || !occ_offsets || !n_halves || !half_occ_offsets || (n_seeds && (!seed_ref || !seed_len || !seed_qpos))) { g_err = "bad arguments"; return MONI_GPU_EINVAL; }
    if (truncated) *truncated = 0;
    *n_halves = 0;
    half_occ_offsets[0] = 0;
    if (n_seeds == 0) { occ_offsets[0] = 0; return 0; }
    if ((!occs && occ_capacity) || (!half_occs && half_occ_capacity)) { g_err = "null output"; return MONI_GPU_EINVAL; }
    if (const int rc = check_occ_args(idx, seed_ref, seed_len, n_seeds, max_per_direction)) return rc;
    // a half of length len >> 1 exists only for len >= 2 * min_len; with min_len == 0 a one-base seed would get a left half of length 0,
    // which the reference's walks assert against
    if (min_len == 0) { g_err = "min_len must be at least 1"; return MONI_GPU_EINVAL; }
    uint64_t want_halves = 0;
    for (uint64_t i = 0; i < n_seeds; ++i)
        if ((uint64_t)seed_len[i] >= (min_len << 1)) want_halves += 2;
    if (want_halves && (!half_parent || !half_ref || !half_len || !half_qpos)) { g_err = "null output"; return MONI_GPU_EINVAL; }
    return guarded([&]() -> int {
        Replica& R = *idx->reps[0];
        std::lock_guard<std::mutex> lock(R.mu);
        CK(cudaSetDevice(R.dev));
        cudaStream_t st = R.stream();
        Lane& L0 = R.lanes[0];
        L0.used = 0;
        L0.take();
        cudaEvent_t* e = L0.events.data();      // 0 start, 1 lists of the seeds on the device, 2 lists of the halves on the device, 3 home
        CK(cudaEventRecord(e[0], st));
        R.occ_ref.reserve(n_seeds * 8);
        R.occ_len.reserve(n_seeds * 4);
        R.pop_qpos.reserve(n_seeds * 4);
        CK(cudaMemcpyAsync(R.occ_ref.p, seed_ref, n_seeds * 8, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(R.occ_len.p, seed_len, n_seeds * 4, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(R.pop_qpos.p, seed_qpos, n_seeds * 4, cudaMemcpyHostToDevice, st));
        OccJob J;
        J.head = J.up = J.down = R.occ_ref.as<uint64_t>();
        J.len = R.occ_len.as<uint32_t>();
        J.n = n_seeds;
        J.cap = max_per_direction;
        J.off = &R.occ_off; J.stage = &R.occ_stage; J.cub = &R.occ_cub; J.flag = &R.occ_flag; J.lists = &R.occ_out; J.above = &R.pop_above;
        bool tr = false;
        const uint64_t total = occ_phase1(R, st, J, occ_offsets, &tr);
        if (total > occ_capacity) {
            g_err = "occ_capacity too small: " + std::to_string(total) + " positions needed";
            return MONI_GPU_ENOMEM;
        }
        occ_phase2(R, st, J, total);
        CK(cudaMemcpyAsync(occs, R.occ_out.p, total * 8, cudaMemcpyDeviceToHost, st));
        CK(cudaEventRecord(e[1], st));
        *n_halves = want_halves;
        if (want_halves) {
            const uint64_t nh = want_halves;
            R.pop_slot.reserve((n_seeds + 1) * 8);
            for (DevBuf* b : {&R.pop_head, &R.pop_up, &R.pop_down, &R.pop_parent}) b->reserve(nh * 8);
            R.pop_len.reserve(nh * 4);
            R.pop_hq.reserve(nh * 4);
            const unsigned blocks = (unsigned)((n_seeds + 1 + kBlock - 1) / kBlock);
            k_seed_qualifies<<<blocks, kBlock, 0, st>>>(R.occ_len.as<uint32_t>(), n_seeds, min_len, R.pop_slot.as<uint64_t>());
            size_t tmp = 0;
            CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp, R.pop_slot.as<uint64_t>(), R.pop_slot.as<uint64_t>(), (int)(n_seeds + 1), st));
            R.occ_cub.reserve(tmp);
            CK(cub::DeviceScan::ExclusiveSum(R.occ_cub.p, tmp, R.pop_slot.as<uint64_t>(), R.pop_slot.as<uint64_t>(), (int)(n_seeds + 1), st));
            k_make_halves<<<blocks, kBlock, 0, st>>>(R.occ_ref.as<uint64_t>(), R.occ_len.as<uint32_t>(), R.pop_qpos.as<uint32_t>(), n_seeds, min_len,
                                                     R.pop_slot.as<uint64_t>(), R.occ_off.as<uint64_t>(), R.occ_out.as<uint64_t>(),
                                                     R.pop_above.as<uint32_t>(), R.pop_head.as<uint64_t>(), R.pop_up.as<uint64_t>(),
                                                     R.pop_down.as<uint64_t>(), R.pop_len.as<uint32_t>(), R.pop_hq.as<uint32_t>(),
                                                     R.pop_parent.as<uint64_t>());
            g_launches += 4;
            CK(cudaGetLastError());
            OccJob H;
            H.head = R.pop_head.as<uint64_t>(); H.up = R.pop_up.as<uint64_t>(); H.down = R.pop_down.as<uint64_t>();
            H.len = R.pop_len.as<uint32_t>();
            H.n = nh;
            H.cap = max_per_direction;
            H.off = &R.pop_hoff; H.stage = &R.occ_stage; H.cub = &R.occ_cub; H.flag = &R.occ_flag; H.lists = &R.pop_hout;
            const uint64_t htotal = occ_phase1(R, st, H, half_occ_offsets, &tr);
            if (htotal > half_occ_capacity) {
                g_err = "half_occ_capacity too small: " + std::to_string(htotal) + " positions needed";
                return MONI_GPU_ENOMEM;
            }
            occ_phase2(R, st, H, htotal);
            CK(cudaEventRecord(e[2], st));
            CK(cudaMemcpyAsync(half_parent, R.pop_parent.p, nh * 8, cudaMemcpyDeviceToHost, st));
            CK(cudaMemcpyAsync(half_ref, R.pop_head.p, nh * 8, cudaMemcpyDeviceToHost, st));
            CK(cudaMemcpyAsync(half_len, R.pop_len.p, nh * 4, cudaMemcpyDeviceToHost, st));
            CK(cudaMemcpyAsync(half_qpos, R.pop_hq.p, nh * 4, cudaMemcpyDeviceToHost, st));
            CK(cudaMemcpyAsync(half_occs, R.pop_hout.p, htotal * 8, cudaMemcpyDeviceToHost, st));
        } else {
            CK(cudaEventRecord(e[2], st));
        }
        CK(cudaEventRecord(e[3], st));
        CK(cudaStreamSynchronize(st));
        if (truncated) *truncated = tr ? 1 : 0;
        R.t[0] = 0; R.t[1] = ev_ms(e[0], e[1]); R.t[2] = ev_ms(e[1], e[2]); R.t[3] = 0; R.t[4] = ev_ms(e[2], e[3]); R.t[5] = ev_ms(e[0], e[3]);
        for (int k = 0; k < 6; ++k) idx->t[k] = R.t[k];
        return 0;
    });
}

int moni_gpu_counters(moni_gpu_index* idx, uint64_t c[8], int reset) {
    if (!idx || !c) { g_err = "bad arguments"; return MONI_GPU_EINVAL; }
    return guarded([&]() -> int {
        for (int k = 0; k < 8; ++k) c[k] = 0;
        for (auto& R : idx->reps) {
            std::lock_guard<std::mutex> lock(R->mu);
            CK(cudaSetDevice(R->dev));
            Counters h;
            CK(cudaMemcpyAsync(&h, R->counters.p, sizeof(h), cudaMemcpyDeviceToHost, R->stream()));
            if (reset) CK(cudaMemsetAsync(R->counters.p, 0, sizeof(Counters), R->stream()));
            CK(cudaStreamSynchronize(R->stream()));
            const unsigned long long* p = &h.bases;
            for (int k = 0; k < 8; ++k) c[k] += p[k];
        }
        return 0;
    });
}

int moni_gpu_timing(const moni_gpu_index* idx, double t[6]) {
    if (!idx || !t) { g_err = "bad arguments"; return MONI_GPU_EINVAL; }
    for (int k = 0; k < 6; ++k) t[k] = idx->t[k];
    return 0;
}

// ---- resident batches ---------------------------------------------------------------------------
int moni_gpu_batch_create(moni_gpu_index* idx, int gpu_slot, const char* seqs, const uint64_t* offsets, uint64_t n_reads, moni_gpu_batch** out) {
    if (!idx || !offsets || !out || gpu_slot < 0 || gpu_slot >= (int)idx->reps.size() || n_reads == 0 || n_reads >= 0x7FFFFFE0ull) {
        g_err = "bad arguments";
        return MONI_GPU_EINVAL;
    }
    if (offsets[n_reads] - offsets[0] >= 0xFFFFFFFFull) { g_err = "a resident batch holds fewer than 2^32 bases"; return MONI_GPU_EINVAL; }
    *out = nullptr;
    return guarded([&]() -> int {
        std::unique_ptr<moni_gpu_batch> b(new moni_gpu_batch);
        b->idx = idx;
        b->rep = idx->reps[gpu_slot].get();
        Replica& R = *b->rep;
        std::lock_guard<std::mutex> lock(R.mu);
        CK(cudaSetDevice(R.dev));
        b->plan = plan_chunk(offsets, 0, n_reads, ~0ull, 0, ~0ull);
        const ChunkPlan& p = b->plan;
        b->s.reserve(p, kModeMs);
        b->s.reserve(p, kModeMems);
        if (p.bases) CK(cudaMemcpyAsync(b->s.seqs.p, seqs + p.base, p.bases, cudaMemcpyHostToDevice, R.stream()));
        CK(cudaMemcpyAsync(b->s.offs.p, offsets, (n_reads + 1) * 8, cudaMemcpyHostToDevice, R.stream()));
        CK(cudaStreamSynchronize(R.stream()));
        *out = b.release();
        return 0;
    });
}

int moni_gpu_batch_run(moni_gpu_batch* b, int what, float kernel_ms[3]) {
    if (!b || (what != 1 && what != 2)) { g_err = "bad arguments"; return MONI_GPU_EINVAL; }
    return guarded([&]() -> int {
        Replica& R = *b->rep;
        std::lock_guard<std::mutex> lock(R.mu);
        CK(cudaSetDevice(R.dev));
        Lane& L = R.lanes[0];
        L.used = 0;
        L.take();
        cudaEvent_t* e = L.events.data();
        CK(cudaEventRecord(e[0], L.stream));
        launch_prepare(b->s, b->plan, L.stream);
        launch_pass1(R, b->s, b->plan, L.stream);
        CK(cudaEventRecord(e[1], L.stream));
        launch_pass2(R, b->s, b->plan, what == 2 ? kModeMems : kModeMs, L.stream);
        CK(cudaEventRecord(e[2], L.stream));
        if (what == 2) launch_compact(b->s, b->plan, 8, L.stream);
        else launch_expand(b->s, b->plan, L.stream);
        CK(cudaEventRecord(e[3], L.stream));
        CK(cudaStreamSynchronize(L.stream));
        b->last = what;
        if (kernel_ms) {
            kernel_ms[0] = ev_ms(e[0], e[1]);
            kernel_ms[1] = ev_ms(e[1], e[2]);
            kernel_ms[2] = ev_ms(e[2], e[3]);
        }
        return 0;
    });
}

int moni_gpu_batch_fetch_ms(moni_gpu_batch* b, uint64_t* pointers, uint64_t* lengths) {
    if (!b || !pointers || !lengths) { g_err = "bad arguments"; return MONI_GPU_EINVAL; }
    if (b->last != 1) { g_err = "batch was not last run in ms mode"; return MONI_GPU_EINVAL; }
    return guarded([&]() -> int {
        Replica& R = *b->rep;
        std::lock_guard<std::mutex> lock(R.mu);
        CK(cudaSetDevice(R.dev));
        if (b->plan.bases) {
            CK(cudaMemcpyAsync(pointers, b->s.out_ptr.p, b->plan.bases * 8, cudaMemcpyDeviceToHost, R.stream()));
            CK(cudaMemcpyAsync(lengths, b->s.out_len.p, b->plan.bases * 8, cudaMemcpyDeviceToHost, R.stream()));
        }
        CK(cudaStreamSynchronize(R.stream()));
        return 0;
    });
}

int moni_gpu_batch_fetch_mems(moni_gpu_batch* b, uint64_t* mem_offsets, uint64_t* mem_pos, uint64_t* mem_len, uint64_t mem_capacity) {
    if (!b || !mem_offsets) { g_err = "bad arguments"; return MONI_GPU_EINVAL; }
    if (b->last != 2) { g_err = "batch was not last run in mems mode"; return MONI_GPU_EINVAL; }
    return guarded([&]() -> int {
        Replica& R = *b->rep;
        std::lock_guard<std::mutex> lock(R.mu);
        CK(cudaSetDevice(R.dev));
        const uint64_t n = b->plan.n;
        CK(cudaMemcpyAsync(mem_offsets, b->s.moff.p, (n + 1) * 8, cudaMemcpyDeviceToHost, R.stream()));
        CK(cudaStreamSynchronize(R.stream()));
        const uint64_t k = mem_offsets[n];
        if (k > mem_capacity) { g_err = "mem_capacity too small: " + std::to_string(k) + " records needed"; return MONI_GPU_ENOMEM; }
        if (k) {
            CK(cudaMemcpyAsync(mem_pos, b->s.mem_pos.p, k * 8, cudaMemcpyDeviceToHost, R.stream()));
            CK(cudaMemcpyAsync(mem_len, b->s.mem_len.p, k * 8, cudaMemcpyDeviceToHost, R.stream()));
        }
        CK(cudaStreamSynchronize(R.stream()));
        return 0;
    });
}

void moni_gpu_batch_free(moni_gpu_batch* b) {
    if (!b) return;
    cudaSetDevice(b->rep->dev);
    b->s.release();
    delete b;
}

}  // extern "C"

// ================================================================================ allocation-free entry point
// moni_gpu_ms_noalloc: moni_gpu_ms for callers that may not allocate (include/moni_gpu.h).  What the two passes exchange -- the
// staged read bytes, restart flags, pointers at the restarts, restart records, restarts per read -- lives in ONE file-scope array
// that every GPU holding the library's module carries until the process exits, instead of the cudaMalloc'ed Scratch of a lane.
// The kernels are the ones above, handed addresses inside the array.
constexpr uint64_t kNoAllocMaxBases = MONI_GPU_NOALLOC_MAX_BASES, kNoAllocMaxReads = MONI_GPU_NOALLOC_MAX_READS;
constexpr uint64_t na_round(uint64_t x) { return (x + 255) & ~255ull; }
constexpr uint64_t kNaRunEntries = kNoAllocMaxBases + 4 * kNoAllocMaxReads + 8;      // ChunkPlan::run_entries()
constexpr uint64_t kNaFlagWords = kNoAllocMaxBases / 32 + kNoAllocMaxReads + 2;      // ChunkPlan::flag_words()
constexpr uint64_t kNaOffStatus = 0;                                                  // one u32, cleared by every call
constexpr uint64_t kNaOffSeqs = 256;                                                  // bases + 16 bytes of slack (load8)
constexpr uint64_t kNaOffFlags = kNaOffSeqs + na_round(kNoAllocMaxBases + 16);
constexpr uint64_t kNaOffRcnt = kNaOffFlags + na_round(kNaFlagWords * 4);
constexpr uint64_t kNaOffRptr = kNaOffRcnt + na_round((kNoAllocMaxReads + 1) * 4);
constexpr uint64_t kNaOffRrec = kNaOffRptr + na_round(kNaRunEntries * 8);
constexpr uint64_t kNaBytes = kNaOffRrec + na_round(kNaRunEntries * 8);
static_assert(kNaBytes <= (32ull << 20), "the allocation-free scratch is meant to stay modest");

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNaBytes];

namespace {

// offsets[] as the caller left them on the device: 0 = offsets[0] <= offsets[1] <= ... <= offsets[n] = n_bases, or status != 0
__global__ void k_noalloc_check(const uint64_t* __restrict__ offs, uint32_t n, uint64_t n_bases, uint32_t* __restrict__ status) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r > n) return;
    const uint64_t a = offs[r];
    bool bad = a > n_bases;
    if (r == 0 && a != 0) bad = true;
    if (r == n && a != n_bases) bad = true;
    if (r < n && offs[r + 1] < a) bad = true;
    if (bad) atomicOr(status, 1u);
}

std::mutex g_noalloc_mu;                    // one caller at a time per process: every call uses the same array
unsigned char* g_noalloc_ptr[64] = {};      // device address of g_noalloc_scratch on CUDA device d (the host-side name is only a handle)

unsigned char* noalloc_scratch_on(int dev) {   // with `dev` current
    if (dev < 0 || dev >= 64) throw BadBatch("device ordinal out of range for the allocation-free scratch");
    if (!g_noalloc_ptr[dev]) {
        void* p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        g_noalloc_ptr[dev] = (unsigned char*)p;
    }
    return g_noalloc_ptr[dev];
}

bool on_device(const void* p, int dev) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return (a.type == cudaMemoryTypeDevice && a.device == dev) || a.type == cudaMemoryTypeManaged;
}

}  // namespace

extern "C" int moni_gpu_ms_noalloc(moni_gpu_index* idx, int gpu_slot, const char* d_seqs, const uint64_t* d_offsets, uint64_t n_reads,
                                   uint64_t n_bases, uint64_t* d_pointers, uint64_t* d_lengths, void* stream) {
    if (!idx || gpu_slot < 0 || gpu_slot >= (int)idx->reps.size() || !d_offsets || (n_bases && (!d_seqs || !d_pointers || !d_lengths))) {
        g_err = "bad arguments";
        return MONI_GPU_EINVAL;
    }
    if (n_reads > kNoAllocMaxReads || n_bases > kNoAllocMaxBases) {
        g_err = "moni_gpu_ms_noalloc takes at most " + std::to_string(kNoAllocMaxReads) + " reads and " + std::to_string(kNoAllocMaxBases) +
                " bases per call (got " + std::to_string(n_reads) + " reads, " + std::to_string(n_bases) +
                " bases): its scratch is a fixed array; split the batch or use moni_gpu_ms";
        return MONI_GPU_EINVAL;
    }
    return guarded([&]() -> int {
        Replica& R = *idx->reps[gpu_slot];
        std::lock_guard<std::mutex> lock(g_noalloc_mu);
        CK(cudaSetDevice(R.dev));
        if (!on_device(d_offsets, R.dev) || (n_bases && (!on_device(d_seqs, R.dev) || !on_device(d_pointers, R.dev) || !on_device(d_lengths, R.dev))))
            throw BadBatch("moni_gpu_ms_noalloc takes device pointers on the GPU of gpu_slot (seqs, offsets, pointers, lengths)");
        unsigned char* const base = noalloc_scratch_on(R.dev);
        cudaStream_t st = (cudaStream_t)stream;
        uint32_t* status = (uint32_t*)(base + kNaOffStatus);
        uint8_t* seqs = base + kNaOffSeqs;
        CK(cudaMemsetAsync(status, 0, 256, st));
        const uint32_t n = (uint32_t)n_reads;
        k_noalloc_check<<<(n + 1 + 255) / 256, 256, 0, st>>>(d_offsets, n, n_bases, status);
        g_launches++;
        CK(cudaGetLastError());
        uint32_t h_status = 0;
        CK(cudaMemcpyAsync(&h_status, status, 4, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        if (h_status) throw BadBatch("offsets must start at 0, be non-decreasing and end at n_bases");
        if (!n_reads || !n_bases) return 0;
        // the reads, with the 16 bytes the 8-byte compares may touch behind the last one (the array keeps the last call's bytes)
        CK(cudaMemcpyAsync(seqs, d_seqs, n_bases, cudaMemcpyDeviceToDevice, st));
        CK(cudaMemsetAsync(seqs + n_bases, 0, 16, st));
        ChunkPlan p;
        p.first = 0; p.n = n_reads; p.base = 0; p.bases = n_bases;
        ChunkMap cm;
        cm.offs = d_offsets; cm.perm = nullptr; cm.base = 0; cm.n_reads = n;
        cm.seq_bytes = n_bases + 16; cm.flag_words = p.flag_words(); cm.run_entries = p.run_entries();
        uint32_t* flags = (uint32_t*)(base + kNaOffFlags);
        uint32_t* rcnt = (uint32_t*)(base + kNaOffRcnt);
        uint64_t* rptr = (uint64_t*)(base + kNaOffRptr);
        uint64_t* rrec = (uint64_t*)(base + kNaOffRrec);
        const dim3 grid((unsigned)((n_reads + kBlock - 1) / kBlock));
        k_ms_pointers<false><<<grid, kBlock, 0, st>>>(R.view, seqs, cm, rptr, flags, rcnt, nullptr);
        RunBufs B;
        B.seqs = seqs; B.flags = flags; B.rptr = rptr; B.rrec = rrec; B.rec = nullptr; B.rec_ref = nullptr;
        B.seq_bytes = cm.seq_bytes; B.flag_words = cm.flag_words; B.run_entries = cm.run_entries; B.rec_entries = 0;
        k_ms_lengths<0, false><<<grid, kBlock, 0, st>>>(R.view, B, cm, rcnt, nullptr, nullptr, 0);
        k_ms_expand<false><<<dim3((unsigned)((n_reads * 32 + kBlock - 1) / kBlock)), kBlock, 0, st>>>(cm, rptr, rrec, flags, rcnt, d_pointers, d_lengths,
                                                                                                    nullptr, nullptr, nullptr);
        g_launches += 3;
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(st));
        return 0;
    });
}
