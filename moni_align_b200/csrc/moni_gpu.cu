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
#include "index_files.hpp"
#include "ms_core.cuh"

using namespace moni;

// ================================================================================ kernels
namespace {

#ifndef MONI_P1_MINBLOCKS
#define MONI_P1_MINBLOCKS 1          // resident CTAs per SM the pass-1 kernel is compiled for (register cap = 65536 / (128 * this))
#endif
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

// How the reads of a chunk map to threads and where a warp keeps its pointers.
//   slot  = global thread index; read = perm ? perm[slot] : slot   (perm sorts ragged chunks by length,
//           longest first, so the 32 reads of a warp retire together and padding stays small)
//   pointers of warp w: ptr_t[wbase(w) + step * 32 + lane], step = number of symbols consumed before
//           (query offset i = m - 1 - step); wbase(w) = w * 32 * uniform_len, or wbase[w] for ragged chunks.
struct ChunkMap {
    const uint64_t* offs;      // n_reads + 1 absolute offsets into the caller's seqs
    const uint32_t* perm;      // or null
    const uint64_t* wbase;     // or null (uniform_len != 0)
    uint64_t base;             // offs[0]: seqs on the device start here
    uint32_t n_reads;
    uint32_t uniform_len;
};

__device__ __forceinline__ uint64_t warp_base(const ChunkMap& cm, uint32_t w) {
    return cm.wbase ? __ldg(cm.wbase + w) : (uint64_t)w * 32u * cm.uniform_len;
}

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

// ext[w] = 32 * (longest read of warp w); ext[n_warps] = 0.  An exclusive scan turns it into wbase.
__global__ void k_warp_extent(const uint64_t* __restrict__ offs, const uint32_t* __restrict__ perm, uint32_t n, uint32_t n_warps,
                              uint64_t* __restrict__ ext) {
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x, w = slot >> 5;
    if (w > n_warps) return;
    uint32_t m = 0;
    if (slot < n) {
        const uint32_t r = perm ? perm[slot] : slot;
        m = (uint32_t)(offs[r + 1] - offs[r]);
    }
    m = warp_max(m);
    if ((threadIdx.x & 31) == 0) ext[w] = w < n_warps ? (uint64_t)m * 32u : 0;
}

// ---- pass 1: MS pointers -------------------------------------------------------------------------
template <bool COUNT>
__global__ void __launch_bounds__(kBlock, MONI_P1_MINBLOCKS) k_ms_pointers(ImageView v, const uint8_t* __restrict__ seqs, ChunkMap cm,
                                                        uint64_t* __restrict__ ptr_t, Counters* g) {
    __shared__ uint8_t tcode[256];
    __shared__ __align__(16) uint8_t tiles[kBlock / 32][32 * kTileStride];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) tcode[i] = v.sym->tcode[i];
    __syncthreads();
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31, w = slot >> 5;
    if ((w << 5) >= cm.n_reads) return;                       // warp-uniform
    uint8_t* tile = tiles[threadIdx.x >> 5];
    uint32_t o = 0, m = 0;                                    // chunk-relative offset (< 2^32) and length
    if (slot < cm.n_reads) {
        const uint32_t r = cm.perm ? __ldg(cm.perm + slot) : slot;
        const uint64_t a = __ldg(cm.offs + r);
        o = (uint32_t)(a - cm.base);
        m = (uint32_t)(__ldg(cm.offs + r + 1) - a);
    }
    const uint32_t mw = warp_max(m);
    uint64_t* out = ptr_t + warp_base(cm, w) + lane;
    ScanState s;
    scan_init(v, s);
    Counters c = {0, 0, 0, 0, 0, 0, 0, 0};
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
                scan_step(v, tcode, tile[lane * kTileStride + t], s, COUNT ? &c : nullptr);
                out[(uint64_t)step * 32u] = s.sample;
            }
        }
    }
    if (COUNT) { c.bases = m; flush_counters(c, g); }
}

// ---- pass 2: MS lengths / MEM records / seeds --------------------------------------------------------
// MODE 0 (ms): the length of query offset i goes to len_t[] next to its pointer (same warp-transposed place, u32), so the
// 32 lanes of a warp write into the same lines; k_ms_untranspose then produces the caller's layout with coalesced
// accesses on both sides.  MODE 1 (mems): the read's MEM records, packed (i << 32) | len, go to rec[] at the read's own
// offset (at most one per base) and counts[read] receives their number.  MODE 2 (seeds): same with
// seed_finder::find_mems' rule (minimum length, N filter) and rec_ref[] = pointers[i].
template <int MODE, bool COUNT>
__global__ void __launch_bounds__(kBlock) k_ms_lengths(ImageView v, const uint8_t* __restrict__ seqs, ChunkMap cm,
                                                       const uint64_t* __restrict__ ptr_t, uint32_t* __restrict__ len_t,
                                                       uint64_t* __restrict__ rec, uint32_t* __restrict__ counts, Counters* g,
                                                       uint64_t* __restrict__ rec_ref, uint32_t min_len) {
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31, w = slot >> 5;
    if (slot >= cm.n_reads) return;
    const uint32_t r = cm.perm ? __ldg(cm.perm + slot) : slot;
    const uint64_t a = __ldg(cm.offs + r);
    const uint64_t o = a - cm.base, m = __ldg(cm.offs + r + 1) - a;
    if (m == 0) { if (MODE) counts[r] = 0; return; }
    const uint64_t at = warp_base(cm, w) + lane + (m - 1) * 32u;              // query offset 0 = last step
    Counters c = {0, 0, 0, 0, 0, 0, 0, 0};
    const uint32_t k = lengths_read<uint32_t, MODE == 2>(v, seqs + o, m, ptr_t + at, -32, nullptr, MODE ? nullptr : len_t + at,
                                                         MODE ? rec + o : nullptr, COUNT ? &c : nullptr,
                                                         MODE == 2 ? rec_ref + o : nullptr, min_len, -32);
    if (MODE) counts[r] = k;
    if (COUNT) flush_counters(c, g);
}

// ms: warp-transposed (pointer, length) -> the caller's layout (u64 each, laid out like seqs).  One warp per 32 reads,
// 32 steps at a time through shared memory: 256-byte loads along the lanes, 256-byte stores along each read.
constexpr int kUtBlock = 64;
// PACKED: 7 bytes per base instead of 16 -- the low 40 bits of the pointer as u32 + u8 (a pointer is a text position < n,
// or 2^64 - k with k < 2^31 after the reference's `sample--` wrapped: both survive `& (2^40 - 1)` when n <= 2^40 - 2^31)
// and the length as u16 (reads shorter than 65536 bases).
template <bool PACKED>
__global__ void __launch_bounds__(kUtBlock) k_ms_untranspose(ChunkMap cm, const uint64_t* __restrict__ ptr_t,
                                                             const uint32_t* __restrict__ len_t, uint64_t* __restrict__ ptr_out,
                                                             uint64_t* __restrict__ len_out, uint32_t* __restrict__ lo_out,
                                                             uint8_t* __restrict__ hi_out, uint16_t* __restrict__ len16_out) {
    __shared__ uint64_t tp[kUtBlock / 32][32 * 33];
    __shared__ uint32_t tl[kUtBlock / 32][32 * 33];
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31, w = slot >> 5;
    if ((w << 5) >= cm.n_reads) return;                       // warp-uniform
    uint64_t* P = tp[threadIdx.x >> 5];
    uint32_t* Q = tl[threadIdx.x >> 5];
    uint64_t o = 0;
    uint32_t m = 0;
    if (slot < cm.n_reads) {
        const uint32_t r = cm.perm ? __ldg(cm.perm + slot) : slot;
        const uint64_t a = __ldg(cm.offs + r);
        o = a - cm.base;
        m = (uint32_t)(__ldg(cm.offs + r + 1) - a);
    }
    const uint32_t mw = warp_max(m);
    const uint64_t base = warp_base(cm, w) + lane;
    for (uint32_t t0 = 0; t0 < mw; t0 += 32) {
        const uint32_t lim = min(32u, mw - t0);
#pragma unroll 8
        for (uint32_t t = 0; t < lim; ++t) {                  // entries of lanes whose read is shorter are never stored below
            P[t * 33 + lane] = __ldg(ptr_t + base + (uint64_t)(t0 + t) * 32u);
            Q[t * 33 + lane] = __ldg(len_t + base + (uint64_t)(t0 + t) * 32u);
        }
        __syncwarp();
#pragma unroll 4
        for (int rr = 0; rr < 32; ++rr) {
            const uint64_t o_r = __shfl_sync(0xFFFFFFFFu, o, rr);
            const uint32_t m_r = __shfl_sync(0xFFFFFFFFu, m, rr);
            const uint32_t step = t0 + lane;
            if (step < m_r) {
                const uint64_t at = o_r + (m_r - 1 - step);
                const uint64_t pv = P[lane * 33 + rr];
                if (PACKED) {
                    lo_out[at] = (uint32_t)pv;
                    hi_out[at] = (uint8_t)(pv >> 32);
                    len16_out[at] = (uint16_t)Q[lane * 33 + rr];
                } else {
                    ptr_out[at] = pv;
                    len_out[at] = Q[lane * 33 + rr];
                }
            }
        }
        __syncwarp();
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
__global__ void __launch_bounds__(kBlock) k_occ_count(ImageView v, const uint64_t* __restrict__ ref, const uint32_t* __restrict__ len,
                                                      uint64_t n_seeds, uint64_t cap, uint64_t* __restrict__ counts,
                                                      uint32_t* __restrict__ truncated) {
    const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (i > n_seeds) return;
    if (i == n_seeds) { counts[i] = 0; return; }
    bool tr = false;
    counts[i] = occ_walk(v, ref[i], len[i], cap, nullptr, &tr);
    if (tr) atomicOr(truncated, 1u);
}

__global__ void __launch_bounds__(kBlock) k_occ_fill(ImageView v, const uint64_t* __restrict__ ref, const uint32_t* __restrict__ len,
                                                     uint64_t n_seeds, uint64_t cap, const uint64_t* __restrict__ off,
                                                     uint64_t* __restrict__ occs) {
    const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (i >= n_seeds) return;
    occ_walk(v, ref[i], len[i], cap, occs + off[i], nullptr);
}

// One walk instead of two when the per-direction cap bounds a seed's list: positions go to a fixed-stride staging area
// (stride = 2 cap + 1; HBM has the room), and a streaming pass packs them once the offsets are known.
__global__ void __launch_bounds__(kBlock) k_occ_walk_staged(ImageView v, const uint64_t* __restrict__ ref, const uint32_t* __restrict__ len,
                                                            uint64_t n_seeds, uint64_t cap, uint64_t stride, uint64_t* __restrict__ counts,
                                                            uint64_t* __restrict__ staging, uint32_t* __restrict__ truncated) {
    const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (i > n_seeds) return;
    if (i == n_seeds) { counts[i] = 0; return; }
    bool tr = false;
    counts[i] = occ_walk(v, ref[i], len[i], cap, staging + i * stride, &tr);
    if (tr) atomicOr(truncated, 1u);
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
    uint32_t lanes = 4;
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

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    void reserve(size_t bytes) {
        if (bytes <= cap) return;
        if (p) CK(cudaFree(p));
        p = nullptr;
        cap = 0;
        const size_t want = bytes + bytes / 8 + 256;
        CK(cudaMalloc(&p, want));
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
    uint64_t n_warps() const { return (n + 31) / 32; }
    bool ragged() const { return uniform_len == 0 && max_len != 0; }
    // entries of the warp-transposed pointer buffer: exact when uniform, else the bound that holds
    // once reads are sorted by length (sum over warps of 32 * longest <= bases + 32 * max_len)
    uint64_t ptr_entries() const { return ragged() ? bases + 32 * max_len + 32 : n_warps() * 32 * (uint64_t)uniform_len + 32; }
};

struct Scratch {   // device buffers of one chunk
    DevBuf seqs, offs, keys_in, keys_out, vals_in, perm, wbase, ptr_t, out_ptr, out_len, rec, counts, moff, mem_pos, mem_len, cub;
    DevBuf rec_ref, mem_ref;     // seeds only
    DevBuf len_t;                // ms only: lengths, warp-transposed like ptr_t
    DevBuf out_hi;               // ms, packed outputs: bits 32..39 of the pointers (out_ptr then holds the u32 halves, out_len u16 lengths)
    void release() {
        for (DevBuf* b : {&seqs, &offs, &keys_in, &keys_out, &vals_in, &perm, &wbase, &ptr_t, &out_ptr, &out_len, &rec, &counts, &moff,
                          &mem_pos, &mem_len, &cub, &rec_ref, &mem_ref, &len_t, &out_hi})
            b->release();
    }
    void reserve(const ChunkPlan& p, bool ms, bool mems, bool seeds = false) {
        if (seeds) { rec_ref.reserve(p.bases * 8 + 16); mem_ref.reserve(p.bases * 8 + 16); }
        seqs.reserve(p.bases + 16);
        offs.reserve((p.n + 1) * 8);
        ptr_t.reserve(p.ptr_entries() * 8);
        if (p.ragged()) {
            keys_in.reserve(p.n * 4); keys_out.reserve(p.n * 4); vals_in.reserve(p.n * 4); perm.reserve(p.n * 4);
            wbase.reserve((p.n_warps() + 1) * 8);
        }
        if (ms) { out_ptr.reserve(p.bases * 8 + 16); out_len.reserve(p.bases * 8 + 16); len_t.reserve(p.ptr_entries() * 4); out_hi.reserve(p.bases + 16); }
        if (mems) {
            rec.reserve(p.bases * 8 + 16);
            counts.reserve((p.n + 1) * 4);
            moff.reserve((p.n + 1) * 16);
            mem_pos.reserve(p.bases * 8 + 16);
            mem_len.reserve(p.bases * 8 + 16);
        }
    }
};

constexpr int kMaxLanes = 8;     // chunks in flight per GPU (option "lanes"): copies of one overlap the kernels of the others
constexpr int kEv = 6;           // events per chunk: start, h2d done, pass 1 done, pass 2 done, compaction done, d2h done

struct Lane {
    cudaStream_t stream = nullptr;
    Scratch scratch;
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
    std::vector<Lane> lanes;
    DevBuf rows, row_pos, crun_row, slots, nodes, pool, sym, counters;
    DevBuf phi, phi_inv, phi_dir, phi_inv_dir;            // `moni align` indexes only
    DevBuf occ_ref, occ_len, occ_off, occ_out, occ_flag, occ_cub, occ_stage;   // moni_gpu_seed_occs scratch
    DevBuf shard_off, shard_pos, shard_len, shard_ref;   // multi-GPU mems: a shard's records wait here for their global offset
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
        shard_off.release(); shard_pos.release(); shard_len.release(); shard_ref.release();
        for (Lane& L : lanes) {
            L.scratch.release();
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
    cm.wbase = p.ragged() ? S.wbase.as<uint64_t>() : nullptr;
    cm.base = p.base;
    cm.n_reads = (uint32_t)p.n;
    cm.uniform_len = p.uniform_len;
    return cm;
}

// Ragged chunks: sort the reads by length (longest first) and lay out the per-warp pointer areas.
void launch_prepare(Scratch& S, const ChunkPlan& p, cudaStream_t st) {
    if (!p.ragged()) return;
    const uint32_t n = (uint32_t)p.n, nw = (uint32_t)p.n_warps();
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
    k_warp_extent<<<(nw * 32 + 32 + 255) / 256, 256, 0, st>>>(S.offs.as<uint64_t>(), S.perm.as<uint32_t>(), n, nw, S.wbase.as<uint64_t>());
    g_launches++;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp, S.wbase.as<uint64_t>(), S.wbase.as<uint64_t>(), (int)(nw + 1), st));
    S.cub.reserve(tmp);
    CK(cub::DeviceScan::ExclusiveSum(S.cub.p, tmp, S.wbase.as<uint64_t>(), S.wbase.as<uint64_t>(), (int)(nw + 1), st));
    g_launches += 2;
    CK(cudaGetLastError());
}

void launch_pass1(Replica& R, Scratch& S, const ChunkPlan& p, cudaStream_t st) {
    if (!p.n) return;
    const dim3 grid((unsigned)((p.n_warps() * 32 + kBlock - 1) / kBlock));
    const ChunkMap cm = chunk_map(S, p);
    if (g_count.load())
        k_ms_pointers<true><<<grid, kBlock, 0, st>>>(R.view, S.seqs.as<uint8_t>(), cm, S.ptr_t.as<uint64_t>(), R.counters.as<Counters>());
    else
        k_ms_pointers<false><<<grid, kBlock, 0, st>>>(R.view, S.seqs.as<uint8_t>(), cm, S.ptr_t.as<uint64_t>(), nullptr);
    g_launches++;
    CK(cudaGetLastError());
}

// mode: 0 = ms, 1 = mems, 2 = seeds
void launch_pass2(Replica& R, Scratch& S, const ChunkPlan& p, int mode, cudaStream_t st, uint32_t min_len = 0) {
    if (!p.n) return;
    const dim3 grid((unsigned)((p.n + kBlock - 1) / kBlock));
    const ChunkMap cm = chunk_map(S, p);
    Counters* g = g_count.load() ? R.counters.as<Counters>() : nullptr;
    const uint8_t* sq = S.seqs.as<uint8_t>();
    const uint64_t* pt = S.ptr_t.as<uint64_t>();
    uint64_t* rec = S.rec.as<uint64_t>();
    uint32_t* cnt = S.counts.as<uint32_t>();
    if (mode == 2) {
        if (g) k_ms_lengths<2, true><<<grid, kBlock, 0, st>>>(R.view, sq, cm, pt, nullptr, rec, cnt, g, S.rec_ref.as<uint64_t>(), min_len);
        else k_ms_lengths<2, false><<<grid, kBlock, 0, st>>>(R.view, sq, cm, pt, nullptr, rec, cnt, nullptr, S.rec_ref.as<uint64_t>(), min_len);
    } else if (mode == 1) {
        if (g) k_ms_lengths<1, true><<<grid, kBlock, 0, st>>>(R.view, sq, cm, pt, nullptr, rec, cnt, g, nullptr, 0);
        else k_ms_lengths<1, false><<<grid, kBlock, 0, st>>>(R.view, sq, cm, pt, nullptr, rec, cnt, nullptr, nullptr, 0);
    } else {
        if (g) k_ms_lengths<0, true><<<grid, kBlock, 0, st>>>(R.view, sq, cm, pt, S.len_t.as<uint32_t>(), nullptr, nullptr, g, nullptr, 0);
        else k_ms_lengths<0, false><<<grid, kBlock, 0, st>>>(R.view, sq, cm, pt, S.len_t.as<uint32_t>(), nullptr, nullptr, nullptr, nullptr, 0);
    }
    g_launches++;
    CK(cudaGetLastError());
}

// ms: pointers and lengths from the warp-transposed buffers to the caller's layout
void launch_untranspose(Scratch& S, const ChunkPlan& p, cudaStream_t st, bool packed = false) {
    if (!p.n) return;
    const dim3 grid((unsigned)((p.n_warps() * 32 + kUtBlock - 1) / kUtBlock));
    if (packed)
        k_ms_untranspose<true><<<grid, kUtBlock, 0, st>>>(chunk_map(S, p), S.ptr_t.as<uint64_t>(), S.len_t.as<uint32_t>(), nullptr, nullptr,
                                                          S.out_ptr.as<uint32_t>(), S.out_hi.as<uint8_t>(), S.out_len.as<uint16_t>());
    else
        k_ms_untranspose<false><<<grid, kUtBlock, 0, st>>>(chunk_map(S, p), S.ptr_t.as<uint64_t>(), S.len_t.as<uint32_t>(), S.out_ptr.as<uint64_t>(),
                                                           S.out_len.as<uint64_t>(), nullptr, nullptr, nullptr);
    g_launches++;
    CK(cudaGetLastError());
}

// counts -> exclusive offsets (n + 1 entries) -> gather.  mem_pos/mem_len hold `bases` entries.
// rec_bytes: 8 = two u64 arrays, 4 = two u32 arrays, 2 = one u32 array of packed (offset:16, length:16) records,
// 12 = seeds (u32 query offset, u32 length, u64 text position)
void launch_compact(Scratch& S, const ChunkPlan& p, int rec_bytes, cudaStream_t st) {
    if (!p.n) return;
    const uint32_t n = (uint32_t)p.n;
    uint64_t* wide = S.moff.as<uint64_t>() + (n + 1);          // second half of moff: widened counts
    k_widen_counts<<<(n + 1 + 255) / 256, 256, 0, st>>>(S.counts.as<uint32_t>(), wide, n);
    g_launches++;
    size_t tmp = 0;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp, wide, S.moff.as<uint64_t>(), (int)(n + 1), st));
    S.cub.reserve(tmp);
    CK(cub::DeviceScan::ExclusiveSum(S.cub.p, tmp, wide, S.moff.as<uint64_t>(), (int)(n + 1), st));
    g_launches += 2;
    const uint64_t threads = (uint64_t)n * 32;
    const unsigned blocks = (unsigned)((threads + kBlock - 1) / kBlock);
    if (rec_bytes == 12)
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
constexpr uint64_t kChunkHardCap = 1ull << 29;

ChunkPlan plan_chunk(const uint64_t* offsets, uint64_t a, uint64_t r1, uint64_t chunk_bases, uint64_t min_reads,
                     uint64_t hard_cap = kChunkHardCap) {
    ChunkPlan p;
    p.first = a;
    p.base = offsets[a];
    uint64_t b = a, lo = ~0ull, hi = 0;
    while (b < r1 && b - a < 0x7FFFFFE0ull) {
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

// Multi-GPU mems: shard g may send its records home once the record totals of shards 0..g-1 are known.
struct Rendezvous {
    std::mutex mu;
    std::condition_variable cv;
    std::vector<int64_t> total;      // -1 = not yet known, -2 = that shard failed
    explicit Rendezvous(int parts) : total(parts, -1) {}
    void publish(int g, int64_t t) {
        { std::lock_guard<std::mutex> lock(mu); total[g] = t; }
        cv.notify_all();
    }
    uint64_t base_of(int g) {
        std::unique_lock<std::mutex> lock(mu);
        uint64_t base = 0;
        for (int h = 0; h < g; ++h) {
            cv.wait(lock, [&] { return total[h] != -1; });
            if (total[h] < 0) throw std::runtime_error("another GPU's shard failed");
            base += (uint64_t)total[h];
        }
        return base;
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

// Processes reads [r0, r1) of a host batch on one replica.  Chunks alternate between the replica's
// lanes (stream + scratch each), so the copies of one chunk overlap the kernels of the other.
// mode 0 = ms, 1 = mems, 2 = seeds.  mems / seeds: records go to out.col[] (contiguous from index 0, capacity out.cap)
// and per-read record offsets (relative to this shard) to out.off[0 .. r1-r0]; *needed receives the shard's record count.
void run_shard(moni_gpu_index* idx, Replica& R, int mode, const char* seqs, const uint64_t* offsets, uint64_t r0, uint64_t r1,
               uint64_t* pointers, uint64_t* lengths, const RecOut& out, uint64_t* needed, Rendezvous* rv = nullptr, int shard = 0) {
    // rv != null: out.off / out.col are the CALLER's whole-batch arrays; this shard's records are kept on
    // the device until the shards before it have published their totals, then copied to their final place.
    const bool mems = mode != 0;
    uint64_t* const rec_off = out.off;
    const uint64_t rec_cap = out.cap;
    std::lock_guard<std::mutex> lock(R.mu);
    CK(cudaSetDevice(R.dev));
    for (double& x : R.t) x = 0;
    // chunks are planned (and their offsets validated) just before they are enqueued, so the host's pass over the
    // offsets overlaps the GPU's work on earlier chunks
    std::vector<ChunkPlan> plans;
    uint64_t next_read = r0;
    const size_t kLanes_ = R.lanes.size();
    for (Lane& L : R.lanes) L.used = 0;
    std::vector<size_t> ev_at;
    uint64_t rec_total = 0;
    if (mems && !rv) rec_off[0] = 0;
    DevBuf* const shard_col[3] = {&R.shard_pos, &R.shard_len, &R.shard_ref};
    if (rv) {
        R.shard_off.reserve((r1 - r0 + 1) * 8);
        for (int k = 0; k < out.ncol; ++k) shard_col[k]->reserve((offsets[r1] - offsets[r0]) * out.width[k] + 16);
    }

    auto enqueue_front = [&](size_t c) {
        const ChunkPlan& p = plans[c];
        Lane& L = R.lanes[c % kLanes_];
        Scratch& S = L.scratch;
        S.reserve(p, !mems, mems, mode == 2);
        ev_at.push_back(L.used);
        L.take();
        cudaEvent_t* e = L.events.data() + ev_at[c];
        CK(cudaEventRecord(e[0], L.stream));
        if (p.bases) CK(cudaMemcpyAsync(S.seqs.p, seqs + p.base, p.bases, cudaMemcpyHostToDevice, L.stream));
        CK(cudaMemcpyAsync(S.offs.p, offsets + p.first, (p.n + 1) * 8, cudaMemcpyHostToDevice, L.stream));
        CK(cudaEventRecord(e[1], L.stream));
        launch_prepare(S, p, L.stream);
        launch_pass1(R, S, p, L.stream);
        CK(cudaEventRecord(e[2], L.stream));
        if (out.fmt == 2 && mems && p.max_len > 0xFFFFull) throw BadBatch("packed MEM records need reads shorter than 65536 bases");
        if (out.fmt == 1 && mems && p.max_len > 0xFFull) throw BadBatch("16-bit MEM records need reads shorter than 256 bases");
        launch_pass2(R, S, p, mode, L.stream, out.min_len);
        CK(cudaEventRecord(e[3], L.stream));
        if (mems) {
            launch_compact(S, p, out.fmt, L.stream);
            CK(cudaEventRecord(e[4], L.stream));
        } else {
            const bool packed = out.fmt == 7;      // moni_gpu_ms_packed: col[] = u32 pointer halves, u8 pointer bits 32..39, u16 lengths
            if (packed && p.max_len > 0xFFFFull) throw BadBatch("packed matching statistics need reads shorter than 65536 bases");
            launch_untranspose(S, p, L.stream, packed);
            CK(cudaEventRecord(e[4], L.stream));
            if (p.bases && packed) {
                DevBuf* const src[3] = {&S.out_ptr, &S.out_hi, &S.out_len};
                for (int q = 0; q < 3; ++q)
                    CK(cudaMemcpyAsync((char*)out.col[q] + p.base * out.width[q], src[q]->p, p.bases * out.width[q], cudaMemcpyDeviceToHost, L.stream));
            } else if (p.bases) {
                CK(cudaMemcpyAsync(pointers + p.base, S.out_ptr.p, p.bases * 8, cudaMemcpyDeviceToHost, L.stream));
                CK(cudaMemcpyAsync(lengths + p.base, S.out_len.p, p.bases * 8, cudaMemcpyDeviceToHost, L.stream));
            }
            CK(cudaEventRecord(e[5], L.stream));
        }
    };
    // mems only: learn the chunk's record count, then send offsets (shifted to shard-relative) and records home
    auto finish_mems = [&](size_t c) {
        const ChunkPlan& p = plans[c];
        Lane& L = R.lanes[c % kLanes_];
        Scratch& S = L.scratch;
        cudaEvent_t* e = L.events.data() + ev_at[c];
        uint64_t k = 0;
        CK(cudaMemcpyAsync(&k, S.moff.as<uint64_t>() + p.n, 8, cudaMemcpyDeviceToHost, L.stream));
        CK(cudaStreamSynchronize(L.stream));
        if (rec_total) {
            k_shift<<<(unsigned)((p.n + 1 + 255) / 256), 256, 0, L.stream>>>(S.moff.as<uint64_t>(), p.n + 1, rec_total);
            g_launches++;
        }
        const bool last = p.first + p.n == r1;
        DevBuf* const chunk_col[3] = {&S.mem_pos, &S.mem_len, &S.mem_ref};
        if (rv) {
            CK(cudaMemcpyAsync(R.shard_off.as<uint64_t>() + (p.first - r0), S.moff.p, (p.n + (last ? 1 : 0)) * 8, cudaMemcpyDeviceToDevice, L.stream));
            for (int q = 0; q < out.ncol && k; ++q)
                CK(cudaMemcpyAsync(shard_col[q]->as<char>() + rec_total * out.width[q], chunk_col[q]->p, k * out.width[q], cudaMemcpyDeviceToDevice, L.stream));
        } else {
            CK(cudaMemcpyAsync(rec_off + (p.first - r0), S.moff.p, (p.n + (last ? 1 : 0)) * 8, cudaMemcpyDeviceToHost, L.stream));
            for (int q = 0; q < out.ncol && k && rec_total + k <= rec_cap; ++q)
                CK(cudaMemcpyAsync((char*)out.col[q] + rec_total * out.width[q], chunk_col[q]->p, k * out.width[q], cudaMemcpyDeviceToHost, L.stream));
        }
        CK(cudaEventRecord(e[5], L.stream));
        rec_total += k;
    };

    // keep lanes - 1 chunks enqueued ahead of the one whose results are being sent home
    size_t fronted = 0;
    try {
        for (size_t c = 0; c < fronted || next_read < r1; ++c) {
            while (next_read < r1 && fronted < c + kLanes_) {
                plans.push_back(plan_chunk(offsets, next_read, r1, idx->opt.chunk_bases, idx->opt.chunk_min_reads));
                next_read += plans.back().n;
                enqueue_front(fronted++);
            }
            if (mems) finish_mems(c);
        }
    } catch (...) {
        // a bad offset found mid-batch (or a CUDA failure): nothing may still be writing the caller's buffers when we return
        for (Lane& L : R.lanes) cudaStreamSynchronize(L.stream);
        throw;
    }
    const size_t nc = plans.size();
    for (Lane& L : R.lanes) CK(cudaStreamSynchronize(L.stream));
    if (nc && getenv("MONI_GPU_TRACE")) {      // per-chunk timeline (ms since the first event): where the pipeline idles
        cudaEvent_t* e0 = R.lanes[0].events.data() + ev_at[0];
        for (size_t c = 0; c < nc; ++c) {
            cudaEvent_t* e = R.lanes[c % kLanes_].events.data() + ev_at[c];
            fprintf(stderr, "[moni-gpu trace] gpu %d chunk %zu lane %zu reads %llu:", R.dev, c, c % kLanes_, (unsigned long long)plans[c].n);
            for (int k = 0; k < kEv; ++k) fprintf(stderr, " %.3f", ev_ms(e0[0], e[k]));
            fprintf(stderr, "   (start h2d pass1 pass2 compact d2h)\n");
        }
    }
    for (size_t c = 0; c < nc; ++c) {
        cudaEvent_t* e = R.lanes[c % kLanes_].events.data() + ev_at[c];
        for (int k = 0; k < 5; ++k) R.t[k] += ev_ms(e[k], e[k + 1]);
        cudaEvent_t* e0 = R.lanes[0].events.data() + ev_at[0];
        R.t[5] = std::max<double>(R.t[5], ev_ms(e0[0], e[5]));
    }
    if (rv) {
        rv->publish(shard, (int64_t)rec_total);
        const uint64_t gbase = rv->base_of(shard);
        Lane& L = R.lanes[0];
        const size_t at = L.used;
        L.take();
        cudaEvent_t* e = L.events.data() + at;
        const uint64_t nr = r1 - r0;
        const bool last_shard = shard + 1 == (int)rv->total.size();
        CK(cudaEventRecord(e[0], L.stream));
        if (gbase && nr) {
            k_shift<<<(unsigned)((nr + 1 + 255) / 256), 256, 0, L.stream>>>(R.shard_off.as<uint64_t>(), nr + 1, gbase);
            g_launches++;
        }
        CK(cudaMemcpyAsync(rec_off + r0, R.shard_off.p, (nr + (last_shard ? 1 : 0)) * 8, cudaMemcpyDeviceToHost, L.stream));
        for (int q = 0; q < out.ncol && rec_total && gbase + rec_total <= rec_cap; ++q)
            CK(cudaMemcpyAsync((char*)out.col[q] + gbase * out.width[q], shard_col[q]->p, rec_total * out.width[q], cudaMemcpyDeviceToHost, L.stream));
        CK(cudaEventRecord(e[1], L.stream));
        CK(cudaStreamSynchronize(L.stream));
        const double ms = ev_ms(e[0], e[1]);
        R.t[4] += ms;
        R.t[5] += ms;
    }
    if (needed) *needed = rec_total;
}

void split_by_bases(const uint64_t* offsets, uint64_t n_reads, int parts, std::vector<uint64_t>& cut) {
    cut.assign(parts + 1, n_reads);
    cut[0] = 0;
    const uint64_t total = offsets[n_reads] - offsets[0];
    for (int g = 1; g < parts; ++g) {
        const uint64_t want = offsets[0] + total / parts * g;
        cut[g] = std::lower_bound(offsets, offsets + n_reads + 1, want) - offsets;
        if (cut[g] < cut[g - 1]) cut[g] = cut[g - 1];
        if (cut[g] > n_reads) cut[g] = n_reads;
    }
}

template <typename F>
int guarded(F&& f) {
    try {
        return f();
    } catch (const CudaError& e) {
        g_err = e.what();
        return MONI_GPU_ECUDA;
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

void collect_times(moni_gpu_index* idx) {
    for (int k = 0; k < 6; ++k) {
        idx->t[k] = 0;
        for (auto& r : idx->reps) idx->t[k] = std::max(idx->t[k], r->t[k]);
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
    else if (k == "lanes" && v >= 1 && v <= kMaxLanes) g_opt.lanes = (uint32_t)v;
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
            CK(cudaSetDevice(R->dev));
            if (idx->opt.l2_fetch) CK(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, idx->opt.l2_fetch));
            R->lanes.resize(idx->opt.lanes);
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
        const int G = (int)idx->reps.size();
        std::vector<uint64_t> cut;
        split_by_bases(offsets, n_reads, G, cut);
        std::vector<std::string> errs(G);
        std::vector<int> rc(G, 0);
        auto work = [&](int g) {
            rc[g] = guarded([&]() -> int {
                run_shard(idx, *idx->reps[g], 0, seqs, offsets, cut[g], cut[g + 1], pointers, lengths, out, nullptr);
                return 0;
            });
            if (rc[g]) errs[g] = g_err;
        };
        if (G == 1) work(0);
        else {
            std::vector<std::thread> th;
            for (int g = 0; g < G; ++g) th.emplace_back(work, g);
            for (auto& t : th) t.join();
        }
        collect_times(idx);
        for (int g = 0; g < G; ++g) if (rc[g]) { g_err = errs[g]; return rc[g]; }
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
        const int G = (int)idx->reps.size();
        std::vector<uint64_t> cut;
        split_by_bases(offsets, n_reads, G, cut);
        uint64_t needed = 0;
        if (G == 1) {
            run_shard(idx, *idx->reps[0], mode, seqs, offsets, 0, n_reads, nullptr, nullptr, out, &needed);
            collect_times(idx);
        } else {
            // every shard keeps its records on its GPU until the totals of the shards before it are known, then copies
            // them straight to their final place in the caller's arrays (input order, no host-side stitching)
            Rendezvous rv(G);
            std::vector<uint64_t> k(G, 0);
            std::vector<int> rc(G, 0);
            std::vector<std::string> errs(G);
            std::vector<std::thread> th;
            for (int g = 0; g < G; ++g)
                th.emplace_back([&, g]() {
                    bool published = false;
                    rc[g] = guarded([&]() -> int {
                        if (cut[g + 1] > cut[g]) {
                            run_shard(idx, *idx->reps[g], mode, seqs, offsets, cut[g], cut[g + 1], nullptr, nullptr, out, &k[g], &rv, g);
                            published = true;
                        }
                        return 0;
                    });
                    if (rc[g]) errs[g] = g_err;
                    if (!published) rv.publish(g, rc[g] ? -2 : 0);     // empty or failed shard: never leave the others waiting
                });
            for (auto& t : th) t.join();
            collect_times(idx);
            for (int g = 0; g < G; ++g) if (rc[g]) { g_err = errs[g]; return rc[g]; }
            for (int g = 0; g < G; ++g) needed += k[g];
            mem_offsets[n_reads] = needed;      // also covers trailing shards without reads
        }
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
    return mems_impl(idx, 1, seqs, offsets, n_reads, rec_out(mem_offsets, mem_pos, 8, mem_len, 8, nullptr, 0, mem_capacity, 8));
}

int moni_gpu_mems32(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads, uint64_t* mem_offsets,
                    uint32_t* mem_pos, uint32_t* mem_len, uint64_t mem_capacity) {
    return mems_impl(idx, 1, seqs, offsets, n_reads, rec_out(mem_offsets, mem_pos, 4, mem_len, 4, nullptr, 0, mem_capacity, 4));
}

int moni_gpu_mems_packed(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads, uint64_t* mem_offsets,
                         uint32_t* mem_rec, uint64_t mem_capacity) {
    return mems_impl(idx, 1, seqs, offsets, n_reads, rec_out(mem_offsets, mem_rec, 4, nullptr, 0, nullptr, 0, mem_capacity, 2));
}

int moni_gpu_mems_packed16(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads, uint64_t* mem_offsets,
                           uint16_t* mem_rec, uint64_t mem_capacity) {
    return mems_impl(idx, 1, seqs, offsets, n_reads, rec_out(mem_offsets, mem_rec, 2, nullptr, 0, nullptr, 0, mem_capacity, 1));
}

int moni_gpu_seeds(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads, uint64_t min_len,
                   uint64_t* seed_offsets, uint64_t* seed_ref, uint32_t* seed_qpos, uint32_t* seed_len, uint64_t seed_capacity) {
    const uint32_t ml = min_len > 0x7FFFFFFFull ? 0x7FFFFFFFu : (uint32_t)min_len;     // reads are shorter than 2^31
    return mems_impl(idx, 2, seqs, offsets, n_reads, rec_out(seed_offsets, seed_qpos, 4, seed_len, 4, seed_ref, 8, seed_capacity, 12, ml));
}

int moni_gpu_seed_occs(moni_gpu_index* idx, const uint64_t* seed_ref, const uint32_t* seed_len, uint64_t n_seeds,
                       uint64_t max_per_direction, uint64_t* occ_offsets, uint64_t* occs, uint64_t occ_capacity, int* truncated) {
    if (!idx || !occ_offsets || (n_seeds && (!seed_ref || !seed_len)) || n_seeds >= 0x7FFFFFF0ull) { g_err = "bad arguments"; return MONI_GPU_EINVAL; }
    if (truncated) *truncated = 0;
    if (n_seeds == 0) { occ_offsets[0] = 0; return 0; }
    if (!occs && occ_capacity) { g_err = "null output"; return MONI_GPU_EINVAL; }
    // the reference asserts len > 0 in find_MEM_above / find_MEM_below (include/aligner/seed_finder.hpp:171,203): a zero-length seed
    // would "match" every suffix and walk the whole suffix array; a cap beyond 2^32 positions per direction cannot be stored anyway
    if (max_per_direction > (1ull << 32)) { g_err = "max_per_direction must be at most 2^32"; return MONI_GPU_EINVAL; }
    for (uint64_t i = 0; i < n_seeds; ++i)
        if (seed_len[i] == 0) { g_err = "seed " + std::to_string(i) + " has length 0 (the reference requires len > 0)"; return MONI_GPU_EINVAL; }
    if (idx->hdr.phi_n == 0) { g_err = "this index has no Phi structures: load <prefix>.thrbv.full.ms (the `moni align` index)"; return MONI_GPU_EINVAL; }
    if (!(idx->hdr.root_ref & kLeafBit)) { g_err = "occurrence enumeration needs the text in HBM (leaf_len = 0 / text_budget)"; return MONI_GPU_EINVAL; }
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
        R.occ_off.reserve((n_seeds + 1) * 8);
        R.occ_flag.reserve(16);
        CK(cudaMemcpyAsync(R.occ_ref.p, seed_ref, n_seeds * 8, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(R.occ_len.p, seed_len, n_seeds * 4, cudaMemcpyHostToDevice, st));
        CK(cudaMemsetAsync(R.occ_flag.p, 0, 4, st));
        CK(cudaEventRecord(e[1], st));
        const unsigned blocks = (unsigned)((n_seeds + 1 + kBlock - 1) / kBlock);
        // staged single walk when 2 cap + 1 positions per seed fit a bounded share of the free HBM; else count, then fill
        uint64_t stride = 0;
        if (max_per_direction <= 4096) {
            const uint64_t want = (2 * max_per_direction + 1) * 8 * n_seeds;
            size_t free_b = 0, total_b = 0;
            CK(cudaMemGetInfo(&free_b, &total_b));
            if (want <= R.occ_stage.cap || want <= std::min<uint64_t>((uint64_t)free_b / 4, 8ull << 30)) stride = 2 * max_per_direction + 1;
        }
        if (stride) {
            R.occ_stage.reserve(stride * 8 * n_seeds);
            k_occ_walk_staged<<<blocks, kBlock, 0, st>>>(R.view, R.occ_ref.as<uint64_t>(), R.occ_len.as<uint32_t>(), n_seeds, max_per_direction,
                                                         stride, R.occ_off.as<uint64_t>(), R.occ_stage.as<uint64_t>(), R.occ_flag.as<uint32_t>());
        } else {
            k_occ_count<<<blocks, kBlock, 0, st>>>(R.view, R.occ_ref.as<uint64_t>(), R.occ_len.as<uint32_t>(), n_seeds, max_per_direction,
                                                   R.occ_off.as<uint64_t>(), R.occ_flag.as<uint32_t>());
        }
        g_launches++;
        size_t tmp = 0;
        CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp, R.occ_off.as<uint64_t>(), R.occ_off.as<uint64_t>(), (int)(n_seeds + 1), st));
        R.occ_cub.reserve(tmp);
        CK(cub::DeviceScan::ExclusiveSum(R.occ_cub.p, tmp, R.occ_off.as<uint64_t>(), R.occ_off.as<uint64_t>(), (int)(n_seeds + 1), st));
        g_launches += 2;
        CK(cudaEventRecord(e[2], st));
        uint32_t flag = 0;
        CK(cudaMemcpyAsync(occ_offsets, R.occ_off.p, (n_seeds + 1) * 8, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(&flag, R.occ_flag.p, 4, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        if (truncated) *truncated = flag ? 1 : 0;
        const uint64_t total = occ_offsets[n_seeds];
        if (total > occ_capacity) {
            g_err = "occ_capacity too small: " + std::to_string(total) + " positions needed";
            return MONI_GPU_ENOMEM;
        }
        R.occ_out.reserve(total * 8 + 16);
        if (stride)
            k_occ_pack<<<(unsigned)((n_seeds * 32 + kBlock - 1) / kBlock), kBlock, 0, st>>>(R.occ_stage.as<uint64_t>(), stride, R.occ_off.as<uint64_t>(),
                                                                                         n_seeds, R.occ_out.as<uint64_t>());
        else
            k_occ_fill<<<(unsigned)((n_seeds + kBlock - 1) / kBlock), kBlock, 0, st>>>(R.view, R.occ_ref.as<uint64_t>(), R.occ_len.as<uint32_t>(), n_seeds,
                                                                                   max_per_direction, R.occ_off.as<uint64_t>(), R.occ_out.as<uint64_t>());
        g_launches++;
        CK(cudaGetLastError());
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
        b->s.reserve(p, true, true);
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
        launch_pass2(R, b->s, b->plan, what == 2 ? 1 : 0, L.stream);
        CK(cudaEventRecord(e[2], L.stream));
        if (what == 2) launch_compact(b->s, b->plan, 8, L.stream);
        else launch_untranspose(b->s, b->plan, L.stream);
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
