// libmoni_gpu.so: CUDA kernels (sm_100a) and the C ABI of include/moni_gpu.h.
//
// Kernel mapping (DESIGN.md "Kernels"): one read per thread.  Every backward step of pass 1 is one
// dependent random 32-byte sector (a MoveRow), so throughput comes from the ~10^5 independent
// reads in flight per GPU, not from cooperation inside a read; the per-read arithmetic lives in
// ms_core.cuh.  No tensor cores, no CPU fallback: without a device every entry point fails.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include <cub/device/device_scan.cuh>

#include "../../include/moni_gpu.h"
#include "device_image.hpp"
#include "index_files.hpp"
#include "ms_core.cuh"

using namespace moni;

// ================================================================================ kernels
namespace {

constexpr int kBlock = 128;

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

// Pass 1: MS pointers.  ptr[] uses the layout of seqs[] (chunk-relative offsets = offs[r] - base).
template <bool COUNT>
__global__ void __launch_bounds__(kBlock) k_ms_pointers(ImageView v, const uint8_t* __restrict__ seqs,
                                                        const uint64_t* __restrict__ offs, uint64_t base,
                                                        uint32_t n_reads, uint64_t* __restrict__ ptr, Counters* g) {
    __shared__ uint8_t tcode[256];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) tcode[i] = v.sym->tcode[i];
    __syncthreads();
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_reads) return;
    const uint64_t o = offs[r] - base, m = offs[r + 1] - offs[r];
    ScanState s;
    scan_init(v, s);
    Counters c = {0, 0, 0, 0, 0, 0, 0, 0};
    for (uint64_t j = m; j-- > 0;) {
        scan_step(v, tcode, seqs[o + j], s, COUNT ? &c : nullptr);
        ptr[o + j] = s.sample;
    }
    if (COUNT) { c.bases = m; flush_counters(c, g); }
}

// Pass 2.  MEMS == false: len[] (u64) per position.  MEMS == true: the read's MEM records are packed
// in place over its own pointers (record k is produced at query offset i >= k, after ptr[i] was
// consumed) and counts[r] receives their number.
template <bool MEMS, bool COUNT>
__global__ void __launch_bounds__(kBlock) k_ms_lengths(ImageView v, const uint8_t* __restrict__ seqs,
                                                       const uint64_t* __restrict__ offs, uint64_t base,
                                                       uint32_t n_reads, uint64_t* ptr, uint64_t* len,
                                                       uint32_t* counts, Counters* g) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_reads) return;
    const uint64_t o = offs[r] - base, m = offs[r + 1] - offs[r];
    Counters c = {0, 0, 0, 0, 0, 0, 0, 0};
    const uint32_t k = lengths_read<uint64_t>(v, seqs + o, m, ptr + o, MEMS ? nullptr : len + o, MEMS ? ptr + o : nullptr,
                                              COUNT ? &c : nullptr);
    if (MEMS) counts[r] = k;
    if (COUNT) flush_counters(c, g);
}

// MEM compaction: one warp per read copies its packed records to their final place.
__global__ void __launch_bounds__(kBlock) k_mem_gather(const uint64_t* __restrict__ packed, const uint64_t* __restrict__ offs,
                                                       uint64_t base, uint32_t n_reads, const uint64_t* __restrict__ moff,
                                                       uint64_t* __restrict__ mem_pos, uint64_t* __restrict__ mem_len) {
    const uint32_t r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (r >= n_reads) return;
    const uint64_t o = offs[r] - base, a = moff[r], cnt = moff[r + 1] - a;
    for (uint64_t k = lane; k < cnt; k += 32) {
        const uint64_t rec = packed[o + k];
        mem_pos[a + k] = rec >> 32;
        mem_len[a + k] = rec & 0xFFFFFFFFull;
    }
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
    uint32_t leaf_len = 128, max_scan = 8;
    bool verbose = false, count = false;
    uint64_t chunk_bases = 1ull << 26;
};
Options g_opt;
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

struct Scratch {   // device buffers of one batch / one chunk
    DevBuf seqs, offs, ptr, len, counts, moff, mem_pos, mem_len, cub;
    void release() {
        seqs.release(); offs.release(); ptr.release(); len.release(); counts.release();
        moff.release(); mem_pos.release(); mem_len.release(); cub.release();
    }
};

struct Replica {
    int dev = 0;
    cudaStream_t stream = nullptr;
    DevBuf rows, row_pos, crun_row, slots, nodes, pool, sym, counters;
    ImageView view;
    Scratch scratch;
    cudaEvent_t ev[8] = {nullptr};
    std::mutex mu;
    double t[6] = {0, 0, 0, 0, 0, 0};
    ~Replica() {
        cudaSetDevice(dev);
        rows.release(); row_pos.release(); crun_row.release(); slots.release(); nodes.release(); pool.release();
        sym.release(); counters.release();
        scratch.release();
        for (auto& e : ev) if (e) cudaEventDestroy(e);
        if (stream) cudaStreamDestroy(stream);
    }
};

template <typename T>
void upload(DevBuf& b, const std::vector<T>& v, cudaStream_t s) {
    b.reserve(std::max<size_t>(v.size() * sizeof(T), 16));
    if (!v.empty()) CK(cudaMemcpyAsync(b.p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, s));
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
    uint64_t n_reads = 0, total = 0, n_mems = 0;
    std::vector<uint64_t> offsets;   // host copy (for sizes)
    int last = 0;
};

namespace {

// ---- one chunk on one replica (device pointers already filled) ---------------------------------
struct ChunkTimes { float pass1 = 0, pass2 = 0, compact = 0; };

void launch_pass1(Replica& R, Scratch& S, uint64_t base, uint32_t n) {
    if (!n) return;
    const dim3 grid((n + kBlock - 1) / kBlock);
    if (g_count.load())
        k_ms_pointers<true><<<grid, kBlock, 0, R.stream>>>(R.view, S.seqs.as<uint8_t>(), S.offs.as<uint64_t>(), base, n,
                                                           S.ptr.as<uint64_t>(), R.counters.as<Counters>());
    else
        k_ms_pointers<false><<<grid, kBlock, 0, R.stream>>>(R.view, S.seqs.as<uint8_t>(), S.offs.as<uint64_t>(), base, n,
                                                            S.ptr.as<uint64_t>(), nullptr);
    g_launches++;
    CK(cudaGetLastError());
}

void launch_pass2(Replica& R, Scratch& S, uint64_t base, uint32_t n, bool mems) {
    if (!n) return;
    const dim3 grid((n + kBlock - 1) / kBlock);
    Counters* g = R.counters.as<Counters>();
    const uint8_t* sq = S.seqs.as<uint8_t>();
    const uint64_t* of = S.offs.as<uint64_t>();
    uint64_t* p = S.ptr.as<uint64_t>();
    if (mems) {
        if (g_count.load()) k_ms_lengths<true, true><<<grid, kBlock, 0, R.stream>>>(R.view, sq, of, base, n, p, nullptr, S.counts.as<uint32_t>(), g);
        else k_ms_lengths<true, false><<<grid, kBlock, 0, R.stream>>>(R.view, sq, of, base, n, p, nullptr, S.counts.as<uint32_t>(), nullptr);
    } else {
        if (g_count.load()) k_ms_lengths<false, true><<<grid, kBlock, 0, R.stream>>>(R.view, sq, of, base, n, p, S.len.as<uint64_t>(), nullptr, g);
        else k_ms_lengths<false, false><<<grid, kBlock, 0, R.stream>>>(R.view, sq, of, base, n, p, S.len.as<uint64_t>(), nullptr, nullptr);
    }
    g_launches++;
    CK(cudaGetLastError());
}

// counts -> exclusive offsets (n + 1 entries) -> gather.  Buffers mem_pos/mem_len must hold `total_bases` entries.
void launch_compact(Replica& R, Scratch& S, uint64_t base, uint32_t n) {
    if (!n) return;
    uint64_t* wide = S.moff.as<uint64_t>() + (n + 1);          // second half of moff: widened counts
    k_widen_counts<<<(n + 1 + 255) / 256, 256, 0, R.stream>>>(S.counts.as<uint32_t>(), wide, n);
    g_launches++;
    size_t tmp = 0;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp, wide, S.moff.as<uint64_t>(), (int)(n + 1), R.stream));
    S.cub.reserve(tmp);
    CK(cub::DeviceScan::ExclusiveSum(S.cub.p, tmp, wide, S.moff.as<uint64_t>(), (int)(n + 1), R.stream));
    g_launches++;
    const uint64_t threads = (uint64_t)n * 32;
    k_mem_gather<<<(unsigned)((threads + kBlock - 1) / kBlock), kBlock, 0, R.stream>>>(
        S.ptr.as<uint64_t>(), S.offs.as<uint64_t>(), base, n, S.moff.as<uint64_t>(), S.mem_pos.as<uint64_t>(), S.mem_len.as<uint64_t>());
    g_launches++;
    CK(cudaGetLastError());
}

void reserve_scratch(Scratch& S, uint64_t n, uint64_t bases, bool mems) {
    S.seqs.reserve(bases + 16);
    S.offs.reserve((n + 1) * 8);
    S.ptr.reserve(bases * 8 + 16);
    if (mems) {
        S.counts.reserve((n + 1) * 4);
        S.moff.reserve((n + 1) * 16);
        S.mem_pos.reserve(bases * 8 + 16);
        S.mem_len.reserve(bases * 8 + 16);
    } else {
        S.len.reserve(bases * 8 + 16);
    }
}

float ev_ms(cudaEvent_t a, cudaEvent_t b) {
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, a, b));
    return ms;
}

// Processes reads [r0, r1) of a host batch on one replica, chunk by chunk.  For mems, records go to
// rec_pos/rec_len (contiguous from index 0, capacity rec_cap) and per-read record offsets (relative
// to this shard) to rec_off[0 .. r1-r0]; *needed receives the number of records of the shard.
void run_shard(moni_gpu_index* idx, Replica& R, bool mems, const char* seqs, const uint64_t* offsets, uint64_t r0, uint64_t r1,
               uint64_t* pointers, uint64_t* lengths, uint64_t* rec_off, uint64_t* rec_pos, uint64_t* rec_len, uint64_t rec_cap,
               uint64_t* needed) {
    std::lock_guard<std::mutex> lock(R.mu);
    CK(cudaSetDevice(R.dev));
    for (double& x : R.t) x = 0;
    Scratch& S = R.scratch;
    uint64_t rec_total = 0;
    if (mems) rec_off[0] = 0;
    std::vector<uint64_t> moff_host;
    const uint64_t chunk_bases = idx->opt.chunk_bases;
    for (uint64_t a = r0; a < r1;) {
        uint64_t b = a + 1;
        while (b < r1 && offsets[b + 1] - offsets[a] <= chunk_bases && b - a < 0x7FFFFFF0ull) ++b;
        const uint64_t n = b - a, base = offsets[a], bases = offsets[b] - base;
        reserve_scratch(S, n, bases, mems);
        CK(cudaEventRecord(R.ev[0], R.stream));
        if (bases) CK(cudaMemcpyAsync(S.seqs.p, seqs + base, bases, cudaMemcpyHostToDevice, R.stream));
        CK(cudaMemcpyAsync(S.offs.p, offsets + a, (n + 1) * 8, cudaMemcpyHostToDevice, R.stream));
        CK(cudaEventRecord(R.ev[1], R.stream));
        launch_pass1(R, S, base, (uint32_t)n);
        CK(cudaEventRecord(R.ev[2], R.stream));
        launch_pass2(R, S, base, (uint32_t)n, mems);
        CK(cudaEventRecord(R.ev[3], R.stream));
        if (mems) {
            launch_compact(R, S, base, (uint32_t)n);
            CK(cudaEventRecord(R.ev[4], R.stream));
            moff_host.resize(n + 1);
            CK(cudaMemcpyAsync(moff_host.data(), S.moff.p, (n + 1) * 8, cudaMemcpyDeviceToHost, R.stream));
            CK(cudaStreamSynchronize(R.stream));
            const uint64_t k = moff_host[n];
            for (uint64_t i = 0; i <= n; ++i) rec_off[a - r0 + i] = rec_total + moff_host[i];
            if (rec_total + k <= rec_cap && k) {
                CK(cudaMemcpyAsync(rec_pos + rec_total, S.mem_pos.p, k * 8, cudaMemcpyDeviceToHost, R.stream));
                CK(cudaMemcpyAsync(rec_len + rec_total, S.mem_len.p, k * 8, cudaMemcpyDeviceToHost, R.stream));
            }
            rec_total += k;
        } else {
            CK(cudaEventRecord(R.ev[4], R.stream));
            if (bases) {
                CK(cudaMemcpyAsync(pointers + base, S.ptr.p, bases * 8, cudaMemcpyDeviceToHost, R.stream));
                CK(cudaMemcpyAsync(lengths + base, S.len.p, bases * 8, cudaMemcpyDeviceToHost, R.stream));
            }
        }
        CK(cudaEventRecord(R.ev[5], R.stream));
        CK(cudaStreamSynchronize(R.stream));
        R.t[0] += ev_ms(R.ev[0], R.ev[1]);
        R.t[1] += ev_ms(R.ev[1], R.ev[2]);
        R.t[2] += ev_ms(R.ev[2], R.ev[3]);
        R.t[3] += ev_ms(R.ev[3], R.ev[4]);
        R.t[4] += ev_ms(R.ev[4], R.ev[5]);
        R.t[5] += ev_ms(R.ev[0], R.ev[5]);
        a = b;
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

int check_batch(const uint64_t* offsets, uint64_t n_reads) {
    for (uint64_t r = 0; r < n_reads; ++r) {
        if (offsets[r + 1] < offsets[r]) { g_err = "offsets must be non-decreasing"; return MONI_GPU_EINVAL; }
        if (offsets[r + 1] - offsets[r] >= 0xFFFFFFFFull) { g_err = "reads of 2^32 bases or more are not supported"; return MONI_GPU_EINVAL; }
    }
    return 0;
}

template <typename F>
int guarded(F&& f) {
    try {
        return f();
    } catch (const CudaError& e) {
        g_err = e.what();
        return MONI_GPU_ECUDA;
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
const char* moni_gpu_version(void) { return "moni-gpu 0.1 (sm_100a)"; }
uint64_t moni_gpu_launch_count(void) { return g_launches.load(); }

int moni_gpu_set_option(const char* key, const char* value) {
    if (!key || !value) { g_err = "null option"; return MONI_GPU_EINVAL; }
    std::lock_guard<std::mutex> lock(g_opt_mu);
    const std::string k(key);
    const long long v = atoll(value);
    if (k == "leaf_len" && v >= 1 && v <= (1 << 20)) g_opt.leaf_len = (uint32_t)v;
    else if (k == "max_scan" && v >= 2 && v <= 4096) g_opt.max_scan = (uint32_t)v;
    else if (k == "verbose") g_opt.verbose = v != 0;
    else if (k == "count") { g_opt.count = v != 0; g_count.store(v != 0); }
    else if (k == "chunk_bases" && v >= 1) g_opt.chunk_bases = (uint64_t)v;
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
void moni_gpu_host_free(void* p) {
    if (p) cudaFreeHost(p);
}

int moni_gpu_index_load(const char* prefix, const int* gpus, int n_gpus, moni_gpu_index** out) {
    if (!prefix || !out || n_gpus < 1) { g_err = "bad arguments"; return MONI_GPU_EINVAL; }
    *out = nullptr;
    return guarded([&]() -> int {
        int ndev = 0;
        CK(cudaGetDeviceCount(&ndev));
        if (ndev < 1) throw CudaError("no CUDA device: libmoni_gpu has no CPU path");
        std::vector<int> devs(n_gpus);
        for (int g = 0; g < n_gpus; ++g) {
            devs[g] = gpus ? gpus[g] : g;
            if (devs[g] < 0 || devs[g] >= ndev) { g_err = "GPU ordinal out of range"; return MONI_GPU_EINVAL; }
        }
        std::unique_ptr<moni_gpu_index> idx(new moni_gpu_index);
        { std::lock_guard<std::mutex> lock(g_opt_mu); idx->opt = g_opt; }
        HostImage img;
        {
            ExportOptions eo;
            eo.leaf_len = idx->opt.leaf_len;
            eo.max_scan = idx->opt.max_scan;
            eo.verbose = idx->opt.verbose;
            FlatRlbwt bwt;
            read_thrbv_ms(std::string(prefix) + ".thrbv.ms", bwt);
            build_move_table(bwt, eo, img);
        }
        {
            ExportOptions eo;
            eo.leaf_len = idx->opt.leaf_len;
            eo.verbose = idx->opt.verbose;
            FlatSlp slp;
            read_plain_slp(std::string(prefix) + ".plain.slp", slp);
            build_grammar_table(slp, eo, img);
        }
        idx->hdr = img.hdr;
        idx->device_bytes = img.bytes() + sizeof(SymbolTables);
        for (int g = 0; g < n_gpus; ++g) {
            std::unique_ptr<Replica> R(new Replica);
            R->dev = devs[g];
            CK(cudaSetDevice(R->dev));
            CK(cudaStreamCreateWithFlags(&R->stream, cudaStreamNonBlocking));
            for (auto& e : R->ev) CK(cudaEventCreate(&e));
            upload(R->rows, img.rows, R->stream);
            upload(R->row_pos, img.row_pos, R->stream);
            upload(R->crun_row, img.crun_row, R->stream);
            upload(R->slots, img.slots, R->stream);
            upload(R->nodes, img.nodes, R->stream);
            upload(R->pool, img.pool, R->stream);
            R->sym.reserve(sizeof(SymbolTables));
            CK(cudaMemcpyAsync(R->sym.p, &img.sym, sizeof(SymbolTables), cudaMemcpyHostToDevice, R->stream));
            R->counters.reserve(sizeof(Counters));
            CK(cudaMemsetAsync(R->counters.p, 0, sizeof(Counters), R->stream));
            CK(cudaStreamSynchronize(R->stream));
            ImageView& v = R->view;
            v.rows = R->rows.as<MoveRow>();
            v.row_pos = R->row_pos.as<uint64_t>();
            v.crun_row = R->crun_row.as<uint32_t>();
            v.slots = R->slots.as<GapSlot>();
            v.nodes = R->nodes.as<SlpNode>();
            v.pool = R->pool.as<uint8_t>();
            v.sym = R->sym.as<SymbolTables>();
            v.hdr = img.hdr;
            idx->reps.push_back(std::move(R));
        }
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

int moni_gpu_ms(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads, uint64_t* pointers, uint64_t* lengths) {
    if (!idx || !offsets || (!seqs && n_reads && offsets[n_reads] != offsets[0])) { g_err = "bad arguments"; return MONI_GPU_EINVAL; }
    if (n_reads == 0) return 0;
    if (!pointers || !lengths) { g_err = "null output"; return MONI_GPU_EINVAL; }
    if (int e = check_batch(offsets, n_reads)) return e;
    return guarded([&]() -> int {
        const int G = (int)idx->reps.size();
        std::vector<uint64_t> cut;
        split_by_bases(offsets, n_reads, G, cut);
        std::vector<std::string> errs(G);
        std::vector<int> rc(G, 0);
        auto work = [&](int g) {
            rc[g] = guarded([&]() -> int {
                run_shard(idx, *idx->reps[g], false, seqs, offsets, cut[g], cut[g + 1], pointers, lengths, nullptr, nullptr, nullptr, 0, nullptr);
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

int moni_gpu_mems(moni_gpu_index* idx, const char* seqs, const uint64_t* offsets, uint64_t n_reads, uint64_t* mem_offsets,
                  uint64_t* mem_pos, uint64_t* mem_len, uint64_t mem_capacity) {
    if (!idx || !offsets || !mem_offsets) { g_err = "bad arguments"; return MONI_GPU_EINVAL; }
    if (n_reads == 0) { mem_offsets[0] = 0; return 0; }
    if (!seqs && offsets[n_reads] != offsets[0]) { g_err = "null seqs"; return MONI_GPU_EINVAL; }
    if ((!mem_pos || !mem_len) && mem_capacity) { g_err = "null output"; return MONI_GPU_EINVAL; }
    if (int e = check_batch(offsets, n_reads)) return e;
    return guarded([&]() -> int {
        const int G = (int)idx->reps.size();
        std::vector<uint64_t> cut;
        split_by_bases(offsets, n_reads, G, cut);
        uint64_t needed = 0;
        if (G == 1) {
            run_shard(idx, *idx->reps[0], true, seqs, offsets, 0, n_reads, nullptr, nullptr, mem_offsets, mem_pos, mem_len, mem_capacity, &needed);
            collect_times(idx);
        } else {
            // every shard fills private buffers; they are stitched together in input order afterwards
            struct Part { std::vector<uint64_t> off, pos, len; uint64_t k = 0; int rc = 0; std::string err; };
            std::vector<Part> parts(G);
            std::vector<std::thread> th;
            for (int g = 0; g < G; ++g)
                th.emplace_back([&, g]() {
                    Part& P = parts[g];
                    P.rc = guarded([&]() -> int {
                        const uint64_t nr = cut[g + 1] - cut[g], cap = offsets[cut[g + 1]] - offsets[cut[g]];
                        P.off.resize(nr + 1);
                        P.pos.resize(cap + 1);
                        P.len.resize(cap + 1);
                        run_shard(idx, *idx->reps[g], true, seqs, offsets, cut[g], cut[g + 1], nullptr, nullptr, P.off.data(), P.pos.data(),
                                  P.len.data(), cap, &P.k);
                        return 0;
                    });
                    if (P.rc) P.err = g_err;
                });
            for (auto& t : th) t.join();
            collect_times(idx);
            for (int g = 0; g < G; ++g) if (parts[g].rc) { g_err = parts[g].err; return parts[g].rc; }
            for (int g = 0; g < G; ++g) {
                const Part& P = parts[g];
                const uint64_t nr = cut[g + 1] - cut[g];
                for (uint64_t i = 0; i <= nr; ++i) mem_offsets[cut[g] + i] = needed + P.off[i];
                if (needed + P.k <= mem_capacity && P.k) {
                    memcpy(mem_pos + needed, P.pos.data(), P.k * 8);
                    memcpy(mem_len + needed, P.len.data(), P.k * 8);
                }
                needed += P.k;
            }
        }
        if (needed > mem_capacity) {
            g_err = "mem_capacity too small: " + std::to_string(needed) + " records needed";
            return MONI_GPU_ENOMEM;
        }
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
            CK(cudaMemcpyAsync(&h, R->counters.p, sizeof(h), cudaMemcpyDeviceToHost, R->stream));
            if (reset) CK(cudaMemsetAsync(R->counters.p, 0, sizeof(Counters), R->stream));
            CK(cudaStreamSynchronize(R->stream));
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
    if (!idx || !offsets || !out || gpu_slot < 0 || gpu_slot >= (int)idx->reps.size() || n_reads == 0 || n_reads >= 0x7FFFFFF0ull) {
        g_err = "bad arguments";
        return MONI_GPU_EINVAL;
    }
    if (int e = check_batch(offsets, n_reads)) return e;
    *out = nullptr;
    return guarded([&]() -> int {
        std::unique_ptr<moni_gpu_batch> b(new moni_gpu_batch);
        b->idx = idx;
        b->rep = idx->reps[gpu_slot].get();
        Replica& R = *b->rep;
        std::lock_guard<std::mutex> lock(R.mu);
        CK(cudaSetDevice(R.dev));
        b->n_reads = n_reads;
        b->offsets.assign(offsets, offsets + n_reads + 1);
        b->total = offsets[n_reads] - offsets[0];
        reserve_scratch(b->s, n_reads, b->total, true);
        b->s.len.reserve(b->total * 8 + 16);
        if (b->total) CK(cudaMemcpyAsync(b->s.seqs.p, seqs + offsets[0], b->total, cudaMemcpyHostToDevice, R.stream));
        CK(cudaMemcpyAsync(b->s.offs.p, offsets, (n_reads + 1) * 8, cudaMemcpyHostToDevice, R.stream));
        CK(cudaStreamSynchronize(R.stream));
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
        const uint64_t base = b->offsets[0];
        CK(cudaEventRecord(R.ev[0], R.stream));
        launch_pass1(R, b->s, base, (uint32_t)b->n_reads);
        CK(cudaEventRecord(R.ev[1], R.stream));
        launch_pass2(R, b->s, base, (uint32_t)b->n_reads, what == 2);
        CK(cudaEventRecord(R.ev[2], R.stream));
        if (what == 2) launch_compact(R, b->s, base, (uint32_t)b->n_reads);
        CK(cudaEventRecord(R.ev[3], R.stream));
        CK(cudaStreamSynchronize(R.stream));
        b->last = what;
        if (kernel_ms) {
            kernel_ms[0] = ev_ms(R.ev[0], R.ev[1]);
            kernel_ms[1] = ev_ms(R.ev[1], R.ev[2]);
            kernel_ms[2] = ev_ms(R.ev[2], R.ev[3]);
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
        if (b->total) {
            CK(cudaMemcpyAsync(pointers, b->s.ptr.p, b->total * 8, cudaMemcpyDeviceToHost, R.stream));
            CK(cudaMemcpyAsync(lengths, b->s.len.p, b->total * 8, cudaMemcpyDeviceToHost, R.stream));
        }
        CK(cudaStreamSynchronize(R.stream));
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
        CK(cudaMemcpyAsync(mem_offsets, b->s.moff.p, (b->n_reads + 1) * 8, cudaMemcpyDeviceToHost, R.stream));
        CK(cudaStreamSynchronize(R.stream));
        const uint64_t k = mem_offsets[b->n_reads];
        if (k > mem_capacity) { g_err = "mem_capacity too small: " + std::to_string(k) + " records needed"; return MONI_GPU_ENOMEM; }
        if (k) {
            CK(cudaMemcpyAsync(mem_pos, b->s.mem_pos.p, k * 8, cudaMemcpyDeviceToHost, R.stream));
            CK(cudaMemcpyAsync(mem_len, b->s.mem_len.p, k * 8, cudaMemcpyDeviceToHost, R.stream));
        }
        CK(cudaStreamSynchronize(R.stream));
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
