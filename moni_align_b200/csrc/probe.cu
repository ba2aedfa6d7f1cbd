// Random-line probe: the measured HBM roofline this path is held against ("measured random-sector
// bandwidth of the box, obtained by a gather micro-benchmark ... ≥ 2^20 independent chains").
//
// Every chain is a dependent walk over a table far larger than L2: the next granule index is a hash of the words just
// loaded, which is the access pattern of an LF step over the move table (one 32-byte row at a random place, the next
// address unknown until it arrives).  B200 fetches a whole 128-byte line from DRAM for such a miss (profiles/
// r01_c_l2fetch_probe*.csv), so the figure of merit is random LINES per second.
//
// Landing shapes (template MODE), each with K independent chains per thread:
//   0  one LDG.256 per access, row lands in registers          (what k_ms_pointers does)
//   1  one LDG.128 per access (first half of the row), registers
//   2  one 16-byte cp.async per access, row half lands in shared memory (no landing registers: more chains per SM)
// Diagnostics only: nothing on the product path calls this.
#include <cuda_runtime.h>

#include <cstdint>
#include <string>

#include "../../include/moni_gpu.h"

namespace {

struct alignas(32) Granule { uint32_t w[8]; };

__device__ __forceinline__ uint32_t mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}
__device__ __forceinline__ uint32_t reduce32(uint32_t h, uint32_t n) { return __umulhi(h, n); }

__global__ void k_probe_fill(uint4* t, uint64_t n16) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n16; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t h = mix32((uint32_t)i * 2654435761u + 12345u);
        t[i] = make_uint4(h, mix32(h), (uint32_t)i, 0x9E3779B9u);
    }
}

template <int MODE, int K, int MINB>
__global__ void __launch_bounds__(256, MINB) k_probe_chain(const Granule* __restrict__ t, uint32_t n_gran, int iters, uint32_t* out) {
    extern __shared__ __align__(16) uint8_t land[];      // MODE 2: 16 bytes per (thread, chain)
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t g[K];
#pragma unroll
    for (int c = 0; c < K; ++c) g[c] = reduce32(mix32(tid * K + c + 1), n_gran);
    uint32_t acc = 0;
    if (MODE == 2) {
        const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(land) + threadIdx.x * 16u;
        // chain c lands at land[(c * 256 + thread) * 16]: conflict-free 16-byte rows across a warp
#pragma unroll
        for (int c = 0; c < K; ++c) {
            asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(sbase + c * 4096u), "l"(t + g[c]));
            asm volatile("cp.async.commit_group;");
        }
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int c = 0; c < K; ++c) {
                asm volatile("cp.async.wait_group %0;" ::"n"(K - 1));
                uint32_t x;
                asm volatile("ld.shared.u32 %0, [%1];" : "=r"(x) : "r"(sbase + c * 4096u));
                acc ^= x;
                g[c] = reduce32(mix32(x ^ g[c] ^ (uint32_t)it), n_gran);
                asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(sbase + c * 4096u), "l"(t + g[c]));
                asm volatile("cp.async.commit_group;");
            }
        }
        asm volatile("cp.async.wait_all;");
    } else {
        for (int it = 0; it < iters; ++it) {
            uint32_t v[K];
#pragma unroll
            for (int c = 0; c < K; ++c) {
                if (MODE == 0) {
                    uint32_t a, b, cc, d, e, f, gg, h;
                    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                                 : "=r"(a), "=r"(b), "=r"(cc), "=r"(d), "=r"(e), "=r"(f), "=r"(gg), "=r"(h) : "l"(t + g[c]));
                    v[c] = a ^ b ^ cc ^ d ^ e ^ f ^ gg ^ h;
                } else {
                    const uint4 x = __ldg(reinterpret_cast<const uint4*>(t + g[c]));
                    v[c] = x.x ^ x.y ^ x.z ^ x.w;
                }
            }
#pragma unroll
            for (int c = 0; c < K; ++c) {
                acc ^= v[c];
                g[c] = reduce32(mix32(v[c] ^ g[c] ^ (uint32_t)it), n_gran);
            }
        }
    }
    if (acc == 0x12345678u) out[0] = acc;
}

template <int MODE, int K, int MINB>
int run_probe(const Granule* t, uint32_t n_gran, int iters, int sms, uint32_t* out, double* lines_per_s, int* threads_per_sm) {
    auto kern = k_probe_chain<MODE, K, MINB>;
    const size_t smem = MODE == 2 ? (size_t)K * 4096 : 0;
    if (smem > 48 * 1024 && cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return -1;
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 256, smem) != cudaSuccess || per_sm < 1) return -1;
    if (per_sm > MINB) per_sm = MINB;       // MINB doubles as the number of resident 256-thread blocks per SM asked for
    const int blocks = sms * per_sm;       // exactly one resident wave, spread evenly over the SMs
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    kern<<<blocks, 256, smem>>>(t, n_gran, 8, out);
    float best = 1e30f;
    for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(a);
        kern<<<blocks, 256, smem>>>(t, n_gran, iters, out);
        cudaEventRecord(b);
        if (cudaEventSynchronize(b) != cudaSuccess) return -1;
        float ms = 0;
        cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
    }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    if (cudaGetLastError() != cudaSuccess) return -1;
    *lines_per_s = (double)blocks * 256.0 * K * iters / (best * 1e-3);
    *threads_per_sm = per_sm * 256;
    return 0;
}

Granule* g_table = nullptr;
uint64_t g_table_cap = 0;
int g_table_dev = -1;
uint32_t* g_out = nullptr;

}  // namespace

extern "C" int moni_gpu_probe_random_lines(int gpu, uint64_t table_bytes, int shape, int chains_per_thread, int min_blocks_per_sm, int iters,
                                           double* lines_per_s, int* threads_per_sm) {
    if (!lines_per_s || table_bytes < (1ull << 20) || table_bytes / 32 >= 0xFFFFFFFFull || iters < 1) return MONI_GPU_EINVAL;
    if (cudaSetDevice(gpu) != cudaSuccess) { cudaGetLastError(); return MONI_GPU_ECUDA; }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, gpu) != cudaSuccess) return MONI_GPU_ECUDA;
    Granule*& table = g_table;                 // kept between calls of one process (a sweep re-uses it)
    uint64_t& table_cap = g_table_cap;
    int& table_dev = g_table_dev;
    uint32_t*& out = g_out;
    if (table_dev != gpu || table_cap < table_bytes) {
        if (table) cudaFree(table);
        if (out) cudaFree(out);
        table = nullptr;
        out = nullptr;
        table_cap = 0;
        if (cudaMalloc(&table, table_bytes) != cudaSuccess || cudaMalloc(&out, 4) != cudaSuccess) { cudaGetLastError(); return MONI_GPU_ENOMEM; }
        k_probe_fill<<<prop.multiProcessorCount * 8, 256>>>((uint4*)table, table_bytes / 16);
        if (cudaDeviceSynchronize() != cudaSuccess) return MONI_GPU_ECUDA;
        table_cap = table_bytes;
        table_dev = gpu;
    }
    const uint32_t n_gran = (uint32_t)(table_bytes / 32);
    const int sms = prop.multiProcessorCount;
    int tps = 0;
    int rc = -1;
#define MONI_PROBE(M, KK, B) \
    if (shape == M && chains_per_thread == KK && min_blocks_per_sm == B) rc = run_probe<M, KK, B>(table, n_gran, iters, sms, out, lines_per_s, &tps);
    MONI_PROBE(0, 1, 4) MONI_PROBE(0, 1, 6) MONI_PROBE(0, 1, 8) MONI_PROBE(0, 2, 4) MONI_PROBE(0, 2, 6) MONI_PROBE(0, 4, 4) MONI_PROBE(0, 4, 3)
    MONI_PROBE(1, 1, 8) MONI_PROBE(1, 2, 8) MONI_PROBE(1, 2, 6) MONI_PROBE(1, 4, 8) MONI_PROBE(1, 4, 6) MONI_PROBE(1, 4, 4) MONI_PROBE(1, 8, 4) MONI_PROBE(1, 8, 6)
    MONI_PROBE(2, 2, 8) MONI_PROBE(2, 4, 8) MONI_PROBE(2, 6, 8) MONI_PROBE(2, 8, 8) MONI_PROBE(2, 8, 6) MONI_PROBE(2, 8, 4) MONI_PROBE(2, 12, 4) MONI_PROBE(2, 16, 4)
    MONI_PROBE(2, 12, 6)
#undef MONI_PROBE
    if (threads_per_sm) *threads_per_sm = tps;
    return rc == 0 ? 0 : MONI_GPU_EINVAL;
}

extern "C" void moni_gpu_probe_release(void) {
    if (g_table_dev < 0) return;
    cudaSetDevice(g_table_dev);
    if (g_table) cudaFree(g_table);
    if (g_out) cudaFree(g_out);
    g_table = nullptr;
    g_out = nullptr;
    g_table_cap = 0;
    g_table_dev = -1;
}
