// gather_bench: measured random-access HBM roofline of this box (SURVEY.md 8(d) denominator).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_build/gather_bench tools/gather_bench.cu
//   tools/_build/gather_bench [max_table_GiB=16] [l2_fetch_granularity] > profiles/random_access_peak.json
//
// Every thread follows dependent chains (next address = hash(loaded data)): the shape of an LF walk.
// Granule G in {32, 64, 128} bytes, aligned.  Load shapes:
//   "2x128"  one thread loads its 32-byte granule with two 16-byte loads (LDG.128 x2)
//   "1x256"  one thread loads its 32-byte granule with one 32-byte load (sm_100 LDG.256)
//   "coop"   G/32 adjacent lanes share one chain and each load 32 bytes of the granule (one LDG.256
//            warp instruction fetches 32*32/G granules), so L1 sees whole-granule requests
// `chains` = independent chains per thread (software pipelining, MLP per thread).
// Reported GB/s count only granule bytes requested (G x accesses / time); table sizes bracket the index
// image (about 1-3 GB) and a table far beyond L2 / TLB reach.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x)                                                                                   \
    do {                                                                                        \
        cudaError_t e = (x);                                                                    \
        if (e != cudaSuccess) {                                                                 \
            fprintf(stderr, "%s: %s (%s:%d)\n", #x, cudaGetErrorString(e), __FILE__, __LINE__); \
            exit(1);                                                                            \
        }                                                                                       \
    } while (0)

struct alignas(32) Sector { uint32_t w[8]; };

__device__ __forceinline__ uint64_t mix(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
    return x;
}
// uniform in [0, n) without a 64-bit division
__device__ __forceinline__ uint64_t reduce(uint64_t h, uint64_t n) { return __umul64hi(h, n); }

__global__ void k_fill(uint4* t, uint64_t n16) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n16; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t h = mix(i);
        t[i] = make_uint4((uint32_t)h, (uint32_t)(h >> 32), (uint32_t)i, 0x9E3779B9u);
    }
}

__device__ __forceinline__ uint32_t ld256(const Sector* p) {
    uint32_t a, b, c, d, e, f, g, h;
    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(a), "=r"(b), "=r"(c), "=r"(d), "=r"(e), "=r"(f), "=r"(g), "=r"(h) : "l"(p));
    return a ^ b ^ c ^ d ^ e ^ f ^ g ^ h;
}
__device__ __forceinline__ uint32_t ld2x128(const Sector* p) {
    const uint4 x = __ldg(reinterpret_cast<const uint4*>(p)), y = __ldg(reinterpret_cast<const uint4*>(p) + 1);
    return x.x ^ x.y ^ x.z ^ x.w ^ y.x ^ y.y ^ y.z ^ y.w;
}

// MODE 0: 2x128 per thread, MODE 1: 1x256 per thread, MODE 2: coop (LANES lanes per granule, 1x256 each)
template <int MODE, int LANES, int CHAINS, int BLOCK = 256>
__global__ void __launch_bounds__(BLOCK) k_chain(const Sector* __restrict__ t, uint64_t n_gran, int iters, uint32_t* out) {
    const uint64_t tid = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    const uint32_t sub = threadIdx.x % LANES;
    uint64_t g[CHAINS];
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) g[c] = reduce(mix((tid / LANES) * CHAINS + c + 1), n_gran);
    uint32_t acc = 0;
    for (int it = 0; it < iters; ++it) {
        uint32_t v[CHAINS];
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) {
            const Sector* p = t + g[c] * LANES + sub;
            v[c] = MODE == 0 ? ld2x128(p) : ld256(p);
        }
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) {
            uint32_t x = v[c];
            if (LANES > 1) {   // the whole granule decides the next address
#pragma unroll
                for (int o = 1; o < LANES; o <<= 1) x ^= __shfl_xor_sync(0xFFFFFFFFu, x, o);
            }
            acc ^= x;
            g[c] = reduce(mix(((uint64_t)x << 20) ^ g[c] ^ (uint64_t)it), n_gran);
        }
    }
    if (acc == 0x12345678u) out[0] = acc;
}

template <typename F>
static double time_ms(F&& launch, int reps) {
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    launch();
    CK(cudaDeviceSynchronize());
    double best = 1e30;
    for (int r = 0; r < reps; ++r) {
        CK(cudaEventRecord(a));
        launch();
        CK(cudaEventRecord(b));
        CK(cudaEventSynchronize(b));
        float ms;
        CK(cudaEventElapsedTime(&ms, a, b));
        if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    CK(cudaEventDestroy(a));
    CK(cudaEventDestroy(b));
    return best;
}

int main(int argc, char** argv) {
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    const double max_gib = argc > 1 ? atof(argv[1]) : 16.0;
    size_t fetch = 0;
    if (argc > 2) CK(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)atoi(argv[2])));   // 32, 64 or 128: a hint
    CK(cudaDeviceGetLimit(&fetch, cudaLimitMaxL2FetchGranularity));
    uint32_t* out;
    CK(cudaMalloc(&out, 4));
    if (argc > 3) {   // probe mode for ncu: one 3 GiB table, one launch of each 32-byte shape and of the 128-byte cooperative shape
        const uint64_t bytes = 3ull << 30;
        Sector* t;
        CK(cudaMalloc(&t, bytes));
        k_fill<<<sms * 8, 256>>>((uint4*)t, bytes / 16);
        const int blocks = sms * 8;
        k_chain<1, 1, 1><<<blocks, 256>>>(t, bytes / 32, 64, out);
        k_chain<0, 1, 1><<<blocks, 256>>>(t, bytes / 32, 64, out);
        k_chain<2, 4, 1><<<blocks, 256>>>(t, bytes / 128, 64, out);
        CK(cudaDeviceSynchronize());
        printf("probe done, l2 fetch granularity %d\n", (int)fetch);
        return 0;
    }
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"l2_fetch_granularity\": %d, \"results\": [", prop.name, sms, (int)fetch);
    bool first = true;
    const double sizes[4] = {0.03, 1.0, 3.0, max_gib};
    for (double gib : sizes) {
        const uint64_t bytes = (uint64_t)(gib * 1073741824.0) & ~127ULL;
        Sector* t;
        CK(cudaMalloc(&t, bytes));
        k_fill<<<sms * 8, 256>>>((uint4*)t, bytes / 16);
        CK(cudaDeviceSynchronize());
        auto emit = [&](const char* shape, int gran, int occ, int chains, double ms, double accesses) {
            printf("%s\n  {\"table_gib\": %.1f, \"shape\": \"%s\", \"granule\": %d, \"threads_per_sm\": %d, \"chains_per_thread\": %d, "
                   "\"ms\": %.3f, \"gacc_per_s\": %.2f, \"gb_per_s\": %.1f}",
                   first ? "" : ",", gib, shape, gran, occ, chains, ms, accesses / ms / 1e6, accesses * gran / ms / 1e6);
            first = false;
            fflush(stdout);
        };
        const int it = 256;
        for (int occ : {1024, 2048}) {
            const int blocks = sms * occ / 256;
            const double th = (double)blocks * 256;
            emit("2x128", 32, occ, 1, time_ms([&] { k_chain<0, 1, 1><<<blocks, 256>>>(t, bytes / 32, it, out); }, 3), th * it);
            emit("1x256", 32, occ, 1, time_ms([&] { k_chain<1, 1, 1><<<blocks, 256>>>(t, bytes / 32, it, out); }, 3), th * it);
            emit("1x256", 32, occ, 2, time_ms([&] { k_chain<1, 1, 2><<<blocks, 256>>>(t, bytes / 32, it, out); }, 3), th * it * 2);
            emit("1x256", 32, occ, 4, time_ms([&] { k_chain<1, 1, 4><<<blocks, 256>>>(t, bytes / 32, it, out); }, 3), th * it * 4);
            emit("coop", 64, occ, 1, time_ms([&] { k_chain<2, 2, 1><<<blocks, 256>>>(t, bytes / 64, it, out); }, 3), th * it / 2);
            emit("coop", 64, occ, 4, time_ms([&] { k_chain<2, 2, 4><<<blocks, 256>>>(t, bytes / 64, it, out); }, 3), th * it * 4 / 2);
            emit("coop", 128, occ, 1, time_ms([&] { k_chain<2, 4, 1><<<blocks, 256>>>(t, bytes / 128, it, out); }, 3), th * it / 4);
            emit("coop", 128, occ, 4, time_ms([&] { k_chain<2, 4, 4><<<blocks, 256>>>(t, bytes / 128, it, out); }, 3), th * it * 4 / 4);
            emit("coop", 128, occ, 8, time_ms([&] { k_chain<2, 4, 8><<<blocks, 256>>>(t, bytes / 128, it, out); }, 3), th * it * 8 / 4);
        }
        // Where is the ceiling?  One 1024-thread block per SM (forced by a 120 KB dynamic shared memory request) on
        // all SMs, then on half of them: a per-SM limit halves the rate, a memory-side limit does not.
        {
            auto kern = k_chain<1, 1, 1, 1024>;
            CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024));
            for (int frac : {1, 2, 4}) {
                const int blocks = sms / frac;
                const double ms = time_ms([&] { kern<<<blocks, 1024, 120 * 1024>>>(t, bytes / 32, it, out); }, 3);
                char shape[32];
                snprintf(shape, sizeof shape, "1x256 on %d SMs", blocks);
                emit(shape, 32, 1024, 1, ms, (double)blocks * 1024 * it);
            }
        }
        CK(cudaFree(t));
    }
    printf("\n]}\n");
    return 0;
}
