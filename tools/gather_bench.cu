// gather_bench: measured random-access HBM roofline of this box (SURVEY.md 8(d) denominator).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_build/gather_bench tools/gather_bench.cu
//   tools/_build/gather_bench [table_GiB=16] > profiles/random_access_peak.json
//
// Two access models over a table much larger than L2, for 32 / 64 / 128-byte aligned granules:
//   "indep"  every thread issues UNROLL independent random granule reads per iteration (bandwidth cap of
//            random sectors: what the memory system can deliver when latency is fully hidden);
//   "chain"  every thread follows ONE dependent chain (next address = hash(loaded data)), the shape of an
//            LF walk: throughput = resident threads / latency, capped by the above.
// Reported GB/s count only the granule bytes requested (granule x accesses / time).
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                                   \
    do {                                                                                        \
        cudaError_t e = (x);                                                                    \
        if (e != cudaSuccess) {                                                                 \
            fprintf(stderr, "%s: %s (%s:%d)\n", #x, cudaGetErrorString(e), __FILE__, __LINE__); \
            exit(1);                                                                            \
        }                                                                                       \
    } while (0)

__device__ __forceinline__ uint64_t mix(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
    return x;
}

__global__ void k_fill(uint4* t, uint64_t n16) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n16; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t h = mix(i);
        t[i] = make_uint4((uint32_t)h, (uint32_t)(h >> 32), (uint32_t)i, 0x9E3779B9u);
    }
}

template <int G16>   // granule in 16-byte units: 2, 4, 8
__device__ __forceinline__ uint32_t load_granule(const uint4* t, uint64_t g) {
    const uint4* p = t + g * G16;
    uint32_t acc = 0;
#pragma unroll
    for (int k = 0; k < G16; ++k) {
        const uint4 v = __ldg(p + k);
        acc ^= v.x ^ v.y ^ v.z ^ v.w;
    }
    return acc;
}

template <int G16, int UNROLL>
__global__ void __launch_bounds__(256) k_indep(const uint4* __restrict__ t, uint64_t n_gran, int iters, uint32_t* out) {
    const uint64_t tid = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    uint64_t s = mix(tid + 1);
    uint32_t acc = 0;
    for (int it = 0; it < iters; ++it) {
        uint64_t g[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) { s = s * 6364136223846793005ULL + 1442695040888963407ULL; g[u] = (s >> 20) % n_gran; }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) acc ^= load_granule<G16>(t, g[u]);
    }
    if (acc == 0x12345678u) out[0] = acc;
}

template <int G16>
__global__ void __launch_bounds__(256) k_chain(const uint4* __restrict__ t, uint64_t n_gran, int iters, uint32_t* out) {
    const uint64_t tid = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    uint64_t g = mix(tid + 1) % n_gran;
    uint32_t acc = 0;
    for (int it = 0; it < iters; ++it) {
        const uint32_t v = load_granule<G16>(t, g);
        acc ^= v;
        g = mix(((uint64_t)v << 20) ^ g ^ (uint64_t)it) % n_gran;      // depends on the loaded data
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
    return best;
}

int main(int argc, char** argv) {
    const double gib = argc > 1 ? atof(argv[1]) : 16.0;
    const uint64_t bytes = (uint64_t)(gib * 1073741824.0) & ~127ULL;
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    uint4* t;
    uint32_t* out;
    CK(cudaMalloc(&t, bytes));
    CK(cudaMalloc(&out, 4));
    k_fill<<<prop.multiProcessorCount * 8, 256>>>(t, bytes / 16);
    CK(cudaDeviceSynchronize());
    const int sms = prop.multiProcessorCount;
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"table_gib\": %.1f, \"results\": [", prop.name, sms, gib);
    bool first = true;
    auto emit = [&](const char* model, int gran, int threads_per_sm, int unroll, double ms, double accesses) {
        printf("%s\n  {\"model\": \"%s\", \"granule\": %d, \"threads_per_sm\": %d, \"unroll\": %d, \"ms\": %.3f, \"gacc_per_s\": %.3f, \"gb_per_s\": %.1f}",
               first ? "" : ",", model, gran, threads_per_sm, unroll, ms, accesses / ms / 1e6, accesses * gran / ms / 1e6);
        first = false;
    };
    for (int occ : {1024, 2048}) {
        const int blocks = sms * occ / 256;
        const double threads = (double)blocks * 256;
        {
            const int it = 64;
            emit("indep", 32, occ, 4, time_ms([&] { k_indep<2, 4><<<blocks, 256>>>(t, bytes / 32, it, out); }, 3), threads * it * 4);
            emit("indep", 64, occ, 4, time_ms([&] { k_indep<4, 4><<<blocks, 256>>>(t, bytes / 64, it, out); }, 3), threads * it * 4);
            emit("indep", 128, occ, 4, time_ms([&] { k_indep<8, 4><<<blocks, 256>>>(t, bytes / 128, it, out); }, 3), threads * it * 4);
        }
        {
            const int it = 256;
            emit("chain", 32, occ, 1, time_ms([&] { k_chain<2><<<blocks, 256>>>(t, bytes / 32, it, out); }, 3), threads * it);
            emit("chain", 64, occ, 1, time_ms([&] { k_chain<4><<<blocks, 256>>>(t, bytes / 64, it, out); }, 3), threads * it);
            emit("chain", 128, occ, 1, time_ms([&] { k_chain<8><<<blocks, 256>>>(t, bytes / 128, it, out); }, 3), threads * it);
        }
    }
    printf("\n]}\n");
    return 0;
}
