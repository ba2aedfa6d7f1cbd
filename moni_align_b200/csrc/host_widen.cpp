// host_widen.cpp -- see host_widen.hpp.  Plain C++ (no CUDA): compiled by the host compiler so that the AVX2 variant can live
// behind a target attribute and a run-time check.
#include "host_widen.hpp"

#include <immintrin.h>
#include <sched.h>

#include <algorithm>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>

namespace moni {

namespace {

// One column: out[i] = (rec(i) >> shift) & mask for i in [0, k), rec = u16 / u32 words.  Streaming stores for the aligned body
// (the destination is written once and not read back by this thread: no write-allocate traffic).
template <typename RecT>
void column_sse2(const RecT* src, uint64_t k, unsigned shift, uint64_t mask, uint64_t* out) {
    uint64_t i = 0;
    while (i < k && ((uintptr_t)(out + i) & 15u)) { out[i] = ((uint64_t)src[i] >> shift) & mask; ++i; }
    for (; i + 2 <= k; i += 2) {
        const uint64_t a = ((uint64_t)src[i] >> shift) & mask, b = ((uint64_t)src[i + 1] >> shift) & mask;
        _mm_stream_si128((__m128i*)(out + i), _mm_set_epi64x((long long)b, (long long)a));
    }
    for (; i < k; ++i) out[i] = ((uint64_t)src[i] >> shift) & mask;
}

__attribute__((target("avx2"))) void column_avx2_u16(const uint16_t* src, uint64_t k, unsigned shift, uint64_t mask, uint64_t* out) {
    uint64_t i = 0;
    while (i < k && ((uintptr_t)(out + i) & 31u)) { out[i] = ((uint64_t)src[i] >> shift) & mask; ++i; }
    const __m256i m = _mm256_set1_epi64x((long long)mask);
    const __m128i sh = _mm_cvtsi32_si128((int)shift);
    for (; i + 4 <= k; i += 4) {
        const __m128i w = _mm_loadl_epi64((const __m128i*)(src + i));
        const __m256i v = _mm256_and_si256(_mm256_srl_epi64(_mm256_cvtepu16_epi64(w), sh), m);
        _mm256_stream_si256((__m256i*)(out + i), v);
    }
    for (; i < k; ++i) out[i] = ((uint64_t)src[i] >> shift) & mask;
}

__attribute__((target("avx2"))) void column_avx2_u32(const uint32_t* src, uint64_t k, unsigned shift, uint64_t mask, uint64_t* out) {
    uint64_t i = 0;
    while (i < k && ((uintptr_t)(out + i) & 31u)) { out[i] = ((uint64_t)src[i] >> shift) & mask; ++i; }
    const __m256i m = _mm256_set1_epi64x((long long)mask);
    const __m128i sh = _mm_cvtsi32_si128((int)shift);
    for (; i + 4 <= k; i += 4) {
        const __m128i w = _mm_loadu_si128((const __m128i*)(src + i));
        const __m256i v = _mm256_and_si256(_mm256_srl_epi64(_mm256_cvtepu32_epi64(w), sh), m);
        _mm256_stream_si256((__m256i*)(out + i), v);
    }
    for (; i < k; ++i) out[i] = ((uint64_t)src[i] >> shift) & mask;
}

constexpr uint64_t kWrapFrom = (1ull << 40) - (1ull << 31);

inline uint64_t decode40(uint32_t lo, uint8_t hi) {
    const uint64_t v = (uint64_t)lo | ((uint64_t)hi << 32);
    return v >= kWrapFrom ? v - (1ull << 40) : v;
}

void pointers_sse2(const uint32_t* lo, const uint8_t* hi, uint64_t k, uint64_t* out) {
    uint64_t i = 0;
    while (i < k && ((uintptr_t)(out + i) & 15u)) { out[i] = decode40(lo[i], hi[i]); ++i; }
    for (; i + 2 <= k; i += 2)
        _mm_stream_si128((__m128i*)(out + i), _mm_set_epi64x((long long)decode40(lo[i + 1], hi[i + 1]), (long long)decode40(lo[i], hi[i])));
    for (; i < k; ++i) out[i] = decode40(lo[i], hi[i]);
}

__attribute__((target("avx2"))) void pointers_avx2(const uint32_t* lo, const uint8_t* hi, uint64_t k, uint64_t* out) {
    uint64_t i = 0;
    while (i < k && ((uintptr_t)(out + i) & 31u)) { out[i] = decode40(lo[i], hi[i]); ++i; }
    const __m256i thr = _mm256_set1_epi64x((long long)(kWrapFrom - 1)), wrap = _mm256_set1_epi64x(1ll << 40);
    for (; i + 4 <= k; i += 4) {
        const __m256i l = _mm256_cvtepu32_epi64(_mm_loadu_si128((const __m128i*)(lo + i)));
        int h4;
        __builtin_memcpy(&h4, hi + i, 4);
        const __m256i h = _mm256_slli_epi64(_mm256_cvtepu8_epi64(_mm_cvtsi32_si128(h4)), 32);
        __m256i v = _mm256_or_si256(l, h);                                  // < 2^40: signed compare is fine
        v = _mm256_sub_epi64(v, _mm256_and_si256(_mm256_cmpgt_epi64(v, thr), wrap));
        _mm256_stream_si256((__m256i*)(out + i), v);
    }
    for (; i < k; ++i) out[i] = decode40(lo[i], hi[i]);
}

const bool g_avx2 = __builtin_cpu_supports("avx2");

void column(const uint16_t* src, uint64_t k, unsigned shift, uint64_t mask, uint64_t* out) {
    if (g_avx2) column_avx2_u16(src, k, shift, mask, out);
    else column_sse2(src, k, shift, mask, out);
}
void column(const uint32_t* src, uint64_t k, unsigned shift, uint64_t mask, uint64_t* out) {
    if (g_avx2) column_avx2_u32(src, k, shift, mask, out);
    else column_sse2(src, k, shift, mask, out);
}

}  // namespace

void widen_records(int cfmt, const void* src, const void* src2, uint64_t k, uint64_t* dpos, uint64_t* dlen) {
    constexpr uint64_t kBlockRecs = 8192;       // the records of a block are read twice (once per column): they stay in L1/L2
    for (uint64_t a = 0; a < k; a += kBlockRecs) {
        const uint64_t n = std::min(kBlockRecs, k - a);
        if (cfmt == 1) {
            const uint16_t* s = (const uint16_t*)src + a;
            column(s, n, 8, 0xFFFFFFFFFFFFFFFFull, dpos + a);
            column(s, n, 0, 0xFFull, dlen + a);
        } else if (cfmt == 2) {
            const uint32_t* s = (const uint32_t*)src + a;
            column(s, n, 16, 0xFFFFFFFFFFFFFFFFull, dpos + a);
            column(s, n, 0, 0xFFFFull, dlen + a);
        } else {
            column((const uint32_t*)src + a, n, 0, 0xFFFFFFFFFFFFFFFFull, dpos + a);
            column((const uint32_t*)src2 + a, n, 0, 0xFFFFFFFFFFFFFFFFull, dlen + a);
        }
    }
    _mm_sfence();
}

void widen_ms(const uint32_t* lo, const uint8_t* hi, const uint16_t* len, uint64_t k, uint64_t* dptr, uint64_t* dlen) {
    if (g_avx2) pointers_avx2(lo, hi, k, dptr);
    else pointers_sse2(lo, hi, k, dptr);
    column(len, k, 0, 0xFFFFFFFFFFFFFFFFull, dlen);
    _mm_sfence();
}

struct WidenPool::Impl {
    std::mutex mu, call_mu;
    std::condition_variable cv, done_cv;
    std::vector<std::thread> workers;
    const std::function<void(int)>* fn = nullptr;
    int n_slices = 0, next = 0, pending = 0;
    uint64_t epoch = 0;
    bool stop = false;

    void loop() {
        std::unique_lock<std::mutex> lock(mu);
        for (;;) {
            cv.wait(lock, [&] { return stop || (fn && next < n_slices); });
            if (stop) return;
            while (fn && next < n_slices) {
                const int s = next++;
                const std::function<void(int)>* f = fn;
                lock.unlock();
                (*f)(s);
                lock.lock();
                if (--pending == 0) { fn = nullptr; done_cv.notify_all(); }
            }
        }
    }
};

WidenPool::WidenPool() : impl_(new Impl) {}
WidenPool::~WidenPool() {
    resize(0);
    delete impl_;
}

WidenPool& WidenPool::get() {
    static WidenPool pool;
    return pool;
}

int WidenPool::default_threads() {
    cpu_set_t set;
    CPU_ZERO(&set);
    int allowed = (int)std::thread::hardware_concurrency();
    if (sched_getaffinity(0, sizeof(set), &set) == 0) allowed = CPU_COUNT(&set);
    if (allowed < 4) return 0;                     // nothing to gain over the copy engine with so few cores
    return std::min(16, allowed * 2 / 3);
}

void WidenPool::resize(int n) {
    std::lock_guard<std::mutex> call(impl_->call_mu);
    {
        std::lock_guard<std::mutex> lock(impl_->mu);
        impl_->stop = true;
    }
    impl_->cv.notify_all();
    for (auto& t : impl_->workers) t.join();
    impl_->workers.clear();
    impl_->stop = false;
    n_ = std::max(n, 0);
    for (int i = 0; i < n_; ++i) impl_->workers.emplace_back([this] { impl_->loop(); });
}

void WidenPool::parallel_for(int n_slices, const std::function<void(int)>& fn) {
    if (n_slices <= 0) return;
    std::lock_guard<std::mutex> call(impl_->call_mu);
    if (n_ == 0) {
        for (int s = 0; s < n_slices; ++s) fn(s);
        return;
    }
    std::unique_lock<std::mutex> lock(impl_->mu);
    impl_->fn = &fn;
    impl_->n_slices = n_slices;
    impl_->next = 0;
    impl_->pending = n_slices;
    impl_->cv.notify_all();
    impl_->done_cv.wait(lock, [&] { return impl_->pending == 0; });
}

void WidenPool::widen(int cfmt, const void* src, const void* src2, uint64_t k, uint64_t* dpos, uint64_t* dlen) {
    if (!k) return;
    // slices of whole 64-record groups, so that every slice but the first starts 512-byte aligned relative to its neighbours
    const uint64_t per = std::max<uint64_t>(1 << 15, ((k + (uint64_t)std::max(n_, 1) * 2 - 1) / ((uint64_t)std::max(n_, 1) * 2) + 63) & ~63ull);
    const int slices = (int)((k + per - 1) / per);
    const int w = cfmt == 1 ? 2 : 4;
    parallel_for(slices, [&](int s) {
        const uint64_t a = (uint64_t)s * per, n = std::min(per, k - a);
        widen_records(cfmt, (const char*)src + a * w, src2 ? (const char*)src2 + a * 4 : nullptr, n, dpos + a, dlen + a);
    });
}

void WidenPool::widen_ms(const uint32_t* lo, const uint8_t* hi, const uint16_t* len, uint64_t k, uint64_t* dptr, uint64_t* dlen) {
    if (!k) return;
    const uint64_t parts = (uint64_t)std::max(n_, 1) * 4;
    const uint64_t per = std::max<uint64_t>(1 << 15, ((k + parts - 1) / parts + 63) & ~63ull);
    parallel_for((int)((k + per - 1) / per), [&](int s) {
        const uint64_t a = (uint64_t)s * per, n = std::min(per, k - a);
        moni::widen_ms(lo + a, hi + a, len + a, n, dptr + a, dlen + a);
    });
}

}  // namespace moni
