// Host-only check of csrc/host_widen.cpp (the widening of packed results into u64 columns): every record format, misaligned
// destinations, several pool sizes, the 40-bit pointer form incl. the wrapped values.  Test binary; prints "ok" or the first mismatch.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../moni_align_b200/csrc/host_widen.hpp"

static uint64_t mix(uint64_t i) { return (i + 1) * 0x9E3779B97F4A7C15ull; }

int main() {
    moni::WidenPool& P = moni::WidenPool::get();
    const uint64_t k = 1000003;
    std::vector<uint16_t> r16(k);
    std::vector<uint32_t> r32(k + 16), lo(k);
    std::vector<uint8_t> hi(k + 8);
    std::vector<uint16_t> ln(k);
    for (uint64_t i = 0; i < k; ++i) {
        r16[i] = (uint16_t)(mix(i) >> 13);
        r32[i] = (uint32_t)(mix(i) >> 7);
        uint64_t v = mix(i) >> 24;                                   // 40 bits
        if (i % 7 == 0) v = (1ull << 40) - 1 - (i % 1000);           // wrapped pointers 2^64 - j
        if (i % 11 == 0) v = (1ull << 40) - (1ull << 31);            // the first wrapped value
        if (i % 13 == 0) v = (1ull << 40) - (1ull << 31) - 1;        // the last plain one
        lo[i] = (uint32_t)v;
        hi[i] = (uint8_t)(v >> 32);
        ln[i] = (uint16_t)(mix(i) >> 40);
    }
    for (uint64_t i = k; i < k + 16; ++i) r32[i] = (uint32_t)mix(i);
    std::vector<uint64_t> a(k + 8), b(k + 8);
    for (int threads : {0, 1, 3, 5}) {
        P.resize(threads);
        for (int shift = 0; shift < 4; ++shift) {
            uint64_t* pa = a.data() + shift;
            uint64_t* pb = b.data() + ((shift + 2) & 3);
            P.widen(1, r16.data(), nullptr, k, pa, pb);
            for (uint64_t i = 0; i < k; ++i)
                if (pa[i] != (uint64_t)(r16[i] >> 8) || pb[i] != (uint64_t)(r16[i] & 0xFF)) { printf("format 1 differs at %llu (threads %d)\n", (unsigned long long)i, threads); return 1; }
            P.widen(2, r32.data(), nullptr, k, pa, pb);
            for (uint64_t i = 0; i < k; ++i)
                if (pa[i] != (uint64_t)(r32[i] >> 16) || pb[i] != (uint64_t)(r32[i] & 0xFFFF)) { printf("format 2 differs at %llu (threads %d)\n", (unsigned long long)i, threads); return 1; }
            P.widen(4, r32.data(), r32.data() + 9, k, pa, pb);
            for (uint64_t i = 0; i < k; ++i)
                if (pa[i] != r32[i] || pb[i] != r32[i + 9]) { printf("format 4 differs at %llu (threads %d)\n", (unsigned long long)i, threads); return 1; }
            P.widen_ms(lo.data(), hi.data(), ln.data(), k, pa, pb);
            for (uint64_t i = 0; i < k; ++i) {
                const uint64_t v = (uint64_t)lo[i] | ((uint64_t)hi[i] << 32);
                const uint64_t w = v >= (1ull << 40) - (1ull << 31) ? v - (1ull << 40) : v;      // u64 arithmetic: 2^64 - j
                if (pa[i] != w || pb[i] != ln[i]) { printf("ms form differs at %llu (threads %d)\n", (unsigned long long)i, threads); return 1; }
            }
        }
        P.widen(1, r16.data(), nullptr, 0, a.data(), b.data());      // nothing to do
        P.widen(2, r32.data(), nullptr, 1, a.data(), b.data());
        if (a[0] != (uint64_t)(r32[0] >> 16)) { puts("single record differs"); return 1; }
    }
    P.resize(0);
    puts("ok");
    return 0;
}
