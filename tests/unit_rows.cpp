// TEST-ONLY: the register view of a MoveRow (RowRegs in moni_align_b200/csrc/ms_core.cuh, what the kernels decode after one
// LDG.E.256) must agree with the struct the exporter fills (MoveRow in device_image.hpp) for every field at its extremes --
// in particular the two 40-bit SA samples, which no test index is large enough to exercise -- and the Phi step and packed
// pointer arithmetic must hold near 2^40.
#include <cstdio>
#include <cstring>
#include <random>

#include "../moni_align_b200/csrc/ms_core.cuh"

using namespace moni;

static void put40(uint8_t* d, uint64_t v) { for (int k = 0; k < 5; ++k) d[k] = (uint8_t)(v >> (8 * k)); }

int main() {
    std::mt19937_64 rng(12345);
    const uint64_t edge40[] = {0, 1, 0xFFFFFFFFull, 0x100000000ull, (1ull << 40) - 1, (1ull << 39), 0x00FF00FF00ull};
    for (int it = 0; it < 200000; ++it) {
        MoveRow r;
        memset(&r, 0, sizeof(r));
        const bool edge = it < 7 * 7;
        r.dest_row = edge ? 0xFFFFFFFFu - (uint32_t)it : (uint32_t)rng();
        r.dest_off = (uint16_t)rng();
        r.len = (uint16_t)rng();
        for (int s = 0; s < 3; ++s) { r.t_rel[s] = (uint16_t)rng(); r.dist_up[s] = (uint8_t)rng(); r.dist_dn[s] = (uint8_t)rng(); }
        r.head = (uint8_t)rng();
        r.flags = (uint8_t)rng();
        const uint64_t a = edge ? edge40[it % 7] : rng() & ((1ull << 40) - 1), b = edge ? edge40[it / 7] : rng() & ((1ull << 40) - 1);
        put40(r.s_start, a);
        put40(r.s_last, b);
        const RowRegs g = load_row(&r, 0);
        bool ok = g.dest_row() == r.dest_row && g.dest_off() == r.dest_off && g.len() == r.len && g.head() == r.head && g.flags() == r.flags &&
                  g.s_start() == a && g.s_last() == b;
        for (int s = 0; s < 3; ++s) ok = ok && g.t_rel(s) == r.t_rel[s] && g.dist_up(s) == r.dist_up[s] && g.dist_dn(s) == r.dist_dn[s];
        if (!ok) { fprintf(stderr, "unit_rows: field mismatch at iteration %d\n", it); return 1; }
    }
    // Phi step over a 3-entry table with keys and values near 2^40 (n just below 2^40), with and without wrap-around
    const uint64_t n = (1ull << 40) - 5;
    const PhiEntry tab[3] = {{10, n - 3}, {1ull << 39, 7}, {n - 2, (1ull << 39) + 1}};
    const uint32_t shift = 38;                                   // buckets of 2^38 positions: (n >> 38) + 2 = 5 words
    uint32_t dir[5];
    for (uint64_t bk = 0; bk < 5; ++bk) { uint32_t j = 0; while (j < 3 && tab[j].key < (bk << shift)) ++j; dir[bk] = j; }
    struct { uint64_t i, want; } cases[] = {
        {11, (n - 3 + 1) % n},                     // predecessor key 10, delta 1
        {10, ((1ull << 39) + 1 + 10 + 1) % n},     // no key < 10: circular predecessor is the last entry, delta = i + 1
        {(1ull << 39) + 5, 7 + 5},                 // predecessor 2^39
        {n - 1, ((1ull << 39) + 1 + 1) % n},       // predecessor n - 2
        {5, ((1ull << 39) + 1 + 6) % n},
    };
    for (auto& c : cases) {
        const uint64_t got = phi_step(tab, dir, shift, 3, n, c.i);
        if (got != c.want) { fprintf(stderr, "unit_rows: phi_step(%llu) = %llu, want %llu\n", (unsigned long long)c.i, (unsigned long long)got, (unsigned long long)c.want); return 1; }
    }
    // packed pointers: v & (2^40 - 1) round-trips text positions below 2^40 - 2^31 and wrapped values 2^64 - k, k < 2^31
    const uint64_t vals[] = {0, 1, (1ull << 40) - (1ull << 31) - 1, 0xFFFFFFFFFFFFFFFFull, 0xFFFFFFFFFFFFFFFFull - 0x7FFFFFFEull, (1ull << 39)};
    for (uint64_t v : vals) {
        const uint64_t enc = v & ((1ull << 40) - 1);
        const uint64_t dec = enc >= (1ull << 40) - (1ull << 31) ? enc - (1ull << 40) : enc;
        if (dec != v) { fprintf(stderr, "unit_rows: packed pointer %llx does not round-trip\n", (unsigned long long)v); return 1; }
    }
    printf("unit_rows OK\n");
    return 0;
}
