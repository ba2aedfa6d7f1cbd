// TEST INFRASTRUCTURE -- drives the UNMODIFIED reference seed_finder<plain_slp_t, ms_pointers<>>::find_mems
// (include/aligner/seed_finder.hpp:126-166 of the reference checkout) over a FASTQ and prints, per read,
//     >name\n(pos,len,idx) (pos,len,idx) ...\n
// where pos = pointers[i] (text position), len = seed length, idx = query offset i.  Built by oracle/build_ref.sh
// into oracle/_ref/bin/ref_seeds; the reference headers are compiled where they lie, nothing is copied.
// The only thing replaced is the lift-over index member (class liftidx, levioSAM + htslib), which find_mems never
// touches: oracle/build_ref.sh generates a three-method stand-in under oracle/_ref/shim/liftstub/liftidx.hpp that
// shadows include/aligner/liftidx.hpp on the include path.
//
// With a fifth argument "occs" every seed is followed by its occurrence list as the reference enumerates it with
// seed_finder::find_MEM_occs (:239-251: the seed's own position, then Phi steps upwards and Phi^-1 steps downwards while
// the bounded LCE reaches the seed length; no seed filtering):   (pos,len,idx:occ,occ,...)
//
// With "pop" instead, the read's MEMs go through seed_finder::populate_seeds (:311-319 -> populate_seed :258-309, seed filtering off):
// every MEM gets its occurrence list and a MEM of at least 2 * min_len bases appends its two halves to the vector -- the left
// half (upper_suffix, len / 2, idx) whose list is upper_suffix, the suffixes above it and the suffixes below lower_suffix at the
// shorter length, and the right half (pos + len / 2, len - len / 2, idx + len / 2) with its own find_MEM_occs list.  The whole
// vector is printed in its final order, same token format.
//
// usage: ref_seeds <index-prefix> <reads.fq> <min_len> <out.seeds> [occs|pop]
//   needs <prefix>.thrbv.full.ms (reference rlebwt_ms_full_build -n), <prefix>.plain.slp and an (empty) <prefix>.ldx
#include <cstdio>
#include <cstdlib>
#include <string>

#include <seed_finder.hpp>

#include <zlib.h>
// kseq's KSEQ_INIT(gzFile, gzread) is already instantiated by the reference headers (seqidx.hpp / common.hpp)

int main(int argc, char** argv) {
    if (argc != 5 && argc != 6) { fprintf(stderr, "usage: ref_seeds prefix reads.fq min_len out [occs]\n"); return 2; }
    const bool with_occs = argc == 6;
    const bool populate = argc == 6 && std::string(argv[5]) == "pop";
    const size_t min_len = (size_t)atoll(argv[3]);
#ifdef REF_SEEDS_LCP   // the sampled-LCP index: <prefix>.thrbv.full.lcp.ms, occurrences through moni_lcp::Phi_lcp / Phi_inv_lcp
    seed_finder<plain_slp_t, moni_lcp<>> sf(argv[1], min_len, false, 5000);
#else
    seed_finder<plain_slp_t, ms_pointers<>> sf(argv[1], min_len, false, 5000);
#endif
    gzFile fp = gzopen(argv[2], "r");
    if (!fp) { fprintf(stderr, "cannot open %s\n", argv[2]); return 1; }
    FILE* out = fopen(argv[4], "w");
    if (!out) { fprintf(stderr, "cannot open %s\n", argv[4]); return 1; }
    kseq_t* seq = kseq_init(fp);
    while (kseq_read(seq) >= 0) {
        std::vector<mem_t> mems;
        sf.find_mems(seq, mems);
        if (populate) sf.populate_seeds(mems);
        fprintf(out, ">%s\n", seq->name.s);
        for (mem_t& m : mems) {
            if (!with_occs) { fprintf(out, "(%zu,%zu,%zu) ", m.pos, m.len, m.idx); continue; }
            if (!populate) sf.find_MEM_occs(m);
            fprintf(out, "(%zu,%zu,%zu:", m.pos, m.len, m.idx);
            for (size_t k = 0; k < m.occs.size(); ++k) fprintf(out, k ? ",%zu" : "%zu", m.occs[k]);
            fprintf(out, ") ");
        }
        fprintf(out, "\n");
    }
    kseq_destroy(seq);
    gzclose(fp);
    fclose(out);
    return 0;
}
