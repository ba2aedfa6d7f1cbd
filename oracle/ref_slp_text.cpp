// TEST INFRASTRUCTURE (ours; compiled by oracle/build_ref.sh against the UNMODIFIED reference headers).
// Writes to stdout the text that the reference's own shaped-grammar decoder produces for a `<prefix>.slp` file:
// SelfShapedSlp<uint32_t, DagcSd, DagcSd, SelSd>::load (thirdparty/ShapedSlp/SelfShapedSlp.hpp:378-397) followed by
// expandSubstr(0, getLen(), ...) (:134-153), the type `ms -q` would instantiate (src/matching_statistics.cpp:164-168).
// With a second argument "charat" every position is read through charAt (:100-131), the accessor on the hot path.
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include <SelectType.hpp>
#include <DirectAccessibleGammaCode.hpp>
#include <SelfShapedSlp.hpp>

using SelSd = SelectSdvec<>;
using DagcSd = DirectAccessibleGammaCode<SelSd>;
using shaped_slp_t = SelfShapedSlp<uint32_t, DagcSd, DagcSd, SelSd>;

int main(int argc, char** argv) {
    if (argc < 2) { fprintf(stderr, "usage: ref_slp_text FILE.slp [charat]\n"); return 2; }
    std::ifstream in(argv[1], std::ios::binary);
    if (!in) { fprintf(stderr, "cannot open %s\n", argv[1]); return 1; }
    shaped_slp_t slp;
    slp.load(in);
    const uint64_t n = slp.getLen();
    std::vector<char> text(n);
    if (argc > 2 && !strcmp(argv[2], "charat")) {
        for (uint64_t i = 0; i < n; ++i) text[i] = slp.charAt(i);
    } else if (n) {
        slp.expandSubstr(0, n, text.data());
    }
    fwrite(text.data(), 1, n, stdout);
    return 0;
}
