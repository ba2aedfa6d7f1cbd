// Test binary (not linked into the library): decodes a grammar file with the product's readers and the product's
// exporter, and writes the text it expands to.   slp_text FILE (plain|shaped) [leaf_len]
// With leaf_len >= the text length the exporter's literal pool IS the text (text mode); with a small leaf_len the
// text is re-read through the node table the kernels walk (grammar mode).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../moni_align_b200/csrc/device_image.hpp"
#include "../moni_align_b200/csrc/index_files.hpp"

using namespace moni;

int main(int argc, char** argv) {
    if (argc < 3) { fprintf(stderr, "usage: slp_text FILE plain|shaped [leaf_len]\n"); return 2; }
    try {
        FlatSlp g;
        if (!strcmp(argv[2], "shaped")) read_shaped_slp(argv[1], g);
        else read_plain_slp(argv[1], g);
        ExportOptions eo;
        eo.leaf_len = argc > 3 ? strtoull(argv[3], nullptr, 10) : g.text_len;
        HostImage img;
        build_grammar_table(g, eo, img);
        std::vector<uint8_t> text(g.text_len);
        for (uint64_t i = 0; i < g.text_len;) {          // root-to-literal walk, one literal block at a time
            uint32_t ref = img.hdr.root_ref;
            uint64_t off = i;
            while (!(ref & kLeafBit)) {
                const SlpNode& nd = img.nodes[ref];
                if (off < nd.left_len) ref = nd.left;
                else { off -= nd.left_len; ref = nd.right; }
            }
            const uint64_t at = ((uint64_t)(ref & ~kLeafBit) << img.hdr.pool_shift) + off;
            // the literal runs to the end of its variable: copy what is certainly inside it (one byte when unsure)
            text[i] = img.pool[at];
            ++i;
        }
        fwrite(text.data(), 1, text.size(), stdout);
        fprintf(stderr, "text %llu rules %llu nodes %llu\n", (unsigned long long)g.text_len, (unsigned long long)g.num_rules,
                (unsigned long long)img.nodes.size());
    } catch (const std::exception& e) {
        fprintf(stderr, "slp_text: %s\n", e.what());
        return 1;
    }
    return 0;
}
