// TEST-ONLY harness: compiles the per-read functions of moni_align_b200/csrc/ms_core.cuh with g++
// and runs them over the host copy of the HBM image, so the exporter and the kernel arithmetic can
// be checked against the oracle on a machine without a GPU.  Never linked into the library.
//
// usage: host_sim <index-prefix> <batch.bin> <out.bin> [leaf_len] [max_scan] [seed_min_len]
//   batch.bin : u64 n_reads, u64 offsets[n_reads+1], u8 seqs[offsets[n_reads]]
//   out.bin   : u64 ptr[total], u64 len[total], u64 mem_offsets[n_reads+1], u64 mem_pos[k], u64 mem_len[k], u64 counters[8]
//               and, when seed_min_len is given: u64 seed_offsets[n_reads+1], u64 seed_ref[s], u64 seed_len[s], u64 seed_qpos[s]
//               and, when the index has Phi structures: u64 occ_offsets[s+1], u64 occs[...]  (occ_walk per seed, cap 2^20)
// If <prefix>.thrbv.ms does not exist, <prefix>.thrbv.full.ms or .thrbv.full.lcp.ms (the `moni align` indexes) is read.
#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <vector>

#include "../moni_align_b200/csrc/index_files.hpp"
#include "../moni_align_b200/csrc/ms_core.cuh"

using namespace moni;

int main(int argc, char** argv) {
    if (argc < 4) { fprintf(stderr, "usage: host_sim prefix batch.bin out.bin [leaf_len] [max_scan]\n"); return 2; }
    try {
        FlatRlbwt bwt;
        FlatSlp slp;
        {   // same choice as moni_gpu_index_load: the ms/mems index, else one of the two `moni align` indexes
            const char* exts[3] = {".thrbv.ms", ".thrbv.full.ms", ".thrbv.full.lcp.ms"};
            int which = 0;
            for (int k = 0; k < 3; ++k) {
                FILE* probe = fopen((std::string(argv[1]) + exts[k]).c_str(), "rb");
                if (probe) { fclose(probe); which = k; break; }
            }
            read_thrbv_ms(std::string(argv[1]) + exts[which], bwt, which != 0);
        }
        read_plain_slp(std::string(argv[1]) + ".plain.slp", slp);
        ExportOptions opt;
        opt.verbose = true;
        if (argc > 4) opt.leaf_len = (uint64_t)atoll(argv[4]);
        if (argc > 5) opt.max_scan = (uint32_t)atoi(argv[5]);
        const bool want_seeds = argc > 6;
        const uint32_t seed_min_len = want_seeds ? (uint32_t)atoll(argv[6]) : 0;
        std::vector<uint64_t> so, s_ref, s_len, s_qpos;
        if (getenv("HOST_SIM_THREADS")) opt.threads = (uint32_t)atoi(getenv("HOST_SIM_THREADS"));
        HostImage img;
        build_move_table(bwt, opt, img);
        {   // hash of the move-table image (rows, row starts, gap slots, run rows): the same for any exporter thread count
            auto fnv = [](uint64_t h, const void* p, size_t k) {
                const uint8_t* b = (const uint8_t*)p;
                for (size_t i = 0; i < k; ++i) { h ^= b[i]; h *= 0x100000001b3ULL; }
                return h;
            };
            uint64_t h = 0xcbf29ce484222325ULL;
            h = fnv(h, img.rows.data(), img.rows.size() * sizeof(MoveRow));
            h = fnv(h, img.row_pos.data(), img.row_pos.size() * sizeof(img.row_pos[0]));
            h = fnv(h, img.slots.data(), img.slots.size() * sizeof(GapSlot));
            h = fnv(h, img.crun_row.data(), img.crun_row.size() * sizeof(img.crun_row[0]));
            fprintf(stderr, "host_sim: move table hash %016llx rows %llu max_scan %u\n", (unsigned long long)h,
                    (unsigned long long)img.hdr.n_rows, img.hdr.max_scan);
        }
        build_grammar_table(slp, opt, img);
        build_phi_tables(bwt, img);
        ImageView v;
        v.rows = img.rows.data(); v.row_pos = img.row_pos.data(); v.crun_row = img.crun_row.data();
        v.slots = img.slots.data(); v.nodes = img.nodes.data(); v.pool = img.pool.data(); v.sym = &img.sym; v.hdr = img.hdr;
        v.phi = img.phi.data(); v.phi_inv = img.phi_inv.data(); v.phi_dir = img.phi_dir.data(); v.phi_inv_dir = img.phi_inv_dir.data();

        FILE* f = fopen(argv[2], "rb");
        if (!f) throw std::runtime_error("cannot open batch");
        uint64_t nr = 0;
        if (fread(&nr, 8, 1, f) != 1) throw std::runtime_error("bad batch");
        std::vector<uint64_t> off(nr + 1);
        if (fread(off.data(), 8, nr + 1, f) != nr + 1) throw std::runtime_error("bad batch");
        std::vector<uint8_t> seqs(off[nr] + 16);   // load8 slack
        if (off[nr] && fread(seqs.data(), 1, off[nr], f) != off[nr]) throw std::runtime_error("bad batch");
        fclose(f);

        const uint64_t total = off[nr];
        std::vector<uint64_t> ptr(total + 1), len64(total + 1), mo(nr + 1), mems;
        Counters cnt = {0, 0, 0, 0, 0, 0, 0, 0};
        for (uint64_t r = 0; r < nr; ++r) {
            const uint64_t o = off[r], m = off[r + 1] - o;
            ScanState s;
            scan_init(v, s);
            // pass 1 the way the kernel runs it: pointers are kept only at the restarts (flags + rptr), ms_core.cuh run_note
            std::vector<uint32_t> flags(m / 32 + 2, 0);
            std::vector<uint64_t> rptr(m + 1), rrec(m + 1), direct(m + 1);
            RunSink rs;
            for (uint64_t j = m; j-- > 0;) {
                const uint64_t old = s.sample;
                scan_step(v, img.sym.tcode, seqs[o + j], s, &cnt);
                direct[j] = s.sample;
                run_note(rs, (uint32_t)j, (uint32_t)m, old, s.sample, rptr.data(), flags.data());
            }
            cnt.bases += m;
            std::vector<uint64_t> rec(m + 1), ref(m + 1);
            uint32_t k = 0;
            // the buffers as pass 2 sees them: a chunk that holds this read alone, at offset 0
            RunBufs B;
            B.seqs = seqs.data() + o;
            B.flags = flags.data();
            B.rptr = rptr.data();
            B.rrec = rrec.data();
            B.rec = rec.data();
            B.rec_ref = ref.data();
            RunRead R;
            R.o = 0; R.m = (uint32_t)m; R.fb = 0; R.cnt = rs.k; R.last = rs.k ? rs.k - 1 : 0;
            if (m) {
                k = runs_read<false>(v, B, R, &cnt);
                expand_runs(m, flags.data(), rptr.data(), rrec.data(), rs.k, ptr.data() + o, len64.data() + o);
                for (uint64_t j = 0; j < m; ++j)
                    if (ptr[o + j] != direct[j]) throw std::runtime_error("expanded pointers differ from the scan's");
            }
            mo[r] = mems.size();
            mems.insert(mems.end(), rec.begin(), rec.begin() + k);
            if (want_seeds) {
                B.rrec = nullptr;
                const uint32_t ks = m ? runs_read<true>(v, B, R, nullptr, seed_min_len) : 0;
                so.push_back(s_ref.size());
                for (uint32_t t = 0; t < ks; ++t) { s_ref.push_back(ref[t]); s_len.push_back(rec[t] & 0xFFFFFFFFu); s_qpos.push_back(rec[t] >> 32); }
            }
        }
        if (want_seeds) so.push_back(s_ref.size());
        mo[nr] = mems.size();
        std::vector<uint64_t> mp(mems.size() + 1), ml(mems.size() + 1);
        for (size_t i = 0; i < mems.size(); ++i) { mp[i] = mems[i] >> 32; ml[i] = mems[i] & 0xFFFFFFFFu; }
        f = fopen(argv[3], "wb");
        if (!f) throw std::runtime_error("cannot open output");
        fwrite(ptr.data(), 8, total, f);
        fwrite(len64.data(), 8, total, f);
        fwrite(mo.data(), 8, nr + 1, f);
        fwrite(mp.data(), 8, mems.size(), f);
        fwrite(ml.data(), 8, mems.size(), f);
        unsigned long long c8[8] = {cnt.bases, cnt.jumps, cnt.restarts, cnt.restart_depth, cnt.chars, cnt.scan_rows, cnt.general_steps, cnt.descents};
        fwrite(c8, 8, 8, f);
        if (want_seeds) {
            fwrite(so.data(), 8, so.size(), f);
            fwrite(s_ref.data(), 8, s_ref.size(), f);
            fwrite(s_len.data(), 8, s_len.size(), f);
            fwrite(s_qpos.data(), 8, s_qpos.size(), f);
            if (img.hdr.phi_n && (img.hdr.root_ref & kLeafBit)) {
                std::vector<uint64_t> oo(1, 0), occs;
                for (size_t t = 0; t < s_ref.size(); ++t) {
                    const uint64_t c = occ_walk(v, s_ref[t], s_len[t], 1ull << 20, nullptr, nullptr);
                    occs.resize(oo.back() + c);
                    occ_walk(v, s_ref[t], s_len[t], 1ull << 20, occs.data() + oo.back(), nullptr);
                    oo.push_back(oo.back() + c);
                }
                fwrite(oo.data(), 8, oo.size(), f);
                fwrite(occs.data(), 8, occs.size(), f);
            }
        }
        fclose(f);
        fprintf(stderr, "host_sim: bases %llu jumps %llu restarts %llu depth %llu chars %llu scan_rows %llu general %llu descents %llu; image %.1f MB\n",
                c8[0], c8[1], c8[2], c8[3], c8[4], c8[5], c8[6], c8[7], (double)img.bytes() / 1e6);
    } catch (const std::exception& e) {
        fprintf(stderr, "host_sim: %s\n", e.what());
        return 1;
    }
    return 0;
}
