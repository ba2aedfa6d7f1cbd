// Drop-in `ms` / `mems` executables over libmoni_gpu.so (compile with -DMONI_CLI_MS or -DMONI_CLI_MEMS).
//
// Same command line, file names and output bytes as the reference binaries:
//   ms   infile [-p patterns] [-o output] [-t threads] [-l len]     src/matching_statistics.cpp:432-484
//   mems infile [-p patterns] [-o output] [-t threads] [-l len]     src/mems.cpp:432-484
// default output prefix = patterns + "_" + basename(infile) (:503-506); outputs <out>.pointers and
// <out>.lengths (ms, text format :569-598) or <out>.mems (mems, src/mems.cpp:585-587), records in input
// order.  `-l` is parsed and ignored exactly like the reference.  `-q` selects the shaped grammar `<infile>.slp`: the reference
// advertises and implements it (:438-440, :178-182, :626) but leaves 'q' out of its getopt string (:447), so there the flag
// is rejected; here it works, with identical results (they do not depend on the grammar encoding).
// `-t` is the number of host threads that format text; the GPUs do the
// computation (extra flag `-g 0,1,..` selects devices; default: device 0, or $MONI_GPUS).
// Reads: FASTA or FASTQ, plain or gzip (zlib), names cut at the first whitespace, sequence bytes used
// verbatim -- the behaviour of kseq.h in the reference (thirdparty/klib/kseq.h).
#include <getopt.h>
#include <libgen.h>
#include <zlib.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/moni_gpu.h"

namespace {

[[noreturn]] void die(const std::string& msg) {
    // the reference's error(): "[ERROR] ..." on stderr, exit(1)   (include/common/common.hpp:108-117)
    fprintf(stderr, "[ERROR] %s\n", msg.c_str());
    exit(1);
}

void info(const char* msg) {
    const time_t t = time(nullptr);
    struct tm tmv;
    localtime_r(&t, &tmv);
    printf("[INFO] %02d:%02d:%02d - Message: %s\n", tmv.tm_hour, tmv.tm_min, tmv.tm_sec, msg);
    fflush(stdout);
}

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// ---- FASTA/FASTQ reader with kseq semantics -----------------------------------------------------
struct SeqReader {
    gzFile f = nullptr;
    std::vector<unsigned char> buf;
    size_t pos = 0, len = 0;
    int last = 0;   // the record marker already consumed ('>' or '@'), 0 = none yet
    bool eof = false;
    explicit SeqReader(const std::string& path) : buf(1 << 20) {
        f = gzopen(path.c_str(), "r");
        if (!f) die("cannot open patterns file " + path);
        gzbuffer(f, 1 << 20);
    }
    ~SeqReader() { if (f) gzclose(f); }
    int getc() {
        if (pos >= len) {
            if (eof) return -1;
            const int k = gzread(f, buf.data(), (unsigned)buf.size());
            if (k <= 0) { eof = true; return -1; }
            len = (size_t)k;
            pos = 0;
        }
        return buf[pos++];
    }
    // appends the rest of the current line (without the newline) to out; returns the delimiter (-1 at EOF)
    int line(std::string* out) {
        for (;;) {
            const int c = getc();
            if (c < 0 || c == '\n') return c;
            if (out && c != '\r') out->push_back((char)c);
        }
    }
    // Reads one record; name cut at first whitespace.  Returns false at end of file.
    bool next(std::string& name, std::string& seq) {
        int c;
        if (last == 0) {
            while ((c = getc()) >= 0 && c != '>' && c != '@') {}
            if (c < 0) return false;
            last = c;
        }
        name.clear();
        seq.clear();
        std::string header;
        line(&header);
        const size_t sp = header.find_first_of(" \t");
        name = header.substr(0, sp);
        // sequence lines until the next marker
        while ((c = getc()) >= 0 && c != '>' && c != '+' && c != '@') {
            if (c == '\n') continue;
            if (c != '\r') seq.push_back((char)c);
            line(&seq);
        }
        if (c == '>' || c == '@') last = c;
        if (c != '+') { if (c < 0) last = 0, eof = true; return true; }
        line(nullptr);                                   // rest of the '+' line
        size_t q = 0;
        std::string qual;
        while (q < seq.size()) {                          // quality: as many characters as bases
            qual.clear();
            const int d = line(&qual);
            q += qual.size();
            if (d < 0) break;
        }
        last = 0;
        return true;
    }
};

inline char* put_u64(char* p, uint64_t v) {
    char tmp[24];
    int k = 0;
    do { tmp[k++] = (char)('0' + v % 10); v /= 10; } while (v);
    while (k) *p++ = tmp[--k];
    return p;
}

struct Batch {
    std::vector<std::string> names;
    std::string seqs;
    std::vector<uint64_t> offsets{0};
    void clear() { names.clear(); seqs.clear(); offsets.assign(1, 0); }
};

// formats reads [a, b) of one batch into `out`
void format_values(const Batch& B, const uint64_t* vals, size_t a, size_t b, std::string& out) {
    out.clear();
    std::vector<char> tmp;
    for (size_t r = a; r < b; ++r) {
        const uint64_t o = B.offsets[r], m = B.offsets[r + 1] - o;
        tmp.resize(m * 21 + 8);
        char* p = tmp.data();
        for (uint64_t i = 0; i < m; ++i) { p = put_u64(p, vals[o + i]); *p++ = ' '; }
        out.push_back('>');
        out += B.names[r];
        out.push_back('\n');
        out.append(tmp.data(), p - tmp.data());
        out.push_back('\n');
    }
}

void format_mems(const Batch& B, const uint64_t* moff, const uint64_t* mpos, const uint64_t* mlen, size_t a, size_t b, std::string& out) {
    out.clear();
    std::vector<char> tmp;
    for (size_t r = a; r < b; ++r) {
        const uint64_t s = moff[r], e = moff[r + 1];
        tmp.resize((e - s) * 46 + 8);
        char* p = tmp.data();
        for (uint64_t k = s; k < e; ++k) {
            *p++ = '(';
            p = put_u64(p, mpos[k]);
            *p++ = ',';
            p = put_u64(p, mlen[k]);
            *p++ = ')';
            *p++ = ' ';
        }
        out.push_back('>');
        out += B.names[r];
        out.push_back('\n');
        out.append(tmp.data(), p - tmp.data());
        out.push_back('\n');
    }
}

template <typename F>
void parallel_format(size_t n_reads, size_t threads, std::vector<std::string>& parts, F&& fn) {
    threads = std::max<size_t>(1, std::min(threads, n_reads ? n_reads : 1));
    parts.resize(threads);
    std::vector<std::thread> th;
    for (size_t t = 0; t < threads; ++t)
        th.emplace_back([&, t]() { fn(n_reads * t / threads, n_reads * (t + 1) / threads, parts[t]); });
    for (auto& x : th) x.join();
}

}  // namespace

int main(int argc, char** argv) {
#if defined(MONI_CLI_MEMS)
    const bool mems = true;
    const char* what = "maximal exact matches";
#else
    const bool mems = false;
    const char* what = "matching statistics";
#endif
    std::string patterns, output, gpus_arg;
    size_t threads = 1;
    const std::string usage = "usage: " + std::string(argv[0]) +
                              " infile [-p patterns] [-o output] [-t threads] [-l len] [-q shaped_slp] [-g gpus]\n\nComputes the " + what +
                              " of the reads in the pattern against the reference index in infile (B200 implementation).\n";
    int c;
    while ((c = getopt(argc, argv, "l:hp:o:t:g:q")) != -1) {
        switch (c) {
            case 'p': patterns = optarg; break;
            case 'o': output = optarg; break;
            case 'l': (void)atoi(optarg); break;          // parsed and unused, like the reference
            case 't': threads = (size_t)std::max(1, atoi(optarg)); break;
            case 'g': gpus_arg = optarg; break;
            case 'q': moni_gpu_set_option("grammar", "shaped"); break;
            case 'h': die(usage);
            default: die("Unknown option.\n" + usage);
        }
    }
    if (argc != optind + 1) die("Invalid number of arguments\n" + usage);
    const std::string index = argv[optind];
    if (gpus_arg.empty() && getenv("MONI_GPUS")) gpus_arg = getenv("MONI_GPUS");
    std::vector<int> gpus;
    for (size_t i = 0; i < gpus_arg.size();) {
        size_t j = gpus_arg.find(',', i);
        if (j == std::string::npos) j = gpus_arg.size();
        gpus.push_back(atoi(gpus_arg.substr(i, j - i).c_str()));
        i = j + 1;
    }
    if (gpus.empty()) gpus.push_back(0);

    info("Construction of the matching statistics data structure");
    double t0 = now_s();
    moni_gpu_index* idx = nullptr;
    if (moni_gpu_index_load(index.c_str(), gpus.data(), (int)gpus.size(), &idx) != 0) die(moni_gpu_last_error());
    printf("[INFO] Elapsed time (s): %g\n", now_s() - t0);

    std::vector<char> ibuf(index.begin(), index.end());
    ibuf.push_back(0);
    std::string out_prefix = patterns + "_" + basename(ibuf.data());
    if (!output.empty()) out_prefix = output;

    FILE* f1 = nullptr;
    FILE* f2 = nullptr;
    if (mems) {
        f1 = fopen((out_prefix + ".mems").c_str(), "wb");
        if (!f1) die("open() file " + out_prefix + ".mems failed");
    } else {
        f1 = fopen((out_prefix + ".pointers").c_str(), "wb");
        if (!f1) die("open() file " + out_prefix + ".pointers failed");
        f2 = fopen((out_prefix + ".lengths").c_str(), "wb");
        if (!f2) die("open() file " + out_prefix + ".lengths failed");
    }

    info("Processing patterns");
    t0 = now_s();
    SeqReader rd(patterns);
    Batch B;
    std::string name, seq;
    std::vector<uint64_t> v1, v2, moff;
    std::vector<std::string> parts1, parts2;
    uint64_t n_reads = 0, n_bases = 0;
    double t_gpu = 0;
    const size_t kBatchBases = 256u << 20, kBatchReads = 4u << 20;
    bool more = true;
    while (more) {
        B.clear();
        while (B.seqs.size() < kBatchBases && B.names.size() < kBatchReads) {
            if (!rd.next(name, seq)) { more = false; break; }
            B.names.push_back(name);
            B.seqs += seq;
            B.offsets.push_back(B.seqs.size());
        }
        const size_t nr = B.names.size();
        if (!nr) break;
        n_reads += nr;
        n_bases += B.seqs.size();
        const size_t total = B.seqs.size();
        v1.resize(total + 1);
        v2.resize(total + 1);
        const double tg = now_s();
        if (mems) {
            moff.resize(nr + 1);
            if (moni_gpu_mems(idx, B.seqs.data(), B.offsets.data(), nr, moff.data(), v1.data(), v2.data(), total) != 0)
                die(moni_gpu_last_error());
            t_gpu += now_s() - tg;
            parallel_format(nr, threads, parts1, [&](size_t a, size_t b, std::string& o) { format_mems(B, moff.data(), v1.data(), v2.data(), a, b, o); });
            for (auto& p : parts1) fwrite(p.data(), 1, p.size(), f1);
        } else {
            if (moni_gpu_ms(idx, B.seqs.data(), B.offsets.data(), nr, v1.data(), v2.data()) != 0) die(moni_gpu_last_error());
            t_gpu += now_s() - tg;
            parallel_format(nr, threads, parts1, [&](size_t a, size_t b, std::string& o) { format_values(B, v1.data(), a, b, o); });
            parallel_format(nr, threads, parts2, [&](size_t a, size_t b, std::string& o) { format_values(B, v2.data(), a, b, o); });
            for (auto& p : parts1) fwrite(p.data(), 1, p.size(), f1);
            for (auto& p : parts2) fwrite(p.data(), 1, p.size(), f2);
        }
    }
    fclose(f1);
    if (f2) fclose(f2);
    printf("[INFO] Elapsed time (s): %g\n", now_s() - t0);
    printf("[INFO] reads %llu bases %llu gpu-call time (s) %g gpus %zu\n", (unsigned long long)n_reads, (unsigned long long)n_bases, t_gpu, gpus.size());
    moni_gpu_index_free(idx);
    return 0;
}
