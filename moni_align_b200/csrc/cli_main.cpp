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
#include <fcntl.h>
#include <getopt.h>
#include <libgen.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/moni_gpu.h"

namespace {

// Output files opened so far: a failed run removes them instead of leaving truncated results behind.
std::mutex g_partial_mu;
std::vector<std::string> g_partial;

[[noreturn]] void die(const std::string& msg) {
    // the reference's error(): "[ERROR] ..." on stderr, exit(1)   (include/common/common.hpp:108-117).  Any pipeline thread may
    // end up here while the others are inside CUDA calls or pwrite, so the process leaves through _exit: no static destructors
    // and no CUDA teardown run beside live work.
    fprintf(stderr, "[ERROR] %s\n", msg.c_str());
    fflush(stderr);
    fflush(stdout);
    {
        std::lock_guard<std::mutex> lk(g_partial_mu);
        for (const std::string& p : g_partial) unlink(p.c_str());
    }
    _exit(1);
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
    std::string header;
    bool eof = false;
    explicit SeqReader(const std::string& path) : buf(1 << 20) {
        f = gzopen(path.c_str(), "r");
        if (!f) die("cannot open patterns file " + path);
        gzbuffer(f, 1 << 20);
    }
    ~SeqReader() { if (f) gzclose(f); }
    bool fill() {
        if (eof) return false;
        const int k = gzread(f, buf.data(), (unsigned)buf.size());
        if (k <= 0) { eof = true; return false; }
        len = (size_t)k;
        pos = 0;
        return true;
    }
    int getc() {
        if (pos >= len && !fill()) return -1;
        return buf[pos++];
    }
    // appends the rest of the current line (without the newline, carriage returns dropped) to out; returns the
    // delimiter (-1 at EOF).  Whole buffer spans at a time: the parser runs at memchr speed, not a call per byte.
    int line(std::string* out) {
        for (;;) {
            if (pos >= len && !fill()) return -1;
            const unsigned char* s = buf.data() + pos;
            const size_t avail = len - pos;
            const unsigned char* nl = (const unsigned char*)memchr(s, '\n', avail);
            const size_t k = nl ? (size_t)(nl - s) : avail;
            if (out && k) {
                if (!memchr(s, '\r', k)) out->append((const char*)s, k);
                else for (size_t i = 0; i < k; ++i) if (s[i] != '\r') out->push_back((char)s[i]);
            }
            pos += k + (nl ? 1 : 0);
            if (nl) return '\n';
        }
    }
    // skips the rest of the current line; n = its characters other than carriage returns
    int skip_line(size_t& n) {
        n = 0;
        for (;;) {
            if (pos >= len && !fill()) return -1;
            const unsigned char* s = buf.data() + pos;
            const size_t avail = len - pos;
            const unsigned char* nl = (const unsigned char*)memchr(s, '\n', avail);
            const size_t k = nl ? (size_t)(nl - s) : avail;
            n += k;
            if (k && memchr(s, '\r', k)) for (size_t i = 0; i < k; ++i) n -= s[i] == '\r';
            pos += k + (nl ? 1 : 0);
            if (nl) return '\n';
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
        header.clear();
        line(&header);
        const size_t sp = header.find_first_of(" \t");
        name.assign(header, 0, sp);
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
        while (q < seq.size()) {                          // quality: as many characters as bases (counted, not kept)
            size_t k = 0;
            const int d = skip_line(k);
            q += k;
            if (d < 0) break;
        }
        last = 0;
        // kseq_read returns -2 for a quality string shorter than the sequence (thirdparty/klib/kseq.h) and the reference's read
        // loop stops there (`while (kseq_read(seq) >= 0)`): the truncated record is not processed
        if (q < seq.size()) { eof = true; return false; }
        return true;
    }
};

// integer -> decimal ASCII, two digits per division
inline char* put_u64(char* p, uint64_t v) {
    static const char kPairs[201] =
        "00010203040506070809101112131415161718192021222324252627282930313233343536373839"
        "40414243444546474849505152535455565758596061626364656667686970717273747576777879"
        "8081828384858687888990919293949596979899";
    char tmp[24];
    int k = 24;
    while (v >= 100) {
        const uint64_t q = v / 100;
        const unsigned r = (unsigned)(v - q * 100);
        v = q;
        tmp[--k] = kPairs[2 * r + 1];
        tmp[--k] = kPairs[2 * r];
    }
    if (v >= 10) { tmp[--k] = kPairs[2 * v + 1]; tmp[--k] = kPairs[2 * v]; }
    else tmp[--k] = (char)('0' + v);
    memcpy(p, tmp + k, 24 - k);
    return p + (24 - k);
}

// A batch travelling through the three stages (parse -> GPU call -> format + write).  Sequence and result buffers are
// pinned so the library's copies run at PCIe speed; results come back in the packed layouts (ms: 9 bytes per restart run,
// expanded to the reference's two vectors by the formatter threads; mems: 4 bytes per MEM) unless a read of 65536 bases or
// more forces the wide calls.
struct Slot {
    std::string names;                       // name bytes back to back
    std::vector<uint32_t> name_end{0};       // n_reads + 1
    std::vector<uint64_t> offsets{0};        // n_reads + 1
    char* seqs = nullptr;                    // pinned
    uint64_t cap = 0;                        // bases the pinned buffers hold
    uint32_t* lo = nullptr;                  // ms: run records (query offset << 16 | length); mems: packed records   (pinned, cap)
    uint8_t* hi = nullptr;                   // ms: pointer bits 32..39 of every run                                   (pinned, cap)
    uint32_t* plo = nullptr;                 // ms: pointer low words of every run                                     (pinned, cap)
    std::vector<uint64_t> moff;              // n_reads + 1: record (mems) or run (ms) offsets
    std::vector<uint64_t> w1, w2;            // wide fallback (reads of >= 65536 bases)
    uint64_t max_len = 0;
    bool wide = false, last = false;
    size_t n_reads() const { return offsets.size() - 1; }
    uint64_t total() const { return offsets.back(); }
    void clear() { names.clear(); name_end.assign(1, 0); offsets.assign(1, 0); max_len = 0; wide = false; last = false; }
    void release() {
        moni_gpu_host_free(seqs); moni_gpu_host_free(lo); moni_gpu_host_free(hi); moni_gpu_host_free(plo);
        seqs = nullptr; lo = nullptr; hi = nullptr; plo = nullptr; cap = 0;
    }
    // grows the pinned buffers to hold `bases` (keeps the sequence bytes already there)
    void ensure(uint64_t bases, bool mems, bool fatal = true) {
        if (bases <= cap) return;
        // doubling up to one full batch (+ room for the read that crosses the limit); beyond that only a single huge read grows it
        const uint64_t soft = (64u << 20) + (1u << 20);
        const uint64_t ncap = bases <= soft ? std::min<uint64_t>(std::max<uint64_t>(cap * 2, bases), soft) : std::max<uint64_t>(bases, cap * 2);
        char* ns = (char*)moni_gpu_host_alloc(ncap + 16);
        if (!ns && !fatal) return;                 // pre-allocation beside the index load: the load reports what is wrong
        if (!ns) die("cannot allocate pinned host memory for the read batch");
        if (seqs) memcpy(ns, seqs, total());
        moni_gpu_host_free(seqs); moni_gpu_host_free(lo); moni_gpu_host_free(hi); moni_gpu_host_free(plo);
        seqs = ns;
        lo = (uint32_t*)moni_gpu_host_alloc((ncap + 4) * 4);
        hi = mems ? nullptr : (uint8_t*)moni_gpu_host_alloc(ncap + 16);
        plo = mems ? nullptr : (uint32_t*)moni_gpu_host_alloc((ncap + 4) * 4);
        if (!lo || (!mems && (!hi || !plo))) {
            if (!fatal) { cap = 0; moni_gpu_host_free(lo); moni_gpu_host_free(hi); moni_gpu_host_free(plo); moni_gpu_host_free(seqs);
                          lo = nullptr; hi = nullptr; plo = nullptr; seqs = nullptr; return; }
            die("cannot allocate pinned host memory for the results");
        }
        cap = ncap;
    }
    void add(const std::string& name, const std::string& seq, bool mems) {
        ensure(total() + seq.size(), mems);
        memcpy(seqs + total(), seq.data(), seq.size());
        names += name;
        name_end.push_back((uint32_t)names.size());
        offsets.push_back(total() + seq.size());
        max_len = std::max<uint64_t>(max_len, seq.size());
    }
};

inline uint64_t unpack_pointer(uint32_t lo, uint8_t hi) {        // include/moni_gpu.h, moni_gpu_ms_packed
    uint64_t v = (uint64_t)lo | (uint64_t)hi << 32;
    if (v >= (1ULL << 40) - (1ULL << 31)) v -= 1ULL << 40;       // the reference's wrapped pointers 2^64 - k
    return v;
}

inline void put_header(const Slot& B, size_t r, std::string& out) {
    out.push_back('>');
    out.append(B.names, B.name_end[r], B.name_end[r + 1] - B.name_end[r]);
    out.push_back('\n');
}

// ms text of reads [a, b): ">name\n" + "v v v \n" per read (src/matching_statistics.cpp:588-598); which = 0 pointers, 1 lengths
void format_ms(const Slot& B, int which, size_t a, size_t b, std::string& out) {
    out.clear();
    std::vector<char> tmp;
    for (size_t r = a; r < b; ++r) {
        const uint64_t o = B.offsets[r], m = B.offsets[r + 1] - o;
        tmp.resize(m * 21 + 8);
        char* p = tmp.data();
        if (B.wide) {
            const uint64_t* v = (which ? B.w2.data() : B.w1.data()) + o;
            for (uint64_t i = 0; i < m; ++i) { p = put_u64(p, v[i]); *p++ = ' '; }
        } else {
            // restart runs -> the reference's vectors (include/moni_gpu.h, moni_gpu_ms_runs): inside a run the pointer grows by
            // one and the length shrinks by one, floor 0 (src/matching_statistics.cpp:246-255)
            for (uint64_t k = B.moff[r], ke = B.moff[r + 1]; k < ke; ++k) {
                const uint64_t i0 = B.lo[k] >> 16, i1 = k + 1 < ke ? B.lo[k + 1] >> 16 : m;
                if (which) {
                    const uint64_t l0 = B.lo[k] & 0xFFFF;
                    for (uint64_t i = i0; i < i1; ++i) { p = put_u64(p, l0 > i - i0 ? l0 - (i - i0) : 0); *p++ = ' '; }
                } else {
                    const uint64_t p0 = unpack_pointer(B.plo[k], B.hi[k]);
                    for (uint64_t i = i0; i < i1; ++i) { p = put_u64(p, p0 + (i - i0)); *p++ = ' '; }
                }
            }
        }
        put_header(B, r, out);
        out.append(tmp.data(), p - tmp.data());
        out.push_back('\n');
    }
}

// mems text: ">name\n" + "(pos,len) " per record + "\n" (src/mems.cpp:585-587)
void format_mems(const Slot& B, size_t a, size_t b, std::string& out) {
    out.clear();
    std::vector<char> tmp;
    for (size_t r = a; r < b; ++r) {
        const uint64_t s = B.moff[r], e = B.moff[r + 1];
        tmp.resize((e - s) * 46 + 8);
        char* p = tmp.data();
        for (uint64_t k = s; k < e; ++k) {
            const uint64_t pos = B.wide ? B.w1[k] : B.lo[k] >> 16, len = B.wide ? B.w2[k] : B.lo[k] & 0xFFFF;
            *p++ = '(';
            p = put_u64(p, pos);
            *p++ = ',';
            p = put_u64(p, len);
            *p++ = ')';
            *p++ = ' ';
        }
        put_header(B, r, out);
        out.append(tmp.data(), p - tmp.data());
        out.push_back('\n');
    }
}

// formats a batch with `threads` workers, then every worker writes its own part at its final file offset
struct Writer {
    int fd = -1;
    uint64_t at = 0;
    std::vector<std::string> parts;
    void open_file(const std::string& path) {
        fd = ::open(path.c_str(), O_WRONLY | O_CREAT | O_TRUNC, 0644);
        if (fd < 0) die("open() file " + path + " failed");
        std::lock_guard<std::mutex> lk(g_partial_mu);
        g_partial.push_back(path);
    }
    void close_file() { if (fd >= 0) ::close(fd); fd = -1; }
    template <typename F>
    void run(const Slot& B, size_t threads, F&& fn) {
        const size_t nr = B.n_reads();
        // split by bases so that workers get equal amounts of text
        threads = std::max<size_t>(1, std::min(threads, nr ? nr : 1));
        parts.resize(threads);
        std::vector<size_t> cut(threads + 1, nr);
        cut[0] = 0;
        for (size_t t = 1; t < threads; ++t) {
            const uint64_t want = B.total() * t / threads;
            cut[t] = (size_t)(std::lower_bound(B.offsets.begin(), B.offsets.end(), want) - B.offsets.begin());
            cut[t] = std::min(std::max(cut[t], cut[t - 1]), nr);
        }
        std::vector<std::thread> th;
        for (size_t t = 0; t < threads; ++t) th.emplace_back([&, t]() { fn(cut[t], cut[t + 1], parts[t]); });
        for (auto& x : th) x.join();
        th.clear();
        std::vector<uint64_t> where(threads);
        for (size_t t = 0; t < threads; ++t) { where[t] = at; at += parts[t].size(); }
        std::atomic<bool> failed{false};
        for (size_t t = 0; t < threads; ++t)
            th.emplace_back([&, t]() {
                const char* p = parts[t].data();
                uint64_t left = parts[t].size(), off = where[t];
                while (left) {
                    const ssize_t k = pwrite(fd, p, left, (off_t)off);
                    if (k <= 0) { failed = true; return; }
                    p += k; left -= (uint64_t)k; off += (uint64_t)k;
                }
            });
        for (auto& x : th) x.join();
        if (failed) die("write failed");
    }
};

}  // namespace

#if defined(MONI_CLI_SELFTEST)
// Host-only check of the reader and the number formatter (tests/test_host_logic.py):
//   parse FILE  -> one line "name<TAB>sequence" per record;   u64 -> every whitespace-separated number of stdin through put_u64
int main(int argc, char** argv) {
    if (argc >= 3 && !strcmp(argv[1], "parse")) {
        SeqReader rd(argv[2]);
        std::string name, seq;
        while (rd.next(name, seq)) {
            fwrite(name.data(), 1, name.size(), stdout);
            fputc('\t', stdout);
            fwrite(seq.data(), 1, seq.size(), stdout);
            fputc('\n', stdout);
        }
        return 0;
    }
    if (argc >= 2 && !strcmp(argv[1], "u64")) {
        unsigned long long v;
        char buf[32];
        while (scanf("%llu", &v) == 1) {
            char* e = put_u64(buf, v);
            *e++ = '\n';
            fwrite(buf, 1, e - buf, stdout);
        }
        return 0;
    }
    return 2;
}
#else
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

    // Batch slots of the processing pipeline (below).  Their pinned buffers are sized once from the file size (a FASTQ holds
    // about one base per two bytes, a gzip four times that) and page-locked on a helper thread while the index loads:
    // page-locking is slow enough that doing it batch by batch would show in the parse stage.
    constexpr int kSlots = 3;
    const uint64_t kBatchBases = 64u << 20, kBatchReads = 2u << 20;
    Slot slots[kSlots];
    uint64_t first_cap = 1u << 20;
    {
        struct stat st;
        if (stat(patterns.c_str(), &st) == 0) {
            const bool gz = patterns.size() > 3 && patterns.compare(patterns.size() - 3, 3, ".gz") == 0;
            first_cap = std::min<uint64_t>(kBatchBases + (1u << 20), std::max<uint64_t>(first_cap, (uint64_t)st.st_size * (gz ? 4 : 1)));
        }
    }
    // a bad -p path is reported now, not after seconds of index loading (an empty path reads nothing, like the reference)
    if (!patterns.empty() && access(patterns.c_str(), R_OK) != 0) die("cannot open patterns file " + patterns);
    const bool early = !getenv("MONI_CLI_NO_PREALLOC");
    std::thread prealloc([&]() {
        if (early) for (auto& sl : slots) sl.ensure(first_cap, mems, false);
    });

    info("Construction of the matching statistics data structure");
    double t0 = now_s();
    moni_gpu_index* idx = nullptr;
    const int load_rc = moni_gpu_index_load(index.c_str(), gpus.data(), (int)gpus.size(), &idx);
    prealloc.join();
    if (load_rc != 0) die(moni_gpu_last_error());
    printf("[INFO] Elapsed time (s): %g\n", now_s() - t0);

    std::vector<char> ibuf(index.begin(), index.end());
    ibuf.push_back(0);
    std::string out_prefix = patterns + "_" + basename(ibuf.data());
    if (!output.empty()) out_prefix = output;

    Writer out1, out2;
    if (mems) {
        out1.open_file(out_prefix + ".mems");
    } else {
        out1.open_file(out_prefix + ".pointers");
        out2.open_file(out_prefix + ".lengths");
    }

    info("Processing patterns");
    t0 = now_s();
    const double t_proc0 = t0;
    // Three stages on three threads, batches of ~64 M bases in a ring of slots: while the GPUs work on one batch the next
    // is being parsed and the previous one formatted and written (the reference splits the file and runs a worker per
    // part instead, src/matching_statistics.cpp:311-426; records come out in input order either way).
    int state[kSlots] = {0, 0, 0};            // 0 = free, 1 = parsed, 2 = results ready
    std::mutex mu;
    std::condition_variable cv;
    auto wait_for = [&](int slot, int want) {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&]() { return state[slot] == want; });
    };
    auto set_state = [&](int slot, int v) {
        { std::lock_guard<std::mutex> lk(mu); state[slot] = v; }
        cv.notify_all();
    };
    uint64_t n_reads = 0, n_bases = 0;
    double t_parse = 0, t_gpu = 0, t_out = 0;

    std::thread parser([&]() {
        SeqReader rd(patterns);
        std::string name, seq;
        bool more = true;
        for (int k = 0; more; k = (k + 1) % kSlots) {
            wait_for(k, 0);
            const double ts = now_s();
            Slot& B = slots[k];
            B.clear();
            B.ensure(first_cap, mems);
            while (B.total() < kBatchBases && B.n_reads() < kBatchReads) {
                if (!rd.next(name, seq)) { more = false; break; }
                B.add(name, seq, mems);
            }
            B.last = !more;
            t_parse += now_s() - ts;
            set_state(k, 1);
        }
    });
    std::thread caller([&]() {
        for (int k = 0;; k = (k + 1) % kSlots) {
            wait_for(k, 1);
            Slot& B = slots[k];
            const size_t nr = B.n_reads();
            const uint64_t total = B.total();
            const double ts = now_s();
            if (nr) {
                B.wide = B.max_len >= 65536;
                int rc = 0;
                if (mems) {
                    B.moff.resize(nr + 1);
                    if (!B.wide) rc = moni_gpu_mems_packed(idx, B.seqs, B.offsets.data(), nr, B.moff.data(), B.lo, total);
                    else {
                        B.w1.resize(total + 1); B.w2.resize(total + 1);
                        rc = moni_gpu_mems(idx, B.seqs, B.offsets.data(), nr, B.moff.data(), B.w1.data(), B.w2.data(), total);
                    }
                } else {
                    if (!B.wide) {
                        B.moff.resize(nr + 1);
                        rc = moni_gpu_ms_runs_packed(idx, B.seqs, B.offsets.data(), nr, B.moff.data(), B.lo, B.plo, B.hi, total);
                        if (rc == MONI_GPU_EINVAL) B.wide = true;          // text too long for 40-bit packed pointers
                    }
                    if (B.wide) {
                        B.w1.resize(total + 1); B.w2.resize(total + 1);
                        rc = moni_gpu_ms(idx, B.seqs, B.offsets.data(), nr, B.w1.data(), B.w2.data());
                    }
                }
                if (rc != 0) die(moni_gpu_last_error());
            }
            t_gpu += now_s() - ts;
            const bool last = B.last;
            set_state(k, 2);
            if (last) break;
        }
    });
    for (int k = 0;; k = (k + 1) % kSlots) {
        wait_for(k, 2);
        Slot& B = slots[k];
        const double ts = now_s();
        if (getenv("MONI_GPU_VERBOSE")) fprintf(stderr, "[moni-gpu] cli batch ready at %.3f s: %zu reads\n", ts - t_proc0, B.n_reads());
        if (B.n_reads()) {
            n_reads += B.n_reads();
            n_bases += B.total();
            if (mems) {
                out1.run(B, threads, [&](size_t a, size_t b, std::string& o) { format_mems(B, a, b, o); });
            } else {
                out1.run(B, threads, [&](size_t a, size_t b, std::string& o) { format_ms(B, 0, a, b, o); });
                out2.run(B, threads, [&](size_t a, size_t b, std::string& o) { format_ms(B, 1, a, b, o); });
            }
        }
        t_out += now_s() - ts;
        const bool last = B.last;
        set_state(k, 0);
        if (last) break;
    }
    const double t_tail0 = now_s();
    parser.join();
    caller.join();
    out1.close_file();
    out2.close_file();
    const double t_tail1 = now_s();
    for (auto& sl : slots) sl.release();
    if (getenv("MONI_GPU_VERBOSE")) fprintf(stderr, "[moni-gpu] cli tail: join + close %.3f s, unpin %.3f s\n", t_tail1 - t_tail0, now_s() - t_tail1);
    printf("[INFO] Elapsed time (s): %g\n", now_s() - t0);
    printf("[INFO] reads %llu bases %llu gpu-call time (s) %g parse (s) %g format+write (s) %g gpus %zu\n", (unsigned long long)n_reads,
           (unsigned long long)n_bases, t_gpu, t_parse, t_out, gpus.size());
    moni_gpu_index_free(idx);
    return 0;
}
#endif
