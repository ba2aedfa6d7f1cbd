"""CPU suite, part 2: the product's host logic without a GPU.
  * the exporter (index reader -> HBM image) and the per-read core of the kernels, compiled for the CPU by
    tests/host_sim.cpp (TEST-ONLY build of moni_align_b200/csrc/ms_core.cuh), against the golden reference outputs;
  * libmoni_gpu.so loads, exports every symbol include/moni_gpu.h declares, and fails loudly without a device."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import CASES, GOLDEN, ROOT, pack, read_fastq, run_host_sim


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("opts", [(), (1, 2), (16, 3), (64, 4), (1000000, 64)])
def test_exporter_and_core_match_reference(case, opts, host_sim, oracle_mod, tmp_path):
    pre = os.path.join(GOLDEN, case)
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs)
    p, l, mo, mp, ml, cnt = run_host_sim(host_sim, pre + ".fa", flat, offs, tmp_path, opts)
    assert oracle_mod.format_ms(names, offs, p) == open(pre + ".pointers", "rb").read()
    assert oracle_mod.format_ms(names, offs, l) == open(pre + ".lengths", "rb").read()
    assert oracle_mod.format_mems(names, mo, mp, ml) == open(pre + ".mems", "rb").read()


@pytest.mark.parametrize("case", CASES)
def test_event_counters_match_oracle(case, host_sim, oracle_mod, tmp_path):
    """B, J, S, C of SURVEY.md 8(d) are properties of (index, reads): the image must count the same events."""
    pre = os.path.join(GOLDEN, case)
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs)
    ix = oracle_mod.OracleIndex(pre + ".fa")
    ix.counters(reset=True)
    ix.ms(flat, offs)
    want = ix.counters()
    cnt = run_host_sim(host_sim, pre + ".fa", flat, offs, tmp_path)[5]
    assert int(cnt[0]) == want["bases"]
    assert int(cnt[1]) == want["jumps"]
    assert int(cnt[2]) == want["restarts"]
    assert int(cnt[4]) == want["chars"]


def test_general_path_is_exercised(host_sim, tmp_path):
    pre = os.path.join(GOLDEN, "withN")
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs)
    cnt = run_host_sim(host_sim, pre + ".fa", flat, offs, tmp_path)[5]
    assert int(cnt[6]) > 100


def test_empty_and_ragged_batches(host_sim, oracle_mod, tmp_path):
    pre = os.path.join(GOLDEN, "pan4")
    ix = oracle_mod.OracleIndex(pre + ".fa")
    seqs = [b"", b"A", b"", b"ACGTACGTTTGA", b""]
    flat, offs = pack(seqs)
    p, l, mo, mp, ml, _ = run_host_sim(host_sim, pre + ".fa", flat, offs, tmp_path)
    wp, wl = ix.ms(flat, offs)
    wmo, wmp, wml = ix.mems(flat, offs)
    assert np.array_equal(p, wp) and np.array_equal(l, wl)
    assert np.array_equal(mo, wmo) and np.array_equal(mp, wmp) and np.array_equal(ml, wml)


def test_reader_rejects_garbage(host_sim, tmp_path):
    import subprocess
    bad = tmp_path / "bad"
    (tmp_path / "bad.thrbv.ms").write_bytes(b"\x00" * 100)
    (tmp_path / "bad.plain.slp").write_bytes(b"\x00" * 10)
    (tmp_path / "b.bin").write_bytes(np.zeros(2, dtype=np.uint64).tobytes())
    r = subprocess.run([host_sim, str(bad), str(tmp_path / "b.bin"), str(tmp_path / "o.bin")], capture_output=True)
    assert r.returncode == 1 and b"host_sim:" in r.stderr


# ------------------------------------------------------------------------------ the C ABI library
def _declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "moni_gpu.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(moni_gpu_[a-z_0-9]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    import moni_align_b200 as m
    L = m.load_library()
    syms = _declared_symbols()
    assert len(syms) >= 19
    for s in syms:
        assert hasattr(L, s), s
    assert sorted(m.EXPORTS) == syms


def test_library_has_no_cpu_fallback():
    """Without a CUDA device the product must fail loudly, never compute on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import moni_align_b200 as m
    with pytest.raises(m.MoniGpuError) as e:
        m.MoniIndex(os.path.join(GOLDEN, "pan4.fa"))
    assert e.value.code == -3
    assert "oracle" not in open(os.path.join(ROOT, "moni_align_b200", "__init__.py")).read().split('"""', 2)[2]


def test_missing_index_file_is_an_error_not_a_crash(host_sim, tmp_path):
    import subprocess
    (tmp_path / "b.bin").write_bytes(np.zeros(2, dtype=np.uint64).tobytes())
    r = subprocess.run([host_sim, str(tmp_path / "nope"), str(tmp_path / "b.bin"), str(tmp_path / "o.bin")], capture_output=True)
    assert r.returncode == 1 and b"cannot open" in r.stderr


# ------------------------------------------------------------------------------ `moni align` front end
@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("min_len,opts", [(0, (0, 8)), (25, (0, 8)), (0, (16, 3)), (25, (64, 4))])
def test_seed_core_matches_reference_find_mems(case, min_len, opts, host_sim, oracle_mod, tmp_path):
    """The kernels' per-read core in seeds mode (8-byte compares, trailing-N counter by bit tricks), on the image exported
    from the `moni align` index file alone, == the reference's seed_finder::find_mems output."""
    from conftest import full_only_prefix
    pre = os.path.join(GOLDEN, case)
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs)
    res = run_host_sim(host_sim, full_only_prefix(case, tmp_path), flat, offs, tmp_path, (*opts, min_len))
    assert oracle_mod.format_ms(names, offs, res[0]) == open(pre + ".pointers", "rb").read()
    so, ref, ln, qpos = res[6:10]
    assert oracle_mod.format_seeds(names, so, ref, ln, qpos) == open(pre + f".seeds{min_len}", "rb").read()


def test_seed_core_on_N_rich_reads(host_sim, oracle_mod, tmp_path):
    """Random reads over {A,C,G,T,N} with long N stretches against the N-rich text: the trailing-N counter must agree with
    the oracle's character-by-character restatement at every 8-byte compare boundary."""
    pre = os.path.join(GOLDEN, "withN")
    ix = oracle_mod.OracleIndex(pre + ".fa")
    text = bytes(ix.char_at(i) for i in range(ix.text_len))
    rng = np.random.default_rng(5)
    seqs = []
    for _ in range(200):
        a = int(rng.integers(0, len(text) - 200))
        s = bytearray(text[a:a + int(rng.integers(1, 200))])
        for _ in range(int(rng.integers(0, 4))):
            q = int(rng.integers(0, len(s)))
            s[q:q + int(rng.integers(1, 12))] = b"N" * min(int(rng.integers(1, 12)), len(s) - q)
        seqs.append(bytes(s))
    flat, offs = pack(seqs)
    for min_len in (0, 3, 20):
        res = run_host_sim(host_sim, pre + ".fa", flat, offs, tmp_path, (0, 8, min_len))
        so, ref, qpos, ln = ix.seeds(flat, offs, min_len)
        assert np.array_equal(res[6], so) and np.array_equal(res[7], ref) and np.array_equal(res[8], ln) and np.array_equal(res[9], qpos)


@pytest.mark.parametrize("case", CASES)
def test_occurrence_walk_matches_reference_find_MEM_occs(case, host_sim, oracle_mod, tmp_path):
    """occ_walk (flattened Phi / Phi^-1 tables, word-wise bounded LCE on the text) on the image exported from the reference's
    .thrbv.full.ms == the reference's find_MEM_occs lists (tests/golden/<case>.occs12)."""
    from conftest import full_only_prefix
    pre = os.path.join(GOLDEN, case)
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs)
    res = run_host_sim(host_sim, full_only_prefix(case, tmp_path), flat, offs, tmp_path, (1 << 30, 8, 12))   # leaf_len >= text: text mode
    assert len(res) == 12
    so, ref, ln, qpos, oo, occs = res[6:12]
    assert oracle_mod.format_occs(names, so, ref, ln, qpos, oo, occs) == open(pre + ".occs12", "rb").read()


def test_occurrence_walk_needs_the_align_index(host_sim, tmp_path):
    """From a plain .thrbv.ms there are no Phi structures: host_sim writes no occurrence section (the library returns EINVAL)."""
    pre = os.path.join(GOLDEN, "pan4")
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs[:5])
    assert len(run_host_sim(host_sim, pre + ".fa", flat, offs, tmp_path, (0, 8, 12))) == 10


@pytest.mark.parametrize("case", CASES)
def test_lcp_align_index_gives_the_same_seeds_and_occurrences(case, host_sim, oracle_mod, tmp_path):
    """.thrbv.full.lcp.ms (the default `moni align` index: one more trailing vector, the sampled LCPs) loads, and seeds and
    occurrence lists are the ones the reference's LCP-based enumeration prints (verified identical to .seeds25 / .occs12 when
    the fixture was generated, tests/golden/make_golden_seeds.py::lcp_variant)."""
    from conftest import full_only_prefix
    pre = os.path.join(GOLDEN, case)
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs)
    res = run_host_sim(host_sim, full_only_prefix(case, tmp_path, lcp=True), flat, offs, tmp_path, (1 << 30, 8, 12))
    so, ref, ln, qpos, oo, occs = res[6:12]
    assert oracle_mod.format_occs(names, so, ref, ln, qpos, oo, occs) == open(pre + ".occs12", "rb").read()
    d = oracle_mod.read_thrbv_ms(os.path.join(GOLDEN, case + ".fa.thrbv.full.lcp.ms"), full=True)
    assert d["slcp"].shape[0] == d["R"]


def test_row_layout_phi_step_and_packed_pointers_at_40_bits(tmp_path):
    """tests/unit_rows.cpp: RowRegs decodes every MoveRow field at its extremes (40-bit samples, u32 rows), phi_step is right
    near 2^40 including the circular predecessor, packed pointers round-trip wrapped values."""
    import subprocess
    exe = str(tmp_path / "unit_rows")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-x", "c++", os.path.join(ROOT, "tests", "unit_rows.cpp"), "-o", exe])
    r = subprocess.run([exe], capture_output=True)
    assert r.returncode == 0, r.stderr


@pytest.fixture(scope="module")
def slp_text_exe(tmp_path_factory):
    """tests/slp_text.cpp: the product's grammar readers + exporter, text written to stdout."""
    import subprocess
    exe = str(tmp_path_factory.mktemp("slp") / "slp_text")
    src = [os.path.join(ROOT, "tests", "slp_text.cpp"), os.path.join(ROOT, "moni_align_b200", "csrc", "index_files.cpp"),
           os.path.join(ROOT, "moni_align_b200", "csrc", "device_image.cpp")]
    subprocess.check_call(["g++", "-O2", "-std=c++17", *src, "-o", exe, "-lpthread"])
    return exe


@pytest.mark.parametrize("case", ["pan4", "withN", "twoletter"])
def test_shaped_grammar_decodes_to_the_reference_text(case, slp_text_exe):
    """<case>.fa.slp (SelfShapedSlp, written by the reference encoder) read by read_shaped_slp -- RecSplit length hash,
    gamma arrays, group tables -- expands to the text the reference's own SelfShapedSlp decoder produced
    (tests/golden/<case>.text.sha256, from oracle/ref_slp_text.cpp), in text mode and through the node table with
    64- and 16-byte literal blocks; the plain grammar of the same case expands to the same bytes."""
    import hashlib
    import subprocess
    want, n = open(os.path.join(GOLDEN, case + ".text.sha256")).read().split()
    shaped, plain = os.path.join(GOLDEN, case + ".fa.slp"), os.path.join(GOLDEN, case + ".fa.plain.slp")
    for leaf in ([], ["64"], ["16"]):
        out = subprocess.check_output([slp_text_exe, shaped, "shaped", *leaf], stderr=subprocess.DEVNULL)
        assert len(out) == int(n) and hashlib.sha256(out).hexdigest() == want
    out = subprocess.check_output([slp_text_exe, plain, "plain"], stderr=subprocess.DEVNULL)
    assert hashlib.sha256(out).hexdigest() == want


def test_shaped_grammar_reader_rejects_damaged_files(slp_text_exe, tmp_path):
    """Truncations and a flipped byte in each section end in an error exit (never a crash or a silent wrong text)."""
    import hashlib
    import subprocess
    src = open(os.path.join(GOLDEN, "pan4.fa.slp"), "rb").read()
    want = open(os.path.join(GOLDEN, "pan4.text.sha256")).read().split()[0]
    p = str(tmp_path / "bad.slp")
    for cut in (0, 7, 40, len(src) // 3, len(src) - 1):
        open(p, "wb").write(src[:cut])
        r = subprocess.run([slp_text_exe, p, "shaped"], capture_output=True)
        assert r.returncode == 1 and b"slp_text:" in r.stderr, cut
    rng = np.random.default_rng(5)
    for at in rng.integers(0, len(src), size=40):
        b = bytearray(src)
        b[at] ^= 0x5A
        open(p, "wb").write(bytes(b))
        r = subprocess.run([slp_text_exe, p, "shaped"], capture_output=True, timeout=60)
        assert r.returncode in (0, 1), (at, r.returncode)          # no signal
        if r.returncode == 0 and hashlib.sha256(r.stdout).hexdigest() != want:
            # damage the format cannot detect (a child offset, an alphabet byte): still a text of the stated length
            assert len(r.stdout) == 23992


@pytest.fixture(scope="module")
def cli_selftest(tmp_path_factory):
    """csrc/cli_main.cpp compiled with -DMONI_CLI_SELFTEST: its reader and number formatter without any GPU call."""
    import subprocess
    exe = str(tmp_path_factory.mktemp("cli") / "cli_selftest")
    pkg = os.path.join(ROOT, "moni_align_b200")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-DMONI_CLI_SELFTEST", os.path.join(pkg, "csrc", "cli_main.cpp"), "-o", exe,
                           "-L" + pkg, "-lmoni_gpu", "-lz", "-lpthread", "-Wl,-rpath," + pkg], stderr=subprocess.DEVNULL)
    return exe


def _parsed(exe, path):
    import subprocess
    out = subprocess.check_output([exe, "parse", str(path)])
    recs = [l.split(b"\t", 1) for l in out.split(b"\n")[:-1]]
    return [r[0] for r in recs], [r[1] for r in recs]


@pytest.mark.parametrize("case", CASES)
def test_cli_reader_parses_the_golden_fastq(case, cli_selftest):
    names, seqs = read_fastq(os.path.join(GOLDEN, case + ".fq"))
    got_names, got_seqs = _parsed(cli_selftest, os.path.join(GOLDEN, case + ".fq"))
    assert got_names == [n if isinstance(n, bytes) else n.encode() for n in names]
    assert got_seqs == [bytes(s) for s in seqs]


def test_cli_reader_kseq_corner_cases(cli_selftest, tmp_path):
    """Spans across the 1 MiB read buffer, CRLF, multi-line FASTA, a quality line that starts with '@', blank lines, a file
    without a final newline, gzip: the record boundaries of kseq.h (thirdparty/klib/kseq.h) that the reference reads with."""
    import gzip
    rng = np.random.default_rng(3)
    big = bytes(rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), size=3_300_000))
    cases = {
        "crlf.fq": (b"@a x\r\nACGT\r\n+\r\nIIII\r\n@b\r\nGG\r\n+\r\nII\r\n", [(b"a", b"ACGT"), (b"b", b"GG")]),
        "multi.fa": (b">s1 desc\tmore\nACG\nTTA\n\n>s2\nGG\n", [(b"s1", b"ACGTTA"), (b"s2", b"GG")]),
        "atqual.fq": (b"@q\nACGT\n+\n@III\n@r\nAC\n+q again\nII\n", [(b"q", b"ACGT"), (b"r", b"AC")]),
        "nonl.fq": (b"@x\nACGTN\n+\nIIIII", [(b"x", b"ACGTN")]),
        "nonl.fa": (b">x\nACGT\n>y\nTT", [(b"x", b"ACGT"), (b"y", b"TT")]),
        "big.fq": (b"@big\n" + big + b"\n+\n" + b"I" * len(big) + b"\n@after\nAC\n+\nII\n", [(b"big", big), (b"after", b"AC")]),
        "biglines.fa": (b">w\n" + b"\n".join(big[i:i + 70_001] for i in range(0, len(big), 70_001)) + b"\n>z\nA\n", [(b"w", big), (b"z", b"A")]),
        "empty.fq": (b"", []),
        "emptyseq.fa": (b">e\n>f\nAC\n", [(b"e", b""), (b"f", b"AC")]),
    }
    for name, (data, want) in cases.items():
        p = tmp_path / name
        p.write_bytes(data)
        names, seqs = _parsed(cli_selftest, p)
        assert list(zip(names, seqs)) == want, name
        gz = tmp_path / (name + ".gz")
        with gzip.open(gz, "wb") as f:
            f.write(data)
        names, seqs = _parsed(cli_selftest, gz)
        assert list(zip(names, seqs)) == want, name + ".gz"


def test_cli_number_formatter(cli_selftest):
    import subprocess
    rng = np.random.default_rng(11)
    vals = [0, 1, 9, 10, 11, 99, 100, 101, 999, 1000, 65535, 65536, 2**32 - 1, 2**32, 2**40 - 1, 2**40, 2**63, 2**64 - 1, 2**64 - 2]
    vals += [10**k for k in range(20)] + [10**k - 1 for k in range(1, 20)]
    vals += [int(x) for x in rng.integers(0, 2**63, size=2000)] + [int(x) >> int(s) for x, s in zip(rng.integers(0, 2**63, size=2000), rng.integers(0, 63, size=2000))]
    out = subprocess.run([cli_selftest, "u64"], input=" ".join(map(str, vals)).encode(), capture_output=True, check=True).stdout
    assert out.decode().split() == [str(v) for v in vals]


@pytest.mark.parametrize("case", CASES)
def test_exporter_image_does_not_depend_on_its_thread_count(case, host_sim, tmp_path):
    """build_move_table cuts its row passes into parts with prefix-computed initial states: the move table (rows, row starts,
    gap slots, run rows) must be the same bytes for 1, 2, 5 and one-row-per-part threads, with and without balancing cuts."""
    import subprocess
    b = tmp_path / "b.bin"
    b.write_bytes(np.uint64(1).tobytes() + np.array([0, 4], dtype=np.uint64).tobytes() + b"ACGT")
    for max_scan in (8, 2):
        seen = set()
        for th in (1, 2, 5, 100000):
            r = subprocess.run([host_sim, os.path.join(GOLDEN, case + ".fa"), str(b), str(tmp_path / "o.bin"), "64", str(max_scan)],
                               env=dict(os.environ, HOST_SIM_THREADS=str(th)), capture_output=True)
            assert r.returncode == 0, r.stderr
            seen.add(re.search(rb"move table hash ([0-9a-f]+) rows (\d+)", r.stderr).groups())
        assert len(seen) == 1, (max_scan, seen)


@pytest.mark.parametrize("case", CASES)
def test_runs_expand_rebuilds_the_reference_vectors(case):
    """moni_gpu_ms_runs_expand (host-side format conversion, no device) on the restarts of the reference's own .pointers /
    .lengths files gives those files back, wrapped pointers included; malformed runs are rejected."""
    import moni_align_b200 as m
    import refparse
    pre = os.path.join(GOLDEN, case)
    _, cnt, ptr = refparse.parse_values(pre + ".pointers")
    _, _, ln = refparse.parse_values(pre + ".lengths")
    offs = np.zeros(cnt.shape[0] + 1, dtype=np.uint64)
    np.cumsum(cnt, out=offs[1:])
    total = int(offs[-1])
    rst = np.ones(total, dtype=bool)
    with np.errstate(over="ignore"):
        rst[1:] = ptr[1:] != ptr[:-1] + np.uint64(1)
    rst[offs[:-1].astype(np.int64)] = True
    at = np.nonzero(rst)[0]
    read_of = np.searchsorted(offs, at, side="right") - 1
    ro = np.zeros(offs.shape[0], dtype=np.uint64)
    np.cumsum(np.bincount(read_of, minlength=cnt.shape[0]), out=ro[1:])
    qpos = (at - offs[read_of].astype(np.int64)).astype(np.uint32)
    p2, l2 = m.expand_runs(offs, ro, qpos, ln[at].astype(np.uint32), ptr[at])
    assert np.array_equal(p2, ptr) and np.array_equal(l2, ln)
    assert at.shape[0] < total                      # runs really are fewer than bases on these fixtures
    bad = qpos.copy()
    bad[0] = 1
    with pytest.raises(m.MoniGpuError):
        m.expand_runs(offs, ro, bad, ln[at].astype(np.uint32), ptr[at])


def test_host_widening_of_packed_results(tmp_path):
    """csrc/host_widen.cpp (moni_gpu_mems / moni_gpu_ms: results cross the link packed and host threads widen them into the u64
    columns): every record format, misaligned destinations, pool sizes 0..5, wrapped 40-bit pointers -- compiled with g++ and run here."""
    import subprocess
    exe = str(tmp_path / "host_widen_test")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-pthread", os.path.join(ROOT, "tests", "host_widen_test.cpp"),
                           os.path.join(ROOT, "moni_align_b200", "csrc", "host_widen.cpp"), "-o", exe])
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0 and out.stdout.strip() == "ok", out.stdout + out.stderr


def test_header_is_plain_c(tmp_path):
    """include/moni_gpu.h is the drop-in boundary: it must compile as strict C99 (plain pointers and sizes, no C++ or torch types), and
    a C program must link against the library with nothing but the header."""
    import subprocess
    src = tmp_path / "abi.c"
    src.write_text('#include "moni_gpu.h"\n#include <stdio.h>\n'
                   'int main(void) { moni_gpu_index* ix = 0; int rc = moni_gpu_index_load("/nonexistent/prefix", 0, 1, &ix);\n'
                   '  printf("%d %s\\n", rc, moni_gpu_version()); return rc == 0; }\n')
    exe = str(tmp_path / "abi")
    lib_dir = os.path.join(ROOT, "moni_align_b200")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"), str(src), "-o", exe,
                           "-L", lib_dir, "-lmoni_gpu", "-Wl,-rpath," + lib_dir])
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr          # the load fails (no such index, or no device here): an error code, not a crash
    assert out.stdout.split()[0].startswith("-") and "moni-gpu" in out.stdout
