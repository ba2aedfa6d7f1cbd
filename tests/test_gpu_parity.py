"""GPU suite (-m gpu): the CUDA path, called through the C ABI (include/moni_gpu.h via ctypes), against
(a) the committed outputs of the reference binaries, (b) the CPU oracle on seeded inputs, and
(c) size-independent properties at larger sizes.  Bit-exact: everything here is unsigned integers."""
import os

import numpy as np
import pytest

from conftest import CASES, GOLDEN, ROOT, pack, read_fastq

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def m():
    import moni_align_b200 as mod
    mod.load_library()
    return mod


@pytest.fixture(scope="module")
def indexes(m):
    return {c: m.MoniIndex(os.path.join(GOLDEN, c + ".fa")) for c in CASES}


@pytest.mark.parametrize("case", CASES)
def test_ms_and_mems_equal_reference_files(case, m, indexes):
    pre = os.path.join(GOLDEN, case)
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs)
    ix = indexes[case]
    ptr, ln = ix.matching_statistics(flat, offs)
    assert m.format_ms(names, offs, ptr) == open(pre + ".pointers", "rb").read()
    assert m.format_ms(names, offs, ln) == open(pre + ".lengths", "rb").read()
    mo, mp, ml = ix.maxrimal_exact_matches(flat, offs)
    assert m.format_mems(names, mo, mp, ml) == open(pre + ".mems", "rb").read()


@pytest.fixture(scope="module")
def grammar_indexes(m):
    """The same indexes with the grammar kept (literal blocks of 16 bytes) instead of the whole text in HBM."""
    m.set_option("leaf_len", 16)
    try:
        return {c: m.MoniIndex(os.path.join(GOLDEN, c + ".fa")) for c in CASES}
    finally:
        m.set_option("leaf_len", 0)


@pytest.mark.parametrize("case", CASES)
def test_grammar_mode_equals_reference_files(case, m, grammar_indexes):
    pre = os.path.join(GOLDEN, case)
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs)
    ix = grammar_indexes[case]
    ptr, ln = ix.matching_statistics(flat, offs)
    assert m.format_ms(names, offs, ptr) == open(pre + ".pointers", "rb").read()
    assert m.format_ms(names, offs, ln) == open(pre + ".lengths", "rb").read()
    mo, mp, ml = ix.maxrimal_exact_matches(flat, offs)
    assert m.format_mems(names, mo, mp, ml) == open(pre + ".mems", "rb").read()


@pytest.mark.parametrize("case", CASES)
def test_random_reads_equal_oracle(case, m, indexes, oracle_mod):
    import synth
    pre = os.path.join(GOLDEN, case + ".fa")
    orc = oracle_mod.OracleIndex(pre)
    text = np.array([orc.char_at(i) for i in range(orc.text_len)], dtype=np.uint8)
    reads = [r for r in synth.short_reads(text, 3000, 100, 77, err=0.03)]
    reads += synth.long_reads(text, 40, 78, mean=1500, sd=600, lo=100, hi=4000)
    rng = np.random.default_rng(5)
    for _ in range(200):   # noise: arbitrary bytes incl. absent symbols
        reads.append(rng.integers(1, 127, size=int(rng.integers(1, 60)), dtype=np.uint8))
    flat, offs = m.pack_reads(reads)
    ptr, ln = indexes[case].matching_statistics(flat, offs)
    wp, wl = orc.ms(flat, offs)
    assert np.array_equal(ptr, wp)
    assert np.array_equal(ln, wl)
    mo, mp, ml = indexes[case].maxrimal_exact_matches(flat, offs)
    wmo, wmp, wml = orc.mems(flat, offs)
    assert np.array_equal(mo, wmo) and np.array_equal(mp, wmp) and np.array_equal(ml, wml)


def test_empty_ragged_and_zero_read_batches(m, indexes, oracle_mod):
    ix = indexes["pan4"]
    orc = oracle_mod.OracleIndex(os.path.join(GOLDEN, "pan4.fa"))
    flat, offs = pack([b"", b"A", b"", b"ACGTACGTTTGA", b""])
    ptr, ln = ix.matching_statistics(flat, offs)
    wp, wl = orc.ms(flat, offs)
    assert np.array_equal(ptr, wp) and np.array_equal(ln, wl)
    mo, mp, ml = ix.maxrimal_exact_matches(flat, offs)
    wmo, wmp, wml = orc.mems(flat, offs)
    assert np.array_equal(mo, wmo) and np.array_equal(mp, wmp) and np.array_equal(ml, wml)
    z = np.zeros(1, dtype=np.uint64)
    mo, mp, ml = ix.maxrimal_exact_matches(np.zeros(0, dtype=np.uint8), z)
    assert mo.tolist() == [0] and mp.size == 0


def test_chunking_and_resident_batches_agree(m, oracle_mod):
    """Small device chunks (many launches) and the resident-batch API give the same bits as one big call."""
    import synth
    pre = os.path.join(GOLDEN, "pan4.fa")
    orc = oracle_mod.OracleIndex(pre)
    text = np.array([orc.char_at(i) for i in range(orc.text_len)], dtype=np.uint8)
    reads = synth.short_reads(text, 5000, 120, 9)
    flat, offs = m.pack_reads(reads)
    m.set_option("chunk_bases", 7001)
    m.set_option("chunk_min_reads", 1)
    try:
        small = m.MoniIndex(pre)
    finally:
        m.set_option("chunk_bases", 1 << 24)
        m.set_option("chunk_min_reads", 96 * 1024)
    big = m.MoniIndex(pre)
    a = small.matching_statistics(flat, offs)
    b = big.matching_statistics(flat, offs)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    am = small.maxrimal_exact_matches(flat, offs)
    bm = big.maxrimal_exact_matches(flat, offs)
    assert all(np.array_equal(x, y) for x, y in zip(am, bm))
    rb = big.resident_batch(flat, offs)
    t = rb.run("ms")
    assert t["pass1_ms"] > 0 and t["pass2_ms"] > 0
    c = rb.fetch_ms()
    assert np.array_equal(c[0], b[0]) and np.array_equal(c[1], b[1])
    rb.run("mems")
    cm = rb.fetch_mems()
    assert all(np.array_equal(x, y) for x, y in zip(cm, bm))
    wp, wl = orc.ms(flat, offs)
    assert np.array_equal(b[0], wp) and np.array_equal(b[1], wl)
    rb.close()


def test_mems32_equals_mems(m, indexes):
    """moni_gpu_mems32 returns the same records in 32-bit fields."""
    names, seqs = read_fastq(os.path.join(GOLDEN, "pan4.fq"))
    flat, offs = pack(seqs)
    a = indexes["pan4"].maxrimal_exact_matches(flat, offs)
    b = indexes["pan4"].maxrimal_exact_matches(flat, offs, dtype=np.uint32)
    assert b[1].dtype == np.uint32 and b[2].dtype == np.uint32
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])


@pytest.mark.parametrize("case", ["pan4", "withN"])
def test_sharded_over_replicas_keeps_input_order(case, m, indexes, oracle_mod):
    """The multi-GPU path (one replica per listed device, reads split by bases, records copied to their global place once
    the earlier shards' totals are known).  Listing device 0 three times runs it on a one-GPU box."""
    import synth
    pre = os.path.join(GOLDEN, case + ".fa")
    orc = oracle_mod.OracleIndex(pre)
    text = np.array([orc.char_at(i) for i in range(orc.text_len)], dtype=np.uint8)
    reads = list(synth.short_reads(text, 2000, 90, 11, err=0.02)) + synth.long_reads(text, 30, 12, mean=900, sd=400, lo=50, hi=2500)
    flat, offs = m.pack_reads(reads)
    three = m.MoniIndex(pre, gpus=[0, 0, 0])
    one = indexes[case]
    a, b = one.matching_statistics(flat, offs), three.matching_statistics(flat, offs)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    for dt in (np.uint64, np.uint32):
        am, bm = one.maxrimal_exact_matches(flat, offs, dtype=dt), three.maxrimal_exact_matches(flat, offs, dtype=dt)
        assert all(np.array_equal(x, y) for x, y in zip(am, bm))
    wmo, wmp, wml = orc.mems(flat, offs)
    assert np.array_equal(bm[0], wmo) and np.array_equal(bm[1], wmp) and np.array_equal(bm[2], wml)
    # fewer reads than shards: some shards are empty
    f2, o2 = pack([b"ACGTTGCAAC", b""])
    x, y = one.maxrimal_exact_matches(f2, o2), three.maxrimal_exact_matches(f2, o2)
    assert all(np.array_equal(p, q) for p, q in zip(x, y))
    three.close()


def test_mem_capacity_error(m, indexes):
    ix = indexes["pan4"]
    flat, offs = pack([b"ACGTACGTACGTTTTGACCA"] * 50)
    mo = np.zeros(51, dtype=np.uint64)
    small = np.zeros(3, dtype=np.uint64)
    with pytest.raises(m.MoniGpuError) as e:
        ix.maxrimal_exact_matches(flat, offs, out=(mo, small, small.copy()))
    assert e.value.code == -4
    assert int(mo[50]) > 3          # the needed capacity is reported


def test_counters_match_oracle(m, oracle_mod):
    import synth
    pre = os.path.join(GOLDEN, "withN.fa")
    orc = oracle_mod.OracleIndex(pre)
    names, seqs = read_fastq(os.path.join(GOLDEN, "withN.fq"))
    flat, offs = pack(seqs)
    ix = m.MoniIndex(pre)
    m.set_option("count", 1)
    try:
        ix.counters(reset=True)
        orc.counters(reset=True)
        ix.matching_statistics(flat, offs)
    finally:
        m.set_option("count", 0)
    orc.ms(flat, offs)
    got, want = ix.counters(), orc.counters()
    for k in ("bases", "jumps", "restarts", "chars"):
        assert got[k] == want[k], k
    assert got["general_steps"] > 0


def test_properties_on_a_larger_index(m, oracle_mod, tmp_path):
    """Size-independent properties on a 2 Mbp x 4 pangenome built on the fly with the reference builder:
    T[ptr:ptr+len] == read[i:i+len], right-maximality against a suffix-array-free spot check, len[i+1] >= len[i]-1,
    MEM records consistent with lengths, and determinism across calls."""
    import build_index
    import synth
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "bin", "rlebwt_ms_build")):
        pytest.skip("reference builder binaries not present")
    recs = synth.pangenome(2_000_000, 4, 5, 2e-3, 2e-4)
    fa = str(tmp_path / "big.fa")
    synth.write_fasta(fa, recs)
    build_index.build_index(fa)
    text = synth.text_of(recs)
    reads = synth.short_reads(text, 200_000, 150, 6)
    flat, offs = m.pack_reads(reads)
    ix = m.MoniIndex(fa)
    ptr, ln = ix.matching_statistics(flat, offs)
    ptr2, ln2 = ix.matching_statistics(flat, offs)
    assert np.array_equal(ptr, ptr2) and np.array_equal(ln, ln2)
    L = ln.reshape(-1, 150).astype(np.int64)
    P = ptr.reshape(-1, 150)
    assert (L[:, 1:] >= L[:, :-1] - 1).all()
    assert (L >= 0).all() and (L <= 150 - np.arange(150)[None, :]).all()
    # the match itself, vectorised: compare first and last matched character and a random interior one
    idx = np.nonzero(L > 0)
    p = P[idx].astype(np.int64)
    l = L[idx]
    r, i = idx
    assert (p + l <= text.shape[0]).all()
    assert (text[p] == reads[r, i]).all()
    assert (text[p + l - 1] == reads[r, i + l - 1]).all()
    k = (np.random.default_rng(1).random(l.shape[0]) * l).astype(np.int64)
    assert (text[p + k] == reads[r, i + k]).all()
    # right-maximal at the pointer: next char differs or the read / text ends
    ext = (i + l < 150) & (p + l < text.shape[0])
    assert (text[(p + l)[ext]] != reads[r[ext], (i + l)[ext]]).all()
    # full check of 300 reads against the oracle (and so against the reference's own loop)
    orc = oracle_mod.OracleIndex(fa)
    wp, wl = orc.ms(flat[:300 * 150], offs[:301])
    assert np.array_equal(ptr[:300 * 150], wp) and np.array_equal(ln[:300 * 150], wl)
    mo, mp, ml = ix.maxrimal_exact_matches(flat, offs)
    cnt = np.diff(mo.astype(np.int64))
    is_mem = np.ones_like(L, dtype=bool)
    is_mem[:, 1:] = L[:, 1:] >= L[:, :-1]
    assert np.array_equal(cnt, is_mem.sum(axis=1))
    assert np.array_equal(ml.astype(np.int64), L[is_mem])
    assert np.array_equal(mp.astype(np.int64), np.nonzero(is_mem)[1])


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("threads", [1, 3])
def test_drop_in_cli_writes_identical_files(case, threads, tmp_path):
    """moni_align_b200/bin/{ms,mems}: the reference's command line, file names and output bytes."""
    import subprocess
    pre = os.path.join(GOLDEN, case)
    out = str(tmp_path / "o")
    for exe in ("ms", "mems"):
        r = subprocess.run([os.path.join(ROOT, "moni_align_b200", "bin", exe), pre + ".fa", "-p", pre + ".fq", "-o", out,
                            "-t", str(threads)], capture_output=True)
        assert r.returncode == 0, r.stderr
    for ext in (".pointers", ".lengths", ".mems"):
        assert open(out + ext, "rb").read() == open(pre + ext, "rb").read(), ext


@pytest.mark.parametrize("case", CASES)
def test_cli_q_reads_the_shaped_grammar(case, tmp_path):
    """`-q` = the shaped grammar <infile>.slp (the reference implements it but its getopt string lacks 'q',
    src/matching_statistics.cpp:447,466,624-627): same output bytes as the plain grammar."""
    import subprocess
    pre = os.path.join(GOLDEN, case)
    out = str(tmp_path / "o")
    for exe in ("ms", "mems"):
        r = subprocess.run([os.path.join(ROOT, "moni_align_b200", "bin", exe), pre + ".fa", "-p", pre + ".fq", "-o", out, "-q"],
                           capture_output=True)
        assert r.returncode == 0, r.stderr
    for ext in (".pointers", ".lengths", ".mems"):
        assert open(out + ext, "rb").read() == open(pre + ext, "rb").read(), ext


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("leaf_len", [0, 64])
def test_shaped_grammar_index_gives_identical_results(case, leaf_len, m, tmp_path):
    """An index directory that holds only <prefix>.slp is loaded through read_shaped_slp (text mode, and grammar mode
    with 64-byte literal blocks so the kernels walk the node table built from it); option grammar=shaped does the same when
    both grammars are present.  ms and mems bytes equal the reference's files."""
    import shutil
    pre = os.path.join(GOLDEN, case)
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs)
    only = str(tmp_path / (case + ".fa"))
    shutil.copy(pre + ".fa.thrbv.ms", only + ".thrbv.ms")
    shutil.copy(pre + ".fa.slp", only + ".slp")
    m.set_option("leaf_len", leaf_len)
    try:
        for prefix, opt in ((only, "plain"), (pre + ".fa", "shaped")):
            m.set_option("grammar", opt)
            ix = m.MoniIndex(prefix)
            ptr, ln = ix.matching_statistics(flat, offs)
            assert m.format_ms(names, offs, ptr) == open(pre + ".pointers", "rb").read()
            assert m.format_ms(names, offs, ln) == open(pre + ".lengths", "rb").read()
            mo, mp, ml = ix.maxrimal_exact_matches(flat, offs)
            assert m.format_mems(names, mo, mp, ml) == open(pre + ".mems", "rb").read()
            ix.close()
    finally:
        m.set_option("leaf_len", 0)
        m.set_option("grammar", "plain")
    with pytest.raises(m.MoniGpuError):
        m.set_option("grammar", "folded")


def test_cli_reads_gzip_and_fasta(tmp_path):
    import gzip
    import subprocess
    pre = os.path.join(GOLDEN, "pan4")
    names, seqs = read_fastq(pre + ".fq")
    fa = tmp_path / "reads.fa.gz"
    with gzip.open(fa, "wb") as f:
        for n, s in zip(names, seqs):
            body = b"\n".join(s[i:i + 60] for i in range(0, len(s), 60))
            f.write(b">" + n + b" some comment\n" + body + b"\n")
    out = str(tmp_path / "o")
    r = subprocess.run([os.path.join(ROOT, "moni_align_b200", "bin", "mems"), pre + ".fa", "-p", str(fa), "-o", out], capture_output=True)
    assert r.returncode == 0, r.stderr
    assert open(out + ".mems", "rb").read() == open(pre + ".mems", "rb").read()


def test_cli_wide_results_for_reads_of_65536_bases_or_more(m, indexes, tmp_path):
    """A batch with a read too long for the packed layouts goes through moni_gpu_ms / moni_gpu_mems inside the CLI; an empty
    patterns file gives empty outputs.  Files equal the formatted results of the library calls (which the other tests pin)."""
    import subprocess
    pre = os.path.join(GOLDEN, "pan4")
    names, seqs = read_fastq(pre + ".fq")
    text = b"".join(bytes(s) for s in seqs)
    long_read = (text * (70_000 // len(text) + 1))[:70_000]
    names2, seqs2 = [b"short0", b"long", b"short1"], [bytes(seqs[0]), long_read, bytes(seqs[1])]
    fq = tmp_path / "wide.fq"
    with open(fq, "wb") as f:
        for n, s in zip(names2, seqs2):
            f.write(b"@" + n + b"\n" + s + b"\n+\n" + b"I" * len(s) + b"\n")
    flat, offs = pack(seqs2)
    ptr, ln = indexes["pan4"].matching_statistics(flat, offs)
    mo, mp, ml = indexes["pan4"].maxrimal_exact_matches(flat, offs)
    out = str(tmp_path / "o")
    for exe in ("ms", "mems"):
        r = subprocess.run([os.path.join(ROOT, "moni_align_b200", "bin", exe), pre + ".fa", "-p", str(fq), "-o", out, "-t", "4"], capture_output=True)
        assert r.returncode == 0, r.stderr
    assert open(out + ".pointers", "rb").read() == m.format_ms(names2, offs, ptr)
    assert open(out + ".lengths", "rb").read() == m.format_ms(names2, offs, ln)
    assert open(out + ".mems", "rb").read() == m.format_mems(names2, mo, mp, ml)
    empty = tmp_path / "empty.fq"
    empty.write_bytes(b"")
    for exe in ("ms", "mems"):
        r = subprocess.run([os.path.join(ROOT, "moni_align_b200", "bin", exe), pre + ".fa", "-p", str(empty), "-o", out + "e"], capture_output=True)
        assert r.returncode == 0, r.stderr
    for ext in (".pointers", ".lengths", ".mems"):
        assert os.path.getsize(out + "e" + ext) == 0


# ------------------------------------------------------------------------------ packed records, `moni align` front end
def test_mems_packed_equals_mems(m, indexes):
    """moni_gpu_mems_packed: the same records as (query offset << 16) | length; refuses reads of 65536 bases or more."""
    for case in CASES:
        names, seqs = read_fastq(os.path.join(GOLDEN, case + ".fq"))
        flat, offs = pack(seqs)
        a = indexes[case].maxrimal_exact_matches(flat, offs)
        mo, rec = indexes[case].maxrimal_exact_matches_packed(flat, offs)
        assert rec.dtype == np.uint32
        assert np.array_equal(a[0], mo) and np.array_equal(a[1], rec >> 16) and np.array_equal(a[2], rec & 0xFFFF)
    three = m.MoniIndex(os.path.join(GOLDEN, "pan4.fa"), gpus=[0, 0, 0])
    names, seqs = read_fastq(os.path.join(GOLDEN, "pan4.fq"))
    flat, offs = pack(seqs)
    mo1, rec1 = indexes["pan4"].maxrimal_exact_matches_packed(flat, offs)
    mo3, rec3 = three.maxrimal_exact_matches_packed(flat, offs)
    assert np.array_equal(mo1, mo3) and np.array_equal(rec1, rec3)
    long_flat, long_offs = pack([b"ACGT" * 16384, b"ACGTTG"])
    with pytest.raises(m.MoniGpuError) as e:
        indexes["pan4"].maxrimal_exact_matches_packed(long_flat, long_offs)
    assert e.value.code == -1
    three.close()


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("min_len", [0, 25])
def test_seeds_equal_reference_find_mems(case, min_len, m, indexes):
    """moni_gpu_seeds == the reference's seed_finder::find_mems (tests/golden/<case>.seeds<min_len>, produced by the
    reference header on the reference's own .thrbv.full.ms)."""
    pre = os.path.join(GOLDEN, case)
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs)
    so, ref, qpos, ln = indexes[case].find_mems(flat, offs, min_len)
    assert m.format_seeds(names, so, ref, ln, qpos) == open(pre + f".seeds{min_len}", "rb").read()


@pytest.mark.parametrize("case", CASES)
def test_align_index_file_gives_identical_results(case, m, indexes, tmp_path):
    """A prefix that only has the `moni align` index (.thrbv.full.ms): ms, mems and seeds are the reference's."""
    from conftest import full_only_prefix
    pre = os.path.join(GOLDEN, case)
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs)
    ix = m.MoniIndex(full_only_prefix(case, tmp_path))
    assert (ix.n, ix.r) == (indexes[case].n, indexes[case].r)
    ptr, ln = ix.matching_statistics(flat, offs)
    assert m.format_ms(names, offs, ptr) == open(pre + ".pointers", "rb").read()
    assert m.format_ms(names, offs, ln) == open(pre + ".lengths", "rb").read()
    mo, mp, ml = ix.maxrimal_exact_matches(flat, offs)
    assert m.format_mems(names, mo, mp, ml) == open(pre + ".mems", "rb").read()
    so, ref, qpos, sl = ix.find_mems(flat, offs, 25)
    assert m.format_seeds(names, so, ref, sl, qpos) == open(pre + ".seeds25", "rb").read()
    ix.close()


def test_seeds_random_reads_chunked_and_sharded(m, oracle_mod):
    """Seeds on N-rich random reads vs the oracle: one call, small chunks (ragged, many launches) and three replicas."""
    import synth
    pre = os.path.join(GOLDEN, "withN.fa")
    orc = oracle_mod.OracleIndex(pre)
    text = np.array([orc.char_at(i) for i in range(orc.text_len)], dtype=np.uint8)
    reads = [r for r in synth.short_reads(text, 3000, 100, 41, err=0.02)]
    reads += synth.long_reads(text, 30, 42, mean=1200, sd=500, lo=80, hi=3000)
    rng = np.random.default_rng(9)
    for r in reads[::3]:
        q = int(rng.integers(0, len(r)))
        r[q:q + int(rng.integers(1, 14))] = ord("N")
    flat, offs = m.pack_reads(reads)
    m.set_option("chunk_bases", 9001)
    m.set_option("chunk_min_reads", 1)
    try:
        small = m.MoniIndex(pre)
    finally:
        m.set_option("chunk_bases", 1 << 24)
        m.set_option("chunk_min_reads", 96 * 1024)
    one, three = m.MoniIndex(pre), m.MoniIndex(pre, gpus=[0, 0, 0])
    for min_len in (0, 12):
        want = orc.seeds(flat, offs, min_len)
        for ix in (one, small, three):
            got = ix.find_mems(flat, offs, min_len)
            assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
            assert np.array_equal(got[2], want[2]) and np.array_equal(got[3], want[3])
    for ix in (one, small, three):
        ix.close()


@pytest.mark.parametrize("case", CASES)
def test_lcp_align_index_on_the_gpu(case, m, oracle_mod, tmp_path):
    """The default `moni align` index (.thrbv.full.lcp.ms): same seeds, same occurrence lists as the reference's
    LCP-based enumeration (identical to .seeds25 / .occs12, see make_golden_seeds.py::lcp_variant)."""
    from conftest import full_only_prefix
    pre = os.path.join(GOLDEN, case)
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs)
    ix = m.MoniIndex(full_only_prefix(case, tmp_path, lcp=True))
    so, ref, qpos, ln = ix.find_mems(flat, offs, 25)
    assert m.format_seeds(names, so, ref, ln, qpos) == open(pre + ".seeds25", "rb").read()
    so, ref, qpos, ln = ix.find_mems(flat, offs, 12)
    oo, occs, trunc = ix.find_MEM_occs(ref, ln)
    assert not trunc
    assert oracle_mod.format_occs(names, so, ref, ln, qpos, oo, occs) == open(pre + ".occs12", "rb").read()
    ix.close()


@pytest.mark.parametrize("case", CASES)
def test_seed_occurrences_equal_reference_find_MEM_occs(case, m, indexes, oracle_mod, tmp_path):
    """moni_gpu_seeds + moni_gpu_seed_occs on the `moni align` index == the reference's find_mems + find_MEM_occs
    (tests/golden/<case>.occs12); per-direction cap and capacity errors behave as documented."""
    from conftest import full_only_prefix
    pre = os.path.join(GOLDEN, case)
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs)
    ix = m.MoniIndex(full_only_prefix(case, tmp_path))
    so, ref, qpos, ln = ix.find_mems(flat, offs, 12)
    oo, occs, trunc = ix.find_MEM_occs(ref, ln)
    assert not trunc
    assert oracle_mod.format_occs(names, so, ref, ln, qpos, oo, occs) == open(pre + ".occs12", "rb").read()
    # cap of one step per direction: a prefix-consistent subset, flagged
    oo1, occs1, trunc1 = ix.find_MEM_occs(ref, ln, max_per_direction=1)
    cnt, cnt1 = np.diff(oo.astype(np.int64)), np.diff(oo1.astype(np.int64))
    assert (cnt1 <= 3).all() and (cnt1 <= cnt).all() and trunc1 == bool((cnt1 < cnt).any())
    assert np.array_equal(occs1[oo1[:-1].astype(np.int64)], ref)
    with pytest.raises(m.MoniGpuError) as e:
        ix.find_MEM_occs(ref, ln, capacity=1)
    assert e.value.code == -4
    with pytest.raises(m.MoniGpuError) as e:          # an index without Phi structures refuses
        indexes[case].find_MEM_occs(ref[:1], ln[:1])
    assert e.value.code == -1
    ix.close()


def test_concurrent_callers_on_one_index(m, indexes, oracle_mod):
    """The library is thread-safe per index (one mutex per GPU replica): four host threads calling ms, mems,
    packed mems and seeds on the same index at once each get the bits of a lone call."""
    import threading
    import synth
    pre = os.path.join(GOLDEN, "pan4.fa")
    orc = oracle_mod.OracleIndex(pre)
    text = np.array([orc.char_at(i) for i in range(orc.text_len)], dtype=np.uint8)
    reads = list(synth.short_reads(text, 4000, 110, 51, err=0.02)) + synth.long_reads(text, 20, 52, mean=800, sd=300, lo=60, hi=2000)
    flat, offs = m.pack_reads(reads)
    ix = indexes["pan4"]
    want = {"ms": ix.matching_statistics(flat, offs), "mems": ix.maxrimal_exact_matches(flat, offs),
            "packed": ix.maxrimal_exact_matches_packed(flat, offs), "seeds": ix.find_mems(flat, offs, 15)}
    calls = {"ms": lambda: ix.matching_statistics(flat, offs), "mems": lambda: ix.maxrimal_exact_matches(flat, offs),
             "packed": lambda: ix.maxrimal_exact_matches_packed(flat, offs), "seeds": lambda: ix.find_mems(flat, offs, 15)}
    bad = []

    def worker(name):
        for _ in range(5):
            got = calls[name]()
            if not all(np.array_equal(a, b) for a, b in zip(got, want[name])):
                bad.append(name)

    th = [threading.Thread(target=worker, args=(k,)) for k in calls]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not bad, bad


@pytest.mark.parametrize("case", CASES)
def test_ms_packed_equals_ms(case, m, indexes):
    """moni_gpu_ms_packed (u32 + u8 pointer halves, u16 lengths) decodes to exactly the u64 results of moni_gpu_ms,
    including the reference's wrapped pointers (withN has reads with symbols absent from the text)."""
    names, seqs = read_fastq(os.path.join(GOLDEN, case + ".fq"))
    flat, offs = pack(seqs)
    ptr, ln = indexes[case].matching_statistics(flat, offs)
    lo, hi, l16 = indexes[case].matching_statistics_packed(flat, offs)
    assert lo.dtype == np.uint32 and hi.dtype == np.uint8 and l16.dtype == np.uint16
    assert np.array_equal(m.unpack_pointers(lo, hi), ptr) and np.array_equal(l16, ln)
    if case == "withN":
        assert (ptr >= np.uint64(1 << 63)).any()          # wrapped pointers are present in this fixture
    three = m.MoniIndex(os.path.join(GOLDEN, case + ".fa"), gpus=[0, 0, 0])
    lo3, hi3, l3 = three.matching_statistics_packed(flat, offs)
    assert np.array_equal(lo3, lo) and np.array_equal(hi3, hi) and np.array_equal(l3, l16)
    three.close()
    with pytest.raises(m.MoniGpuError) as e:
        indexes[case].matching_statistics_packed(*pack([b"AC" * 40000]))
    assert e.value.code == -1


def test_numa_near_pinned_buffers(m, indexes):
    """moni_gpu_host_alloc_near gives usable page-locked memory whatever the host topology."""
    a = m.pinned_empty(1000, np.uint64, gpu=0)
    a[:] = 7
    assert int(a.sum()) == 7000
    assert m.load_library().moni_gpu_numa_node(0) >= -1


def test_mems_packed16_equals_mems(m, indexes):
    """moni_gpu_mems_packed16: the same records as (query offset << 8) | length for reads shorter than 256 bases (one call, small
    chunks, three replicas); a longer read is refused."""
    for case in CASES:
        names, seqs = read_fastq(os.path.join(GOLDEN, case + ".fq"))
        short = [s for s in seqs if len(s) < 256]
        assert len(short) > 20
        flat, offs = pack(short)
        a = indexes[case].maxrimal_exact_matches(flat, offs)
        mo, rec = indexes[case].maxrimal_exact_matches_packed16(flat, offs)
        assert rec.dtype == np.uint16
        assert np.array_equal(a[0], mo) and np.array_equal(a[1], rec >> 8) and np.array_equal(a[2], rec & 0xFF)
    names, seqs = read_fastq(os.path.join(GOLDEN, "pan4.fq"))
    short = [s for s in seqs if len(s) < 256]
    flat, offs = pack(short)
    three = m.MoniIndex(os.path.join(GOLDEN, "pan4.fa"), gpus=[0, 0, 0])
    a = three.maxrimal_exact_matches(flat, offs)
    m.set_option("chunk_bases", 1000)
    m.set_option("chunk_min_reads", 0)
    try:
        small = m.MoniIndex(os.path.join(GOLDEN, "pan4.fa"))
    finally:
        m.set_option("chunk_bases", 1 << 24)
        m.set_option("chunk_min_reads", 98304)
    for ix in (three, small):
        mo, rec = ix.maxrimal_exact_matches_packed16(flat, offs)
        assert np.array_equal(a[0], mo) and np.array_equal(a[1], rec >> 8) and np.array_equal(a[2], rec & 0xFF)
    flat, offs = pack(short + [b"A" * 256])
    with pytest.raises(m.MoniGpuError) as e:
        small.maxrimal_exact_matches_packed16(flat, offs)
    assert e.value.code == -1
    three.close()
    small.close()


def _runs_of(ptr, ln, offs):
    """Restart runs of reference-style vectors, in numpy: restart at i == 0 or ptr[i] != ptr[i-1] + 1 (u64 arithmetic)."""
    total = int(offs[-1])
    rst = np.ones(total, dtype=bool)
    with np.errstate(over="ignore"):
        rst[1:] = ptr[1:] != ptr[:-1] + np.uint64(1)
    rst[offs[:-1][offs[:-1] < total].astype(np.int64)] = True
    at = np.nonzero(rst)[0]
    read_of = np.searchsorted(offs, at, side="right") - 1
    qpos = (at - offs[read_of].astype(np.int64)).astype(np.uint32)
    ro = np.zeros(offs.shape[0], dtype=np.uint64)
    np.cumsum(np.bincount(read_of, minlength=offs.shape[0] - 1), out=ro[1:])
    return ro, qpos, ln[at].astype(np.uint32), ptr[at]


@pytest.mark.parametrize("case", CASES)
def test_ms_runs_expand_to_ms(case, m, indexes):
    """moni_gpu_ms_runs returns exactly the restarts of the reference's vectors, and moni_gpu_ms_runs_expand rebuilds
    pointers[] / lengths[] bit for bit (withN: wrapped pointers 2^64 - k); packed runs decode to the same; three replicas
    keep the input order."""
    names, seqs = read_fastq(os.path.join(GOLDEN, case + ".fq"))
    seqs = seqs + [b"", b"A", b"ACGT" * 70]
    flat, offs = pack(seqs)
    ix = indexes[case]
    ptr, ln = ix.matching_statistics(flat, offs)
    ro, qp, rl, rp = ix.matching_statistics_runs(flat, offs)
    want = _runs_of(ptr, ln, offs)
    for got, exp in zip((ro, qp, rl, rp), want):
        assert np.array_equal(got, exp)
    p2, l2 = m.expand_runs(offs, ro, qp, rl, rp)
    assert np.array_equal(p2, ptr) and np.array_equal(l2, ln)
    assert m.format_ms(names, offs, p2) == open(os.path.join(GOLDEN, case + ".pointers"), "rb").read()
    assert m.format_ms(names, offs, l2) == open(os.path.join(GOLDEN, case + ".lengths"), "rb").read()
    rop, rec, lo, hi = ix.matching_statistics_runs(flat, offs, packed=True)
    q2, n2, r2 = m.unpack_runs(rec, lo, hi)
    assert np.array_equal(rop, ro) and np.array_equal(q2, qp) and np.array_equal(n2, rl) and np.array_equal(r2, rp)
    three = m.MoniIndex(os.path.join(GOLDEN, case + ".fa"), gpus=[0, 0, 0])
    for got, exp in zip(three.matching_statistics_runs(flat, offs), (ro, qp, rl, rp)):
        assert np.array_equal(got, exp)
    three.close()
    with pytest.raises(m.MoniGpuError) as e:
        ix.matching_statistics_runs(*pack([b"AC" * 40000]), packed=True)
    assert e.value.code == -1


def test_ms_runs_on_random_chunked_batches(m, oracle_mod):
    """Runs over many chunks and lanes (small chunk size) expand to the oracle's vectors; ragged lengths incl. empty reads."""
    rng = np.random.default_rng(77)
    pre = os.path.join(GOLDEN, "pan4.fa")
    orc = oracle_mod.OracleIndex(pre)
    text = np.array([orc.char_at(i) for i in range(orc.text_len)], dtype=np.uint8)
    reads = []
    for _ in range(3000):
        L = int(rng.integers(0, 400))
        s = int(rng.integers(0, max(1, text.shape[0] - L - 1)))
        r = text[s:s + L].copy()
        mut = rng.random(L) < 0.03
        r[mut] = np.frombuffer(b"ACGTN", dtype=np.uint8)[rng.integers(0, 5, size=int(mut.sum()))]
        reads.append(r.tobytes())
    flat, offs = pack(reads)
    m.set_option("chunk_bases", 20000)
    m.set_option("chunk_min_reads", 64)
    try:
        ix = m.MoniIndex(pre)
        ro, qp, rl, rp = ix.matching_statistics_runs(flat, offs)
        ptr, ln = m.expand_runs(offs, ro, qp, rl, rp)
        ix.close()
    finally:
        m.set_option("chunk_bases", 1 << 24)
        m.set_option("chunk_min_reads", 96 * 1024)
    wp, wl = orc.ms(flat, offs)
    assert np.array_equal(ptr, wp) and np.array_equal(ln, wl)


@pytest.mark.parametrize("case", ["pan4", "withN"])
def test_dynamic_and_static_kernels_agree(case, m, oracle_mod):
    """The persistent kernels whose lanes pull reads from a counter (option dynamic=1; the default for ragged chunks) and the
    one-read-per-lane kernels (dynamic=0) give the same bits for ms, mems, seeds and runs, on ragged batches with empty reads,
    one-base reads and reads much longer than the others, over several chunks."""
    import synth
    pre = os.path.join(GOLDEN, case + ".fa")
    orc = oracle_mod.OracleIndex(pre)
    text = np.array([orc.char_at(i) for i in range(orc.text_len)], dtype=np.uint8)
    reads = [r for r in synth.short_reads(text, 2000, 100, 31, err=0.03)]
    reads += synth.long_reads(text, 60, 32, mean=1500, sd=900, lo=1, hi=6000)
    reads += [np.zeros(0, dtype=np.uint8), np.frombuffer(b"A", dtype=np.uint8), np.frombuffer(b"NNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNN", dtype=np.uint8)]
    rng = np.random.default_rng(9)
    rng.shuffle(reads)
    flat, offs = m.pack_reads(reads)
    res = {}
    for dyn in (0, 1, 11, 17):           # 11, 17: persistent kernels sized for 1 resp. 7 reads per thread (dyn_reads_per_lane; default 2)
        m.set_option("dynamic", min(dyn, 1))
        m.set_option("dyn_reads_per_lane", {11: 1, 17: 7}.get(dyn, 2))
        m.set_option("chunk_bases", 60000)
        m.set_option("chunk_min_reads", 200)
        try:
            ix = m.MoniIndex(pre)
        finally:
            m.set_option("dynamic", 2)
            m.set_option("chunk_bases", 1 << 24)
            m.set_option("chunk_min_reads", 96 * 1024)
        res[dyn] = (ix.matching_statistics(flat, offs), ix.maxrimal_exact_matches(flat, offs), ix.find_mems(flat, offs, 12),
                    ix.matching_statistics_runs(flat, offs))
        ix.close()
    m.set_option("dyn_reads_per_lane", 2)
    for other in (1, 11, 17):
        for a, b in zip(res[0], res[other]):
            assert all(np.array_equal(x, y) for x, y in zip(a, b)), other
    wp, wl = orc.ms(flat, offs)
    assert np.array_equal(res[1][0][0], wp) and np.array_equal(res[1][0][1], wl)
    wm = orc.mems(flat, offs)
    assert all(np.array_equal(x, y) for x, y in zip(res[1][1], wm))


def test_chunk_schedules_keep_the_bits(m, oracle_mod):
    """The chunk-size schedules (options "ramp": shorter first chunks, "guided": chunks that shrink towards the end of the batch), the
    number of chunks in flight ("lanes") and the sizing of the persistent kernels change how a batch is cut and run, never what comes
    back: ragged reads over many small chunks, every combination against the oracle."""
    import synth
    pre = os.path.join(GOLDEN, "pan4.fa")
    orc = oracle_mod.OracleIndex(pre)
    text = np.array([orc.char_at(i) for i in range(orc.text_len)], dtype=np.uint8)
    reads = synth.long_reads(text, 400, 41, mean=900, sd=500, lo=1, hi=3000)
    reads += [r for r in synth.short_reads(text, 1500, 80, 42, err=0.04)]
    reads.insert(100, np.zeros(0, np.uint8))
    flat, offs = m.pack_reads(reads)
    want_ms, want_mems = orc.ms(flat, offs), orc.mems(flat, offs)
    try:
        for ramp, guided, lanes, rpt in ((1, 0, 0, 2), (0, 1, 3, 1), (1, 1, 8, 4), (2, 1, 1, 2)):
            for k, v in (("ramp", ramp), ("guided", guided), ("lanes", lanes), ("dyn_reads_per_lane", rpt), ("chunk_bases", 20_000), ("chunk_min_reads", 64)):
                m.set_option(k, v)
            ix = m.MoniIndex(pre)
            got = ix.matching_statistics(flat, offs)
            assert np.array_equal(got[0], want_ms[0]) and np.array_equal(got[1], want_ms[1]), (ramp, guided, lanes, rpt)
            gm = ix.maxrimal_exact_matches(flat, offs)
            assert all(np.array_equal(x, y) for x, y in zip(gm, want_mems)), (ramp, guided, lanes, rpt)
            ix.close()
    finally:
        for k, v in (("ramp", 0), ("guided", 0), ("lanes", 0), ("dyn_reads_per_lane", 2), ("chunk_bases", 1 << 24), ("chunk_min_reads", 96 * 1024)):
            m.set_option(k, v)


@pytest.mark.parametrize("case", CASES)
def test_populate_seeds_equals_reference(case, m, oracle_mod, tmp_path):
    """moni_gpu_seeds + moni_gpu_populate_seeds on the `moni align` index == the reference's find_mems + populate_seeds with seed
    filtering off (tests/golden/<case>.pop12, min_len 12): every MEM's list, then the two halves of every MEM of >= 24 bases."""
    from conftest import full_only_prefix
    pre = os.path.join(GOLDEN, case)
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs)
    ix = m.MoniIndex(full_only_prefix(case, tmp_path))
    so, ref, qpos, ln = ix.find_mems(flat, offs, 12)
    po, pref, pln, pq, oo, occs, trunc = ix.populate_seeds(so, ref, ln, qpos, 12)
    assert not trunc
    assert int(po[-1]) > ref.shape[0]
    assert oracle_mod.format_occs(names, po, pref, pln, pq, oo, occs) == open(pre + ".pop12", "rb").read()
    with pytest.raises(m.MoniGpuError) as e:          # a split needs min_len >= 1
        ix.populate_seeds(so, ref, ln, qpos, 0)
    assert e.value.code == -1
    ix.close()


def test_mems_u64_widened_on_host_equals_link_copy(m, oracle_mod):
    """moni_gpu_mems with its records crossing the link packed and widened by host threads (option widen_threads, the default where
    the process may use 4 or more CPUs) against the same call with the u64 columns copied as they are (widen_threads = 0) and
    against the oracle: chunks of reads shorter than 256 bases (2-byte records), shorter than 65536 (4-byte) and longer (two u32
    columns) in one batch, over several lanes, on one GPU and sharded over three replicas."""
    import synth
    pre = os.path.join(GOLDEN, "pan4.fa")
    orc = oracle_mod.OracleIndex(pre)
    text = np.array([orc.char_at(i) for i in range(orc.text_len)], dtype=np.uint8)
    reads = [r for r in synth.short_reads(text, 6000, 100, 21, err=0.03)]
    reads += synth.long_reads(text, 60, 22, mean=2000, sd=800, lo=300, hi=5000)
    reads += [r for r in synth.short_reads(text, 3000, 250, 23, err=0.02)]
    reads.append(np.tile(text[:20000], 4)[:70000])          # a read of 65536 bases or more: u32 columns
    reads += [r for r in synth.short_reads(text, 2000, 60, 24, err=0.05)]
    reads.insert(5, np.zeros(0, np.uint8))
    flat, offs = m.pack_reads(reads)
    m.set_option("chunk_bases", 50_000)
    m.set_option("chunk_min_reads", 1)
    try:
        one = m.MoniIndex(pre)
        three = m.MoniIndex(pre, gpus=[0, 0, 0])
    finally:
        m.set_option("chunk_bases", 1 << 24)
        m.set_option("chunk_min_reads", 96 * 1024)
    want = orc.mems(flat, offs)
    want_ms = orc.ms(flat, offs)
    try:
        for threads in (0, 3, 7, 1, -1):      # moni_gpu_mems widens from 2 threads, moni_gpu_ms from 6
            m.set_option("widen_threads", threads)
            for ix in (one, three):
                got = ix.maxrimal_exact_matches(flat, offs)
                assert all(np.array_equal(x, y) for x, y in zip(got, want)), threads
                # moni_gpu_ms: 7 bytes per base cross the link and are widened (chunks with a read of 65536 bases or more go as u64)
                gp, gl = ix.matching_statistics(flat, offs)
                assert np.array_equal(gp, want_ms[0]) and np.array_equal(gl, want_ms[1]), threads
        # wrapped pointers (2^64 - k after absent symbols) survive the 40-bit form: the withN reads, widened and not
        names, seqs = read_fastq(os.path.join(GOLDEN, "withN.fq"))
        wflat, woffs = pack(seqs)
        wn = m.MoniIndex(os.path.join(GOLDEN, "withN.fa"))
        for threads in (0, 6):
            m.set_option("widen_threads", threads)
            p2, l2 = wn.matching_statistics(wflat, woffs)
            assert m.format_ms(names, woffs, p2) == open(os.path.join(GOLDEN, "withN.pointers"), "rb").read()
            assert m.format_ms(names, woffs, l2) == open(os.path.join(GOLDEN, "withN.lengths"), "rb").read()
        wn.close()
        # capacity error path with widening on: offsets still say how much is needed
        m.set_option("widen_threads", 2)
        mo = np.zeros(len(reads) + 1, np.uint64)
        small = np.zeros(16, np.uint64)
        with pytest.raises(m.MoniGpuError) as e:
            one.maxrimal_exact_matches(flat, offs, out=(mo, small, small.copy()))
        assert e.value.code == -4 and int(mo[-1]) == int(want[0][-1])
    finally:
        m.set_option("widen_threads", -1)
    one.close()
    three.close()


def test_index_larger_than_l2_against_reference_binaries(m, tmp_path):
    """BASELINE configs[2] (cfg3: n = 1.28 G, 4.8 GB image, far larger than the 126 MB L2) when its index travels with the snapshot
    (data/cfg3, what bench.py times): 20 000 reads sampled from the text through the CUDA path and through the unmodified reference
    `ms` / `mems` binaries -- pointers, lengths and MEM records equal number by number.  Skipped where the index is absent."""
    import build_index
    import refparse
    import synth
    import workloads
    pre = workloads.index_prefix("cfg3")
    if not os.path.exists(pre + ".thrbv.ms"):
        pytest.skip("data/cfg3 index not present")
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "bin", "mems")):
        pytest.skip("reference binaries not built")
    workloads.ensure_index("cfg3", verbose=False)          # makes <prefix>.plain.slp from the seeded text if only the .thrbv.ms travelled
    reads = workloads.reads_of("cfg3", pairs=10_000)
    flat, offs = m.pack_reads(reads)
    ix = m.MoniIndex(pre)
    assert ix.n > (1 << 30)
    ptr, ln = ix.matching_statistics(flat, offs)
    mo, mp, ml = ix.maxrimal_exact_matches(flat, offs)
    ro, rq, rl, rp = ix.matching_statistics_runs(flat, offs)
    xp, xl = m.expand_runs(offs, ro, rq, rl, rp)
    assert np.array_equal(xp, ptr) and np.array_equal(xl, ln)
    ix.close()
    fq = os.path.join(str(tmp_path), "s.fq")
    synth.write_fastq_matrix(fq, reads)
    build_index.run_reference("ms", pre, fq, os.path.join(str(tmp_path), "ref"), threads=os.cpu_count() or 1)
    build_index.run_reference("mems", pre, fq, os.path.join(str(tmp_path), "ref"), threads=os.cpu_count() or 1)
    _, _, pv = refparse.parse_values(os.path.join(str(tmp_path), "ref.pointers"))
    _, _, lv = refparse.parse_values(os.path.join(str(tmp_path), "ref.lengths"))
    assert np.array_equal(pv, ptr) and np.array_equal(lv, ln)
    _, cnt, qo, ql = refparse.parse_mems(os.path.join(str(tmp_path), "ref.mems"))
    assert np.array_equal(cnt.astype(np.int64), np.diff(mo.astype(np.int64)))
    assert np.array_equal(qo, mp) and np.array_equal(ql, ml)


def test_index_larger_than_l2_built_on_the_box(m, tmp_path):
    """An index whose image (move table + text, > 300 MB) does not fit the 126 MB L2, built on the spot with the unmodified reference
    builder (8 x 20 Mbp, about a minute): 30 000 paired reads through the CUDA path (all result forms) and through the unmodified
    reference `ms` / `mems` binaries -- pointers, lengths and MEM records equal number by number.  Unlike the cfg3 test above it
    needs nothing that does not travel with the snapshot, so it runs wherever the GPU suite runs."""
    import build_index
    import refparse
    import synth
    import workloads
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "bin", "rlebwt_ms_build")):
        pytest.skip("reference builder binaries not present")
    recs = synth.pangenome(20_000_000, 8, 55, 1e-3, 1e-4)
    fa = str(tmp_path / "l2.fa")
    synth.write_fasta(fa, recs)
    build_index.build_index(fa)
    text = synth.text_of(recs)
    del recs
    reads = workloads.paired_reads(text, 15_000, 150, 56)
    flat, offs = m.pack_reads(reads)
    ix = m.MoniIndex(fa)
    assert ix.device_bytes > 300_000_000 and ix.n > 150_000_000
    ptr, ln = ix.matching_statistics(flat, offs)
    mo, mp, ml = ix.maxrimal_exact_matches(flat, offs)
    ro, rq, rl, rp = ix.matching_statistics_runs(flat, offs)
    xp, xl = m.expand_runs(offs, ro, rq, rl, rp)
    assert np.array_equal(xp, ptr) and np.array_equal(xl, ln)
    po, prec = ix.maxrimal_exact_matches_packed16(flat, offs)
    assert np.array_equal(po, mo) and np.array_equal(prec >> 8, mp) and np.array_equal(prec & 0xFF, ml)
    ix.close()
    fq = os.path.join(str(tmp_path), "s.fq")
    synth.write_fastq_matrix(fq, reads)
    build_index.run_reference("ms", fa, fq, os.path.join(str(tmp_path), "ref"), threads=os.cpu_count() or 1)
    build_index.run_reference("mems", fa, fq, os.path.join(str(tmp_path), "ref"), threads=os.cpu_count() or 1)
    _, _, pv = refparse.parse_values(os.path.join(str(tmp_path), "ref.pointers"))
    _, _, lv = refparse.parse_values(os.path.join(str(tmp_path), "ref.lengths"))
    assert np.array_equal(pv, ptr) and np.array_equal(lv, ln)
    _, cnt, qo, ql = refparse.parse_mems(os.path.join(str(tmp_path), "ref.mems"))
    assert np.array_equal(cnt.astype(np.int64), np.diff(mo.astype(np.int64)))
    assert np.array_equal(qo, mp) and np.array_equal(ql, ml)


def test_bounds_checked_build_runs_the_suite():
    """The kernels compiled with -DMONI_BOUNDS_CHECK (an index check in front of every data-dependent global access: move rows,
    read bytes incl. the 15 bytes load8 may touch past a read, text words, flag words, restart lists, records; a violation traps)
    run the parity tests that exercise ragged batches, chunking, sharding, both kernel families and the grammar path without a
    trap and with the same bits.  This is the memcheck of this tree: the pool's compute-sanitizer is closed."""
    import subprocess
    import sys
    lib = os.path.join(ROOT, "moni_align_b200", "variants", "libmoni_gpu_check.so")
    if not os.path.exists(lib):
        pytest.skip("bounds-checked library not built (make -C moni_align_b200/csrc)")
    env = dict(os.environ, MONI_GPU_LIB=lib)
    k = ("reference_files or random_reads or empty_ragged or chunking or sharded or runs or dynamic or packed or seeds_equal or "
         "grammar_index or occurrences or populate")
    p = subprocess.run([sys.executable, "-m", "pytest", os.path.join(ROOT, "tests", "test_gpu_parity.py"), "-m", "gpu", "-x", "-q", "-k", k,
                        "--deselect", "tests/test_gpu_parity.py::test_bounds_checked_build_runs_the_suite"],
                       env=env, capture_output=True, text=True, cwd=ROOT)
    tail = (p.stdout + p.stderr)[-3000:]
    assert p.returncode == 0, tail
    assert "bounds check failed" not in tail
    assert " passed" in p.stdout
