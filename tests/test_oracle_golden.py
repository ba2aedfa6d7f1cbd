"""CPU suite, part 1: the oracle (oracle/oracle.py + oracle/ms_oracle.c) is pinned against outputs of the
UNMODIFIED reference binaries committed under tests/golden/ (made by tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest

from conftest import CASES, GOLDEN, pack, read_fastq


def _u40(path, k):
    a = np.fromfile(path, dtype=np.uint8).reshape(-1, 5 * k)
    out = []
    for j in range(k):
        v = np.zeros(a.shape[0], dtype=np.uint64)
        for b in range(5):
            v |= a[:, 5 * j + b].astype(np.uint64) << np.uint64(8 * b)
        out.append(v)
    return out


@pytest.mark.parametrize("case", CASES)
def test_parser_matches_builder_intermediates(case, oracle_mod):
    """.thrbv.ms decoded without sdsl == the builder's raw run files (reference test/src/thr_test.cpp:166-183
    checks the same equivalence between thr_bv and the raw .thr_pos)."""
    pre = os.path.join(GOLDEN, case + ".fa")
    d = oracle_mod.read_thrbv_ms(pre + ".thrbv.ms")
    heads = np.fromfile(pre + ".bwt.heads", dtype=np.uint8)
    heads[heads <= 1] = 1
    assert np.array_equal(heads, d["heads"])
    (lens,) = _u40(pre + ".bwt.len", 1)
    assert np.array_equal(lens, d["lens"])
    n = d["n"]
    (thr,) = _u40(pre + ".thr_pos", 1)
    for c in range(256):
        mine = d["thr"][c]
        theirs = thr[(heads == c) & (thr > 0)]
        assert np.array_equal(mine, theirs), c
    for ext, key in ((".ssa", "samples_start"), (".esa", "samples_last")):
        _, sa = _u40(pre + ext, 2)
        want = np.where(sa > 0, sa - np.uint64(1), np.uint64(n - 1))
        assert np.array_equal(want, d[key])


@pytest.mark.parametrize("case", CASES)
def test_slp_expands_to_text_length(case, oracle_mod):
    s = oracle_mod.read_plain_slp(os.path.join(GOLDEN, case + ".fa.plain.slp"))
    d = oracle_mod.read_thrbv_ms(os.path.join(GOLDEN, case + ".fa.thrbv.ms"))
    assert s["text_len"] == d["n"] - 1
    assert s["explen"][s["start_var"] - s["alph_size"]] == s["text_len"]


@pytest.mark.parametrize("case", CASES)
def test_oracle_equals_reference_outputs(case, oracle_mod):
    pre = os.path.join(GOLDEN, case)
    ix = oracle_mod.OracleIndex(pre + ".fa")
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs)
    ptr, ln = ix.ms(flat, offs)
    assert oracle_mod.format_ms(names, offs, ptr) == open(pre + ".pointers", "rb").read()
    assert oracle_mod.format_ms(names, offs, ln) == open(pre + ".lengths", "rb").read()
    mo, mp, ml = ix.mems(flat, offs)
    assert oracle_mod.format_mems(names, mo, mp, ml) == open(pre + ".mems", "rb").read()


def test_oracle_lengths_are_true_matching_statistics(oracle_mod):
    """Independent check (brute force on the text): T[ptr:ptr+len] == read[i:i+len] and the match is
    right-maximal; positions where the pointer was produced by an absent symbol are skipped."""
    pre = os.path.join(GOLDEN, "pan4")
    ix = oracle_mod.OracleIndex(pre + ".fa")
    text = bytes(ix.char_at(i) for i in range(ix.text_len))
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs[:40])
    ptr, ln = ix.ms(flat, offs)
    for r in range(40):
        s = seqs[r]
        for i in range(len(s)):
            p, l = int(ptr[int(offs[r]) + i]), int(ln[int(offs[r]) + i])
            assert text[p:p + l] == s[i:i + l]
            assert s[i:i + l + 1] not in text or i + l == len(s)


# ---------------------------------------------------------------------------- `moni align` front end
@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("min_len", [0, 25])
def test_oracle_seeds_equal_reference_find_mems(case, min_len, oracle_mod):
    """oracle_find_mems == the reference's seed_finder::find_mems (run through oracle/ref_seeds_harness.cpp on the
    reference's own .thrbv.full.ms; committed as tests/golden/<case>.seeds<min_len>)."""
    pre = os.path.join(GOLDEN, case)
    ix = oracle_mod.OracleIndex(pre + ".fa")
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs)
    so, ref, qpos, ln = ix.seeds(flat, offs, min_len)
    assert oracle_mod.format_seeds(names, so, ref, ln, qpos) == open(pre + f".seeds{min_len}", "rb").read()


@pytest.mark.parametrize("case", CASES)
def test_full_index_file_holds_the_same_query_structures(case, oracle_mod):
    """.thrbv.full.ms (rlebwt_ms_full_build -n) == .thrbv.ms (rlebwt_ms_build) on everything the MS query reads."""
    pre = os.path.join(GOLDEN, case + ".fa")
    a = oracle_mod.read_thrbv_ms(pre + ".thrbv.ms")
    b = oracle_mod.read_thrbv_ms(pre + ".thrbv.full.ms", full=True)
    for key in ("n", "R", "terminator_position"):
        assert a[key] == b[key]
    for key in ("F", "heads", "starts", "samples_start", "samples_last", "count"):
        assert np.array_equal(a[key], b[key]), key
    for c in range(256):
        assert np.array_equal(a["thr"][c], b["thr"][c])
    assert b["pred"].shape[0] == b["R"] and b["pred_start"].shape[0] == b["R"]


def test_seed_rule_drops_all_N_matches(oracle_mod):
    """The N filter of seed_finder.hpp:148 is exercised by the fixtures: withN has reads whose only matches are N runs."""
    pre = os.path.join(GOLDEN, "withN")
    ix = oracle_mod.OracleIndex(pre + ".fa")
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs)
    so, ref, qpos, ln = ix.seeds(flat, offs, 0)
    mo, mp, ml = ix.mems(flat, offs)
    r = seqs.index(b"NNNNNNNNNNNNNNNNNNNN")
    assert int(mo[r + 1] - mo[r]) > 0 and int(ml[int(mo[r])]) > 0      # `mems` reports the N match ...
    assert int(so[r + 1] - so[r]) == 0                                  # ... the seed rule drops it


@pytest.mark.parametrize("case", CASES)
def test_oracle_occurrences_equal_reference_find_MEM_occs(case, oracle_mod, tmp_path):
    """oracle_find_mem_occs (Phi / Phi^-1 walks with the bounded LCE) == the reference's seed_finder::find_MEM_occs with seed
    filtering off, on the reference's own .thrbv.full.ms (tests/golden/<case>.occs12, min_len 12)."""
    from conftest import full_only_prefix
    pre = os.path.join(GOLDEN, case)
    ix = oracle_mod.OracleIndex(full_only_prefix(case, tmp_path))
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs)
    so, ref, qpos, ln = ix.seeds(flat, offs, 12)
    oo, occs = ix.mem_occs(ref, ln)
    assert int(oo[-1]) > len(ref)            # some seeds do have further occurrences
    assert oracle_mod.format_occs(names, so, ref, ln, qpos, oo, occs) == open(pre + ".occs12", "rb").read()


@pytest.mark.parametrize("case", CASES)
def test_oracle_populate_seeds_equals_reference(case, oracle_mod, tmp_path):
    """The restatement of seed_finder::populate_seeds (half-MEM splitting, seed filtering off) == the reference's own, on its own
    .thrbv.full.ms (tests/golden/<case>.pop12: ref_seeds ... pop, min_len 12): the read's MEMs with their lists, then the halves."""
    from conftest import full_only_prefix
    pre = os.path.join(GOLDEN, case)
    ix = oracle_mod.OracleIndex(full_only_prefix(case, tmp_path))
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs)
    so, ref, qpos, ln = ix.seeds(flat, offs, 12)
    po, pref, pln, pq, oo, occs = ix.populate_seeds(so, ref, ln, qpos, 12)
    assert int(po[-1]) > len(ref)            # some MEMs are long enough to be split
    assert oracle_mod.format_occs(names, po, pref, pln, pq, oo, occs) == open(pre + ".pop12", "rb").read()
