"""The BASELINE.json workloads as seeded synthetic stand-ins.  Input preparation only.

cfg2 -- "mouse chr19 FASTA+VCF subset (-H12) index, `moni mems` on 1M synthetic 150 bp paired-end reads".
The mouse FASTA/VCF are not in the reference checkout (.MISSING_LARGE_BLOBS), so the text is a 61.4 Mbp uniform
ACGT reference plus 6 haplotypes (3 samples x 2) with SNPs at 5e-3 and indels at 5e-4: 7 records, ~430 Mbp,
r ~ 68.7 M runs.  The index is built ONCE with the unmodified reference builder (tools/build_index.py, ~13 min of
pfp_thresholds) and cached under data/cfg2/; only `ref.fa.thrbv.ms` and `ref.fa.plain.slp` are kept, the text is
regenerated from its seed in seconds whenever reads have to be sampled.

small -- a 4 x 2 Mbp pangenome for quick functional runs of bench.py (not a bench line).
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import build_index  # noqa: E402
import synth  # noqa: E402

WORKLOADS = {
    # BASELINE.json configs[0]: "data/yeast.fasta index, `moni ms` on 100k synthetic 150 bp reads (1% errors)".  data/yeast.fasta is
    # not in the reference checkout (.MISSING_LARGE_BLOBS): yeast-scale stand-in of SURVEY.md 8(d), 12.16 Mbp of uniform ACGT.
    "cfg1": dict(base_len=12_160_000, haplotypes=1, seed=101, snp=0.0, indel=0.0, read_seed=102, read_len=150,
                 pairs=50_000, insert=(400, 50), err=0.01,
                 desc="cfg1 stand-in: 12.16 Mbp single genome, ms on 100k x 150 bp reads, 1% substitutions, half reverse-complemented"),
    "cfg2": dict(base_len=61_400_000, haplotypes=7, seed=201, snp=5e-3, indel=5e-4, read_seed=203, read_len=150,
                 pairs=1_000_000, insert=(400, 50), err=0.01,
                 desc="cfg2 stand-in: 61.4 Mbp x 7 haplotypes, mems on 1M x 2 x 150 bp paired-end reads"),
    # cfg2 scaled so that the two index files fit the 512 MiB gpurun snapshot (the full-size index takes ~13 min of
    # the reference's pfp_thresholds to build and cannot be shipped): same haplotype model, same reads.
    "cfg2s": dict(base_len=28_000_000, haplotypes=7, seed=201, snp=5e-3, indel=5e-4, read_seed=203, read_len=150,
                  pairs=1_000_000, insert=(400, 50), err=0.01,
                  desc="cfg2 stand-in scaled to the snapshot limit: 28 Mbp x 7 haplotypes, mems on 1M x 2 x 150 bp paired-end reads"),
    # BASELINE.json configs[2]: "synthetic 32-haplotype chr21-scale pangenome, MS+MEMs on 10M 150 bp reads, 1 B200".  The index
    # (~1.28 Gbp) is too large for the gpurun snapshot, so it is built on the GPU box by tests/scale/scale_run.py.
    "cfg3": dict(base_len=40_000_000, haplotypes=32, seed=301, snp=1e-3, indel=1e-4, read_seed=302, read_len=150,
                 pairs=5_000_000, bench_pairs=1_000_000, insert=(400, 50), err=0.01,
                 desc="cfg3: 40 Mbp x 32 haplotypes (SNP 1e-3, indel 1e-4), 10M x 150 bp reads, 1% substitutions"),
    # BASELINE.json configs[3]/[4]: "synthetic 128-haplotype human-chromosome-scale pangenome (~6 Gbp text), 100M reads" and
    # "same index, 10 kbp ONT-like reads".  A 6 Gbp index takes the reference's CPU builder hours and ~100 GB of RAM, so the
    # haplotype count and the read counts are kept and the chromosome is scaled down: 128 x 4 Mbp = 512 Mbp.
    "cfg4s": dict(base_len=4_000_000, haplotypes=128, seed=401, snp=1e-3, indel=1e-4, read_seed=402, read_len=150,
                  pairs=50_000_000, bench_pairs=1_000_000, insert=(400, 50), err=0.01,
                  desc="cfg4 scaled: 4 Mbp x 128 haplotypes (SNP 1e-3, indel 1e-4), 100M x 150 bp reads, 1% substitutions"),
    # n >= 2^32: the configs[3]/[4] regime of 40-bit positions (reference include/common/common.hpp:64-65, include/ms/ms_pointers.hpp:164-200)
    # on a text the reference's CPU builder finishes in tens of minutes: 4096 near-identical copies of a 1.1 Mbp base, n = 4.5 G.
    "cfg4x": dict(base_len=1_100_000, haplotypes=4096, seed=411, snp=1e-4, indel=1e-5, read_seed=412, read_len=150,
                  pairs=1_000_000, bench_pairs=1_000_000, insert=(400, 50), err=0.01,
                  desc="cfg4x: 1.1 Mbp x 4096 haplotypes (SNP 1e-4, indel 1e-5), n = 4.5 G > 2^32, 150 bp reads, 1% substitutions"),
    "small": dict(base_len=2_000_000, haplotypes=4, seed=5, snp=2e-3, indel=2e-4, read_seed=6, read_len=150,
                  pairs=100_000, insert=(400, 50), err=0.01, desc="functional check: 2 Mbp x 4 haplotypes"),
}


def index_prefix(name: str) -> str:
    return os.path.join(ROOT, "data", name, "ref.fa")


def text_of(name: str) -> np.ndarray:
    w = WORKLOADS[name]
    return synth.text_of(synth.pangenome(w["base_len"], w["haplotypes"], w["seed"], w["snp"], w["indel"]))


def ensure_index(name: str, verbose: bool = True) -> str:
    """Returns the index prefix, building it with the reference builder if it is not cached."""
    pre = index_prefix(name)
    if os.path.exists(pre + ".thrbv.ms") and os.path.exists(pre + ".plain.slp"):
        return pre
    w = WORKLOADS[name]
    os.makedirs(os.path.dirname(pre), exist_ok=True)
    if verbose:
        print(f"[workloads] building the {name} index with the reference builder (one-off, cached in data/{name})",
              file=sys.stderr, flush=True)
    recs = synth.pangenome(w["base_len"], w["haplotypes"], w["seed"], w["snp"], w["indel"])
    synth.write_fasta(pre, recs)
    del recs
    build_index.build_index(pre, verbose=verbose)
    os.remove(pre)
    return pre


def paired_reads(text: np.ndarray, pairs: int, length: int, seed: int, insert=(400, 50), err: float = 0.01,
                 chunk: int = 250_000) -> np.ndarray:
    """(2 * pairs, length) uint8: rows [0, pairs) are R1 (forward), rows [pairs, 2 pairs) the matching R2
    (reverse complement of the far end of the fragment); each base substituted w.p. err."""
    rng = np.random.default_rng(seed)
    n = text.shape[0]
    out = np.empty((2 * pairs, length), dtype=np.uint8)
    ar = np.arange(length)
    for s in range(0, pairs, chunk):
        k = min(chunk, pairs - s)
        ins = np.clip(rng.normal(insert[0], insert[1], size=k).astype(np.int64), length, None)
        start = rng.integers(0, n - int(ins.max()) - 1, size=k)
        r1 = text[start[:, None] + ar[None, :]]
        r2 = text[(start + ins - length)[:, None] + ar[None, :]]
        r2 = synth._COMP[r2[:, ::-1]]
        for blk in (r1, r2):
            mask = rng.random(blk.shape) < err
            if mask.any():
                blk[mask] = synth._different_base(blk[mask], rng)
        out[s:s + k] = r1
        out[pairs + s:pairs + s + k] = r2
    return out


def reads_of(name: str, text: np.ndarray | None = None, pairs: int | None = None, seed_shift: int = 0) -> np.ndarray:
    w = WORKLOADS[name]
    if text is None:
        text = text_of(name)
    return paired_reads(text, pairs or w["pairs"], w["read_len"], w["read_seed"] + seed_shift, w["insert"], w["err"])
