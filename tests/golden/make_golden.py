"""Generates the committed golden fixtures by running the UNMODIFIED reference binaries
(oracle/_ref/bin, built from /root/reference by oracle/build_ref.sh) in THIS container.

    python tests/golden/make_golden.py

For each case: a tiny seeded FASTA -> newscanNT.x / pfp_thresholds / rlebwt_ms_build / SlpEncBuild
(index files `<case>.fa.thrbv.ms`, `<case>.fa.plain.slp` + the builder intermediates the parser tests
read) and a FASTQ with regular and edge-case reads -> reference `ms` and `mems` outputs
(`<case>.pointers`, `<case>.lengths`, `<case>.mems`).  Everything is a few hundred KB.
"""
import os
import shutil
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import build_index  # noqa: E402
import synth  # noqa: E402

KEEP = [".thrbv.ms", ".plain.slp", ".bwt.heads", ".bwt.len", ".thr_pos", ".ssa", ".esa"]


def edge_reads(text: np.ndarray) -> list:
    t = text.tobytes()
    return [
        b"A", b"N", b"NNNNNNNNNNNNNNNNNNNN", t[100:160].lower(), t[200:230] + b"N" + t[231:260],
        t[300:330] + b"#" + t[331:360], t[400:430] + b"~" + t[431:460], b"~" + t[500:540], b"#A#C#G#T~A~C~G~T",
        t[0:300], t[-120:], t[-1:] + t[:50], b"ACGT" * 40, b"T" * 70, t[1000:1100] + t[3000:3100],
        b"acgtNACGT" + t[50:90] + b"n", b"\x01" + t[10:40], t[600:640] + b"\x01" + t[650:690],
    ]


def make_case(name: str, records, seed: int, n_reads: int, trivial_grammar: bool = False, length: int = 150):
    fa = os.path.join(HERE, name + ".fa")
    synth.write_fasta(fa, records)
    build_index.build_index(fa, trivial_grammar=trivial_grammar, keep=True)
    text = synth.text_of(records)
    # newscan upper-cases the text
    text = np.frombuffer(text.tobytes().upper(), dtype=np.uint8)
    reads = [r for r in synth.short_reads(text, n_reads, min(length, text.shape[0] // 2), seed)]
    reads += synth.long_reads(text, 6, seed + 1, mean=900, sd=300, lo=200, hi=2000)
    reads += edge_reads(text)
    fq = os.path.join(HERE, name + ".fq")
    synth.write_fastq_list(fq, reads)
    for cmd in ("ms", "mems"):
        build_index.run_reference(cmd, fa, fq, os.path.join(HERE, name), threads=1)
    for ext in build_index.INTERMEDIATES + [".build.log"]:
        if ext not in KEEP:
            try:
                os.remove(fa + ext)
            except FileNotFoundError:
                pass
    os.remove(fa)
    for f in os.listdir(HERE):
        if f.endswith(".log"):
            os.remove(os.path.join(HERE, f))


def cases():
    """(name, records, make_case keyword arguments) of the three fixtures, in generation order (the rng is shared)."""
    rng = np.random.default_rng(7)
    # 1: small pangenome, ACGT only
    yield "pan4", synth.pangenome(6000, 4, 11, 1e-2, 2e-3), dict(seed=12, n_reads=60)
    # 2: text with N stretches and a fifth frequent symbol -> exercises the general path
    recs = synth.pangenome(5000, 3, 21, 2e-2, 2e-3)
    for r in recs:
        for _ in range(12):
            p = int(rng.integers(0, r.shape[0] - 40))
            r[p:p + int(rng.integers(1, 30))] = ord("N")
        r[rng.integers(0, r.shape[0], size=60)] = ord("R")
    yield "withN", recs, dict(seed=22, n_reads=40)
    # 3: two-letter text (fewer than four table symbols), trivial grammar, long runs
    ab = np.frombuffer(b"AC", dtype=np.uint8)[(rng.random(3000) < 0.2).astype(np.int64)]
    yield "twoletter", [np.concatenate([ab, ab[:2000], ab[500:2500]])], dict(seed=32, n_reads=30, trivial_grammar=True, length=80)


def main():
    for name, records, kw in cases():
        make_case(name, records, **kw)


if __name__ == "__main__":
    main()
    print(sorted(os.listdir(HERE)))
