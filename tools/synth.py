"""Seeded synthetic genomes and reads for the `moni ms` / `moni mems` workloads.

Input preparation only — nothing here is on the measured path.  Everything is driven by
``numpy.random.default_rng(seed)`` so the same (name, seed) pair produces the same bytes here
and on the GPU box.
"""
from __future__ import annotations

import os
import numpy as np

ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)
_COMP = np.arange(256, dtype=np.uint8)
for _a, _b in zip(b"ACGTacgt", b"TGCAtgca"):
    _COMP[_a] = _b


def random_genome(n: int, rng: np.random.Generator) -> np.ndarray:
    """i.i.d. uniform ACGT text of n bases (uint8 ASCII)."""
    return ACGT[rng.integers(0, 4, size=n, dtype=np.uint8)]


def _different_base(orig: np.ndarray, rng: np.random.Generator) -> np.ndarray:
    """For each ASCII base in orig draw a uniformly random *different* base."""
    code = np.searchsorted(ACGT, orig)  # ACGT is sorted in ASCII
    code = (code + rng.integers(1, 4, size=orig.shape[0])) % 4
    return ACGT[code]


def mutate(base: np.ndarray, rng: np.random.Generator, snp: float, indel: float,
           max_indel: int = 10) -> np.ndarray:
    """One mutated copy of `base`: SNPs at rate `snp`, indels (length 1..max_indel, geometric,
    half insertions / half deletions) at rate `indel`."""
    out = base.copy()
    n = out.shape[0]
    k = rng.binomial(n, snp)
    if k:
        pos = rng.choice(n, size=k, replace=False)
        out[pos] = _different_base(out[pos], rng)
    k = rng.binomial(n, indel)
    if k == 0:
        return out
    pos = np.sort(rng.choice(n, size=k, replace=False))
    lens = np.minimum(rng.geometric(0.5, size=k), max_indel)
    is_ins = rng.random(k) < 0.5
    pieces = []
    prev = 0
    for p, ln, ins in zip(pos.tolist(), lens.tolist(), is_ins.tolist()):
        if p < prev:
            continue
        pieces.append(out[prev:p])
        if ins:
            pieces.append(random_genome(ln, rng))
            prev = p
        else:
            prev = min(n, p + ln)
    pieces.append(out[prev:])
    return np.concatenate(pieces)


def pangenome(base_len: int, haplotypes: int, seed: int, snp: float, indel: float) -> list[np.ndarray]:
    """`haplotypes` records: a uniform base genome followed by haplotypes-1 independently mutated copies."""
    rng = np.random.default_rng(seed)
    base = random_genome(base_len, rng)
    recs = [base]
    for _ in range(haplotypes - 1):
        recs.append(mutate(base, rng, snp, indel))
    return recs


def write_fasta(path: str, records: list[np.ndarray], width: int = 80) -> None:
    with open(path, "wb") as f:
        for i, rec in enumerate(records):
            f.write(b">rec%d\n" % i)
            n = rec.shape[0]
            full = (n // width) * width
            if full:
                body = np.empty((full // width, width + 1), dtype=np.uint8)
                body[:, :width] = rec[:full].reshape(-1, width)
                body[:, width] = 10
                f.write(body.tobytes())
            if n > full:
                f.write(rec[full:].tobytes() + b"\n")


def text_of(records: list[np.ndarray]) -> np.ndarray:
    """The indexed text: records concatenated without separators (what `newscanNT.x -f` parses)."""
    return np.concatenate(records)


def short_reads(text: np.ndarray, n_reads: int, length: int, seed: int, err: float = 0.01,
                revcomp: float = 0.5) -> np.ndarray:
    """(n_reads, length) uint8 matrix: substrings of `text`, each base substituted w.p. `err`
    by a different base, a fraction `revcomp` reverse-complemented."""
    rng = np.random.default_rng(seed)
    n = text.shape[0]
    starts = rng.integers(0, n - length + 1, size=n_reads)
    idx = starts[:, None] + np.arange(length)[None, :]
    reads = text[idx]
    mask = rng.random(reads.shape) < err
    if mask.any():
        reads[mask] = _different_base(reads[mask], rng)
    rc = rng.random(n_reads) < revcomp
    if rc.any():
        reads[rc] = _COMP[reads[rc][:, ::-1]]
    return reads


def long_reads(text: np.ndarray, n_reads: int, seed: int, mean: float = 10000.0, sd: float = 3000.0,
               lo: int = 1000, hi: int = 30000, sub: float = 0.05, ins: float = 0.05,
               dele: float = 0.05) -> list[np.ndarray]:
    """ONT-like ragged reads (config 5): length ~ clip(N(mean, sd), lo, hi); substitution, insertion
    and deletion errors at the given per-base rates."""
    rng = np.random.default_rng(seed)
    n = text.shape[0]
    out = []
    for _ in range(n_reads):
        ln = int(min(max(rng.normal(mean, sd), lo), hi))
        ln = min(ln, n)
        s = int(rng.integers(0, n - ln + 1))
        r = text[s:s + ln].copy()
        m = rng.random(ln)
        subs = m < sub
        if subs.any():
            r[subs] = _different_base(r[subs], rng)
        keep = ~((m >= sub) & (m < sub + dele))
        insm = (m >= sub + dele) & (m < sub + dele + ins)
        # deletions drop the base, insertions add a random base after it
        rep = keep.astype(np.int64) + insm.astype(np.int64)
        r2 = np.repeat(r, rep)
        # positions of inserted bases: second copy of each inserted-after base
        ends = np.cumsum(rep) - 1
        ipos = ends[insm & keep]
        if ipos.size:
            r2[ipos] = random_genome(ipos.size, rng)
        if r2.size == 0:
            r2 = r[:1]
        out.append(r2)
    return out


def write_fastq_matrix(path: str, reads: np.ndarray, prefix: str = "r") -> None:
    """Uncompressed FASTQ with names r<k>; fixed-length reads given as a matrix."""
    n, m = reads.shape
    qual = b"I" * m
    with open(path, "wb") as f:
        chunk = 100000
        for s in range(0, n, chunk):
            parts = []
            blk = reads[s:s + chunk]
            for k in range(blk.shape[0]):
                parts.append(b"@%s%d\n" % (prefix.encode(), s + k))
                parts.append(blk[k].tobytes())
                parts.append(b"\n+\n")
                parts.append(qual)
                parts.append(b"\n")
            f.write(b"".join(parts))


def write_fastq_list(path: str, reads: list, names: list | None = None) -> None:
    """Uncompressed FASTQ for ragged reads (list of uint8 arrays or bytes)."""
    with open(path, "wb") as f:
        for k, r in enumerate(reads):
            b = r.tobytes() if isinstance(r, np.ndarray) else bytes(r)
            nm = names[k] if names is not None else b"r%d" % k
            if isinstance(nm, str):
                nm = nm.encode()
            f.write(b"@" + nm + b"\n" + b + b"\n+\n" + b"I" * len(b) + b"\n")


def concat_reads(reads) -> tuple[np.ndarray, np.ndarray]:
    """(seqs uint8[total], offsets uint64[n+1]) — the batch layout of the C ABI."""
    if isinstance(reads, np.ndarray) and reads.ndim == 2:
        n, m = reads.shape
        return np.ascontiguousarray(reads).reshape(-1), (np.arange(n + 1, dtype=np.uint64) * np.uint64(m))
    lens = np.fromiter((len(r) for r in reads), dtype=np.uint64, count=len(reads))
    offs = np.zeros(len(reads) + 1, dtype=np.uint64)
    np.cumsum(lens, out=offs[1:])
    if len(reads) == 0:
        return np.zeros(0, dtype=np.uint8), offs
    seqs = np.concatenate([np.frombuffer(bytes(r), dtype=np.uint8) if not isinstance(r, np.ndarray) else r
                           for r in reads]) if offs[-1] else np.zeros(0, dtype=np.uint8)
    return seqs, offs
