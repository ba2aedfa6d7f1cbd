"""Parsers of the reference's text outputs (.pointers / .lengths: src/matching_statistics.cpp:569-598; .mems: src/mems.cpp:585-587)
into arrays, so that a result computed elsewhere can be compared with them number by number.  Checker-side helper (bench.py's
parity_sample, tests/scale): nothing on the product path imports it."""
from __future__ import annotations

import numpy as np


def _numbers(blob: bytes) -> np.ndarray:
    if not blob.strip():
        return np.zeros(0, dtype=np.uint64)
    return np.array(blob.split(), dtype=np.uint64)


def parse_values(path: str):
    """(names, counts uint64[n], values uint64[total]) of a .pointers / .lengths file."""
    names, counts, parts = [], [], []
    with open(path, "rb") as f:
        lines = f.read().split(b"\n")
    for k in range(0, len(lines) - 1, 2):
        names.append(lines[k][1:])
        parts.append(lines[k + 1])
        counts.append(lines[k + 1].count(b" "))
    vals = _numbers(b" ".join(parts))
    return names, np.asarray(counts, dtype=np.uint64), vals


def parse_mems(path: str):
    """(names, counts uint64[n], query offsets uint64[k], lengths uint64[k]) of a .mems file."""
    names, counts, parts = [], [], []
    with open(path, "rb") as f:
        lines = f.read().split(b"\n")
    for k in range(0, len(lines) - 1, 2):
        names.append(lines[k][1:])
        parts.append(lines[k + 1])
        counts.append(lines[k + 1].count(b"("))
    blob = b" ".join(parts).replace(b"(", b" ").replace(b")", b" ").replace(b",", b" ")
    flat = _numbers(blob)
    return names, np.asarray(counts, dtype=np.uint64), flat[0::2], flat[1::2]
