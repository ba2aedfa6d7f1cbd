import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
for p in (ROOT, os.path.join(ROOT, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)

CASES = ["pan4", "withN", "twoletter"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def read_fastq(path):
    """4-line FASTQ -> (names, seqs); name = text up to the first whitespace, like kseq."""
    lines = open(path, "rb").read().split(b"\n")
    names, seqs = [], []
    for i in range(0, len(lines) - 3, 4):
        if not lines[i].startswith(b"@"):
            break
        names.append(lines[i][1:].split()[0])
        seqs.append(lines[i + 1])
    return names, seqs


def pack(seqs):
    lens = np.fromiter((len(s) for s in seqs), dtype=np.uint64, count=len(seqs))
    offs = np.zeros(len(seqs) + 1, dtype=np.uint64)
    np.cumsum(lens, out=offs[1:])
    flat = np.frombuffer(b"".join(seqs), dtype=np.uint8).copy() if int(offs[-1]) else np.zeros(0, dtype=np.uint8)
    return flat, offs


@pytest.fixture(scope="session")
def host_sim():
    """g++ build of tests/host_sim.cpp (the per-read core + exporter on the CPU, test-only)."""
    out = os.path.join(ROOT, "tests", "_build", "host_sim")
    srcs = [os.path.join(ROOT, "tests", "host_sim.cpp"), os.path.join(ROOT, "moni_align_b200", "csrc", "index_files.cpp"),
            os.path.join(ROOT, "moni_align_b200", "csrc", "device_image.cpp")]
    deps = srcs + [os.path.join(ROOT, "moni_align_b200", "csrc", f) for f in ("ms_core.cuh", "device_image.hpp", "index_files.hpp")]
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(d) for d in deps):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-pthread", "-x", "c++", *srcs, "-o", out])
    return out


def run_host_sim(exe, prefix, flat, offs, tmpdir, extra=()):
    b = os.path.join(str(tmpdir), "batch.bin")
    o = os.path.join(str(tmpdir), "out.bin")
    with open(b, "wb") as f:
        f.write(np.uint64(len(offs) - 1).tobytes())
        f.write(offs.astype(np.uint64).tobytes())
        f.write(flat.tobytes())
    subprocess.check_call([exe, prefix, b, o, *map(str, extra)], stderr=subprocess.DEVNULL)
    raw = np.fromfile(o, dtype=np.uint64)
    total, nr = int(offs[-1]), len(offs) - 1
    p, l = raw[:total], raw[total:2 * total]
    mo = raw[2 * total:2 * total + nr + 1]
    k = int(mo[-1])
    base = 2 * total + nr + 1
    res = (p, l, mo, raw[base:base + k], raw[base + k:base + 2 * k], raw[base + 2 * k:base + 2 * k + 8])
    if len(extra) < 3:
        return res
    at = base + 2 * k + 8              # seeds: offsets, ref, len, qpos
    so = raw[at:at + nr + 1]
    ks = int(so[-1]) if nr >= 0 and so.shape[0] else 0
    at += nr + 1
    res = res + (so, raw[at:at + ks], raw[at + ks:at + 2 * ks], raw[at + 2 * ks:at + 3 * ks])
    at += 3 * ks
    if at < raw.shape[0]:              # occurrence lists (index with Phi structures, text mode)
        oo = raw[at:at + ks + 1]
        res = res + (oo, raw[at + ks + 1:at + ks + 1 + int(oo[-1])])
    return res


def full_only_prefix(case, tmpdir, lcp=False):
    """A prefix under tmpdir that holds ONLY a `moni align` index of a golden case (+ the grammar): .thrbv.full.ms, or the
    sampled-LCP variant .thrbv.full.lcp.ms."""
    import shutil
    ext = ".thrbv.full.lcp.ms" if lcp else ".thrbv.full.ms"
    pre = os.path.join(str(tmpdir), case + ("_lcp.fa" if lcp else "_full.fa"))
    shutil.copy(os.path.join(GOLDEN, case + ".fa" + ext), pre + ext)
    shutil.copy(os.path.join(GOLDEN, case + ".fa.plain.slp"), pre + ".plain.slp")
    return pre


@pytest.fixture(scope="session")
def oracle_mod():
    from oracle import oracle
    oracle.build()
    return oracle
