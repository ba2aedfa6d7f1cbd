"""Index production with the UNMODIFIED reference builder (input preparation, not the hot path).

    FASTA --newscanNT.x -f--> .parse/.dict/.occ --pfp_thresholds -r--> .bwt.heads/.bwt.len/.ssa/.esa/.thr_pos
          --rlebwt_ms_build--> <fa>.thrbv.ms         (reference: pipeline/moni.in:419-491)
    FASTA --tools/mkgrammar--> .R/.C --SlpEncBuild -e PlainSlp_FblcFblc -f Bigrepair--> <fa>.plain.slp

The binaries come from oracle/_ref/bin (built from /root/reference by oracle/build_ref.sh; they
travel to the GPU box with the snapshot).  Both the CPU reference and the GPU library consume the
SAME two files.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "bin")
MKGRAMMAR = os.path.join(ROOT, "tools", "_build", "mkgrammar")

INTERMEDIATES = [".parse", ".parse_old", ".dict", ".occ", ".last", ".bwt.heads", ".bwt.len", ".ssa", ".esa",
                 ".thr", ".thr_pos", ".slcp", ".R", ".C", ".sai"]


def ref_tool(name: str) -> str:
    p = os.path.join(REF_BIN, name)
    if not os.path.exists(p):
        raise FileNotFoundError(
            f"{p} missing: run oracle/build_ref.sh where /root/reference is mounted (the built files travel with gpurun)")
    return p


def build_mkgrammar() -> str:
    src = os.path.join(ROOT, "tools", "mkgrammar.cpp")
    if not os.path.exists(MKGRAMMAR) or os.path.getmtime(MKGRAMMAR) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(MKGRAMMAR), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", src, "-o", MKGRAMMAR])
    return MKGRAMMAR


def _run(cmd, log, cwd=None):
    t0 = time.time()
    with open(log, "ab") as lf:
        lf.write(("\n$ " + " ".join(cmd) + "\n").encode())
        lf.flush()
        r = subprocess.run(cmd, stdout=lf, stderr=subprocess.STDOUT, cwd=cwd)
    if r.returncode != 0:
        tail = open(log, "rb").read()[-2000:].decode(errors="replace")
        raise RuntimeError(f"command failed ({r.returncode}): {' '.join(cmd)}\n{tail}")
    return time.time() - t0


def build_index(fasta: str, trivial_grammar: bool = False, keep: bool = False, wsize: int = 10,
                mod: int = 100, verbose: bool = False) -> dict:
    """Builds <fasta>.thrbv.ms and <fasta>.plain.slp next to the FASTA; returns timings."""
    log = fasta + ".build.log"
    times = {}
    if not (os.path.exists(fasta + ".thrbv.ms")):
        times["newscan"] = _run([ref_tool("newscanNT.x"), fasta, "-w", str(wsize), "-p", str(mod), "-f"], log)
        parse_words = os.path.getsize(fasta + ".parse") // 4
        dict_bytes = os.path.getsize(fasta + ".dict")
        exe = "pfp_thresholds64" if (parse_words >= 2**31 - 1 or dict_bytes >= 2**31 - 4) else "pfp_thresholds"
        times["pfp_thresholds"] = _run([ref_tool(exe), fasta, "-w", str(wsize), "-r"], log)
        times["rlebwt_ms_build"] = _run([ref_tool("rlebwt_ms_build"), fasta], log)
    if not os.path.exists(fasta + ".plain.slp"):
        mk = build_mkgrammar()
        cmd = [mk, fasta, fasta] + (["--trivial"] if trivial_grammar else [])
        times["mkgrammar"] = _run(cmd, log)
        times["SlpEncBuild"] = _run([ref_tool("SlpEncBuild"), "-i", fasta, "-o", fasta + ".plain.slp",
                                     "-e", "PlainSlp_FblcFblc", "-f", "Bigrepair"], log)
    if not keep:
        for ext in INTERMEDIATES:
            try:
                os.remove(fasta + ext)
            except FileNotFoundError:
                pass
    if verbose:
        print("build_index:", fasta, {k: round(v, 2) for k, v in times.items()}, file=sys.stderr)
    return times


def run_reference(cmd: str, index_prefix: str, patterns: str, out_prefix: str, threads: int = 1) -> dict:
    """Runs the reference `ms` or `mems` binary; returns {'wall': s, 'phase': s} where phase is the
    reference's own "Processing patterns" timer (src/matching_statistics.cpp:500-524: the 4th
    'Elapsed time (s)' line of its output)."""
    assert cmd in ("ms", "mems")
    t0 = time.time()
    r = subprocess.run([ref_tool(cmd), index_prefix, "-p", patterns, "-o", out_prefix, "-t", str(threads)],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    wall = time.time() - t0
    out = r.stdout.decode(errors="replace")
    if r.returncode != 0:
        raise RuntimeError(f"reference {cmd} failed:\n{out[-2000:]}")
    el = [float(l.rsplit(":", 1)[1]) for l in out.splitlines() if "Elapsed time (s)" in l]
    return {"wall": wall, "phase": el[3] if len(el) > 3 else None, "elapsed_lines": el, "stdout": out}


if __name__ == "__main__":
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("fasta")
    ap.add_argument("--trivial-grammar", action="store_true")
    ap.add_argument("--keep", action="store_true")
    a = ap.parse_args()
    print(build_index(a.fasta, a.trivial_grammar, a.keep, verbose=True))
