#!/usr/bin/env python
"""Parity + throughput of the BASELINE.json configurations that are not the bench line (run on the GPU box):

    python tests/scale/config_runs.py [--workload cfg2s] [--out gpurun_out/config_runs.json]

  cfg1-like  `ms`   on 100k x 150 bp reads (1 % substitutions, half reverse-complemented)
  cfg5-like  `ms` + `mems` on ragged ONT-like reads (mean 10 kbp, 15 % errors): warp divergence, long scans, the
             length-sorted ragged path
Each case: results through the C ABI with pinned host buffers, wall-clock throughput, and a differential check of a
read sample against the UNMODIFIED reference binaries (oracle/_ref/bin/ms, mems) on the same index files -- the
files the reference writes are compared byte for byte with the ones formatted from the GPU results.
Checking at scale (test infrastructure: it runs the reference binaries and the CPU oracle as the checker); nothing here is on the product path.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import build_index  # noqa: E402
import synth  # noqa: E402
import workloads  # noqa: E402


def reference_files(prefix, reads, cmd, threads):
    """Runs the reference binary on `reads` (list of arrays); returns {ext: bytes} and its processing-phase seconds."""
    with tempfile.TemporaryDirectory() as td:
        fq = os.path.join(td, "s.fq")
        synth.write_fastq_list(fq, reads)
        out = os.path.join(td, "o")
        r = build_index.run_reference(cmd, prefix, fq, out, threads=threads)
        exts = [".pointers", ".lengths"] if cmd == "ms" else [".mems"]
        return {e: open(out + e, "rb").read() for e in exts}, r["phase"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg2s")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "config_runs.json"))
    ap.add_argument("--short-reads", type=int, default=100_000)
    ap.add_argument("--long-reads", type=int, default=20_000)
    ap.add_argument("--check-short", type=int, default=20_000)
    ap.add_argument("--check-long", type=int, default=300)
    args = ap.parse_args()
    import moni_align_b200 as m

    prefix = workloads.ensure_index(args.workload)
    text = workloads.text_of(args.workload)
    cores = os.cpu_count() or 1
    ix = m.MoniIndex(prefix, gpus=[0])
    results = {"workload": workloads.WORKLOADS[args.workload]["desc"], "index_n": ix.n, "index_runs": ix.r, "host_cores": cores, "cases": []}

    def run_case(name, reads, check_n, do_ms, do_mems):
        flat, offs = m.pack_reads(reads)
        total, n = int(offs[-1]), len(offs) - 1
        names = [b"r%d" % k for k in range(n)]
        h_seqs = m.pinned_empty(total, np.uint8)
        h_seqs[:] = flat
        h_offs = m.pinned_empty(n + 1, np.uint64)
        h_offs[:] = offs
        case = {"name": name, "reads": n, "bases": total, "max_len": int(np.diff(offs.astype(np.int64)).max())}
        sample = [np.asarray(r) for r in (reads[:check_n] if not isinstance(reads, np.ndarray) else list(reads[:check_n]))]
        s_off = offs[:check_n + 1]
        if do_ms:
            out = (m.pinned_empty(total, np.uint64), m.pinned_empty(total, np.uint64))
            ix.matching_statistics(h_seqs, h_offs, out=out)
            t = time.perf_counter()
            ptr, ln = ix.matching_statistics(h_seqs, h_offs, out=out)
            dt = time.perf_counter() - t
            ref, ref_s = reference_files(prefix, sample, "ms", cores)
            ok = (m.format_ms(names[:check_n], s_off, ptr) == ref[".pointers"] and m.format_ms(names[:check_n], s_off, ln) == ref[".lengths"])
            case["ms"] = {"reads_per_s": n / dt, "bases_per_s": total / dt, "ms": dt * 1e3, "timing": ix.timing(),
                          "identical_to_reference_files": bool(ok), "checked_reads": check_n,
                          "reference_reads_per_s": check_n / ref_s, "reference_bases_per_s": int(s_off[-1]) / ref_s, "reference_threads": cores}
            print(name, "ms", case["ms"], file=sys.stderr, flush=True)
        if do_mems:
            out = (m.pinned_empty(n + 1, np.uint64), m.pinned_empty(total, np.uint32), m.pinned_empty(total, np.uint32))
            ix.maxrimal_exact_matches(h_seqs, h_offs, out=out)
            t = time.perf_counter()
            mo, mp, ml = ix.maxrimal_exact_matches(h_seqs, h_offs, out=out)
            dt = time.perf_counter() - t
            ref, ref_s = reference_files(prefix, sample, "mems", cores)
            ok = m.format_mems(names[:check_n], mo, mp, ml) == ref[".mems"]
            case["mems"] = {"reads_per_s": n / dt, "bases_per_s": total / dt, "ms": dt * 1e3, "timing": ix.timing(),
                            "identical_to_reference_files": bool(ok), "checked_reads": check_n, "records": int(mo[n]),
                            "reference_reads_per_s": check_n / ref_s, "reference_bases_per_s": int(s_off[-1]) / ref_s, "reference_threads": cores}
            print(name, "mems", case["mems"], file=sys.stderr, flush=True)
        results["cases"].append(case)
        m.release_pinned()

    short = synth.short_reads(text, args.short_reads, 150, 102, err=0.01)
    run_case("cfg1-like: ms on 150 bp reads, 1% substitutions", short, args.check_short, True, False)
    longr = synth.long_reads(text, args.long_reads, 501)
    run_case("cfg5-like: ragged ONT-like reads, mean 10 kbp, 15% errors", longr, args.check_long, True, True)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(results, open(args.out, "w"), indent=1)
    print(json.dumps(results))


if __name__ == "__main__":
    main()
