#!/usr/bin/env python
"""BASELINE.json configs[4] at its stated size on the GPU box: 1 M ONT-like reads (mean 10 kbp, 15 % errors: 5 % substitutions,
5 % insertions, 5 % deletions) on the 128-haplotype index.

    python tests/scale/long_run.py [--workload cfg4s] [--reads 1000000] [--out gpurun_out/long_run.json]

Stages (each records its outcome as soon as it is known):
  reads     generated in parallel, 5000 reads per task, seed 501 + task (the same bytes for any worker count)
  e2e       moni_gpu_mems_packed over ALL reads with pinned host buffers, once with one read per lane (dynamic=0) and once with
            lanes pulling reads (dynamic=1; the default for ragged batches): wall time, bases/s, hash of the complete result;
            moni_gpu_ms_runs_packed on a prefix of the reads (expanded on the host and compared with moni_gpu_ms on a sample)
  parity    the first 20 000 reads as FASTQ through the UNMODIFIED reference `ms` / `mems` (-t nproc) and through the drop-in
            moni_align_b200/bin/{ms,mems}: output files must be byte-identical ("cfg 5: first 20 k")
Checking at scale (test infrastructure: runs the reference binaries as the checker); nothing here is on the product path.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile
import time
from concurrent.futures import ProcessPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
sys.path.insert(0, os.path.join(ROOT, "tests", "scale"))
import synth  # noqa: E402
import workloads  # noqa: E402

RES: dict = {}
OUT = None
TASK = 5000
_TEXT = None


def save():
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    with open(OUT, "w") as f:
        json.dump(RES, f, indent=1)


def log(*a):
    print("[long_run]", *a, file=sys.stderr, flush=True)


def _task(args):
    k, count = args
    lr = synth.long_reads(_TEXT, count, 501 + k)
    lens = np.fromiter((r.shape[0] for r in lr), dtype=np.int64, count=len(lr))
    return k, lens, np.concatenate(lr)


def main():
    global OUT, _TEXT
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg4s")
    ap.add_argument("--reads", type=int, default=1_000_000)
    ap.add_argument("--parity-reads", type=int, default=20_000)
    ap.add_argument("--runs-reads", type=int, default=200_000)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "long_run.json"))
    ap.add_argument("--opt", action="append", default=[], help="library option key=value, repeatable (e.g. lanes=8)")
    ap.add_argument("--dynamic", default="0,1", help="comma list of the dynamic settings to time")
    args = ap.parse_args()
    OUT = args.out
    import moni_align_b200 as m
    from scale_run import cli_vs_reference, hash64
    cores = os.cpu_count() or 1
    for kv in args.opt:
        k_, _, v_ = kv.partition("=")
        m.set_option(k_, v_)
    RES["options"] = args.opt
    w = workloads.WORKLOADS[args.workload]
    RES["workload"] = w["desc"] + f"; {args.reads} ONT-like reads (mean 10 kbp, 5/5/5 % sub/ins/del)"
    RES["host_cores"] = cores
    t0 = time.time()
    prefix = workloads.ensure_index(args.workload)
    RES["index_files_s"] = round(time.time() - t0, 1)
    _TEXT = workloads.text_of(args.workload)

    t0 = time.time()
    tasks = [(k, min(TASK, args.reads - k * TASK)) for k in range((args.reads + TASK - 1) // TASK)]
    parts = [None] * len(tasks)
    with ProcessPoolExecutor(max_workers=cores) as ex:       # workers are forked: they see _TEXT
        for k, lens, flat in ex.map(_task, tasks, chunksize=1):
            parts[k] = (lens, flat)
    lens = np.concatenate([p[0] for p in parts])
    n = lens.shape[0]
    total = int(lens.sum())
    h_seqs = m.pinned_empty(total, np.uint8)
    at = 0
    for _, flat in parts:
        h_seqs[at:at + flat.shape[0]] = flat
        at += flat.shape[0]
    del parts
    h_offs = m.pinned_empty(n + 1, np.uint64)
    h_offs[0] = 0
    np.cumsum(lens, out=h_offs[1:])
    RES["reads"] = {"n": int(n), "bases": total, "max_len": int(lens.max()), "mean_len": float(lens.mean()), "prepare_s": round(time.time() - t0, 1)}
    log(RES["reads"])
    save()

    cap = total + 1024               # 15 % errors: a MEM record every other base (0.51 per base measured); one per base always suffices
    mo = m.pinned_empty(n + 1, np.uint64)
    rec = m.pinned_empty(cap, np.uint32)
    RES["e2e"] = {}
    hashes = {}
    # 0 / 1, or 1xx = persistent kernels sized for xx reads per thread, or L<lanes>R<reads per thread> (e.g. L8R3)
    def parse_dyn(tok):
        if tok.startswith("L"):                 # L<lanes>R<reads per thread>[P<ramp 0|1>]
            l_, r_ = tok[1:].split("R")
            ramp, guided = 0, 1
            if "G" in r_:
                r_, g_ = r_.split("G")
                guided = int(g_)
            if "P" in r_:
                r_, p_ = r_.split("P")
                ramp = int(p_)
            m.set_option("ramp", ramp)
            m.set_option("guided", guided)
            return tok, 1, int(r_), int(l_)
        v = int(tok)
        return v, min(v, 1), (v - 100 if v >= 100 else 2), 0
    dyns = []
    for tok in args.dynamic.split(","):
        dyn, dflag, rpt, lanes = parse_dyn(tok)
        dyns.append(dyn)
    for tok in args.dynamic.split(","):
        dyn, dflag, rpt, lanes = parse_dyn(tok)
        m.set_option("dynamic", dflag)
        for k_, v_ in (("dyn_reads_per_lane", rpt), ("lanes", lanes)):
            try:
                m.set_option(k_, v_)
            except m.MoniGpuError:          # an A/B library of an older tree (MONI_GPU_LIB) does not know the option
                pass
        t = time.time()
        ix = m.MoniIndex(prefix, gpus=[0])
        load_s = time.time() - t
        m.set_option("dynamic", 2)
        try:
            m.set_option("lanes", 0)
        except m.MoniGpuError:
            pass
        try:
            ix.maxrimal_exact_matches_packed(h_seqs, h_offs, out=(mo, rec))
            t = time.perf_counter()
            mo_, rec_ = ix.maxrimal_exact_matches_packed(h_seqs, h_offs, out=(mo, rec))
            dt = time.perf_counter() - t
            hashes[dyn] = hash64(mo_, rec_)
            RES["e2e"][f"mems_packed_dynamic{dyn}"] = {"wall_ms": dt * 1e3, "reads_per_s": n / dt, "bases_per_s": total / dt, "records": int(mo[n]),
                                                      "timing": ix.timing(), "hash": hashes[dyn], "index_load_s": round(load_s, 1)}
        except Exception as e:
            RES["e2e"][f"mems_packed_dynamic{dyn}"] = {"error": f"{type(e).__name__}: {e}"}
        log(dyn, json.dumps(RES["e2e"][f"mems_packed_dynamic{dyn}"])[:600])
        save()
        if dyn != dyns[-1]:
            ix.close()
    RES["e2e"]["same_hash_dynamic_0_1"] = bool(len(hashes) == 2 and len(set(hashes.values())) == 1) if len(dyns) == 2 else None

    # ms as restart runs on a prefix; expanded on the host and compared with moni_gpu_ms on a sample of it
    try:
        k = min(n, args.runs_reads)
        tk = int(h_offs[k])
        rcap = tk + 1024
        rlo, rhi = m.pinned_empty(rcap, np.uint32), m.pinned_empty(rcap, np.uint8)
        ix.matching_statistics_runs(h_seqs[:tk], h_offs[:k + 1], out=(mo[:k + 1], rec[:rcap], rlo, rhi), packed=True)
        t = time.perf_counter()
        ro, rr, lo_, hi_ = ix.matching_statistics_runs(h_seqs[:tk], h_offs[:k + 1], out=(mo[:k + 1], rec[:rcap], rlo, rhi), packed=True)
        dt = time.perf_counter() - t
        RES["e2e"]["ms_runs_packed"] = {"reads": int(k), "bases": tk, "wall_ms": dt * 1e3, "bases_per_s": tk / dt, "runs": int(ro[k]),
                                        "d2h_bytes": int(9 * int(ro[k]) + 8 * (k + 1)), "timing": ix.timing()}
        ks = min(k, 2000)
        ts = int(h_offs[ks])
        q, l, p = m.unpack_runs(rr[:int(ro[ks])], lo_[:int(ro[ks])], hi_[:int(ro[ks])])
        xp, xl = m.expand_runs(h_offs[:ks + 1], ro[:ks + 1], q, l, p)
        wp, wl = ix.matching_statistics(h_seqs[:ts], h_offs[:ks + 1])
        RES["e2e"]["ms_runs_packed"]["expands_to_ms_on_2000_reads"] = bool(np.array_equal(xp, wp) and np.array_equal(xl, wl))
    except Exception as e:
        import traceback
        traceback.print_exc()
        RES["e2e"]["ms_runs_packed"] = {"error": f"{type(e).__name__}: {e}"}
    log(json.dumps(RES["e2e"].get("ms_runs_packed"))[:600])
    save()
    ix.close()

    try:
        k = min(n, args.parity_reads)
        lst = [h_seqs[int(h_offs[r]):int(h_offs[r + 1])].copy() for r in range(k)]
        with tempfile.TemporaryDirectory() as td:
            RES["parity"] = cli_vs_reference(prefix, lst, td, "long", cores)
    except Exception as e:
        import traceback
        traceback.print_exc()
        RES["parity"] = {"error": f"{type(e).__name__}: {e}"}
    save()
    print(json.dumps(RES))


if __name__ == "__main__":
    main()
