#!/usr/bin/env python
"""The `moni align` seed stage at scale on the GPU box:

    python tests/scale/occ_run.py [--prefix data/cfg4s_full/ref.fa] [--workload cfg4s] [--pairs 1000000]

  seeds        moni_gpu_seeds (seed_finder::find_mems, min_len 25) over the read batch, pinned host buffers
  occurrences  moni_gpu_seed_occs (seed_finder::find_MEM_occs, unfiltered) for every seed: Phi / Phi^-1 walks + bounded LCE
  checks       a sample of seeds against the CPU oracle's restatement; for ALL occurrences T[occ : occ+len] == T[pos : pos+len]
               at three probe offsets, the seed's own position first, no position twice within a seed
The index prefix holds the reference builder's `.thrbv.full.ms` (rlebwt_ms_full_build -n) and `.plain.slp`.
Checking at scale (test infrastructure: it runs the reference binaries and the CPU oracle as the checker); nothing here is on the product path.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import workloads  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--prefix", default=os.path.join(ROOT, "data", "cfg4s_full", "ref.fa"))
    ap.add_argument("--workload", default="cfg4s")
    ap.add_argument("--pairs", type=int, default=1_000_000)
    ap.add_argument("--min-len", type=int, default=25)
    ap.add_argument("--cap", type=int, default=4096)
    ap.add_argument("--oracle-seeds", type=int, default=3000)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "occ_run.json"))
    args = ap.parse_args()
    import moni_align_b200 as m
    res = {"workload": workloads.WORKLOADS[args.workload]["desc"], "prefix": os.path.relpath(args.prefix, ROOT)}
    t = time.time()
    ix = m.MoniIndex(args.prefix, gpus=[0])
    res["load"] = {"seconds": round(time.time() - t, 1), "n": ix.n, "r": ix.r, "hbm_image_bytes": ix.device_bytes}
    text = workloads.text_of(args.workload)
    reads = workloads.reads_of(args.workload, text=text, pairs=args.pairs)
    n_reads, L = reads.shape
    total = n_reads * L
    h_seqs = m.pinned_empty(total, np.uint8)
    h_seqs[:] = reads.reshape(-1)
    h_offs = m.pinned_empty(n_reads + 1, np.uint64)
    h_offs[:] = np.arange(n_reads + 1, dtype=np.uint64) * np.uint64(L)
    cap = total // 8
    out = (m.pinned_empty(n_reads + 1, np.uint64), m.pinned_empty(cap, np.uint64), m.pinned_empty(cap, np.uint32), m.pinned_empty(cap, np.uint32))
    ix.find_mems(h_seqs, h_offs, args.min_len, out=out)
    t = time.perf_counter()
    so, ref, qpos, ln = ix.find_mems(h_seqs, h_offs, args.min_len, out=out)
    dt = time.perf_counter() - t
    n_seeds = int(so[-1])
    res["seeds"] = {"reads": int(n_reads), "min_len": args.min_len, "records": n_seeds, "wall_ms": dt * 1e3, "reads_per_s": n_reads / dt,
                    "timing": ix.timing()}
    print("[occ_run] seeds", json.dumps(res["seeds"]), file=sys.stderr, flush=True)

    before = m.launch_count()
    oo, occs, trunc = ix.find_MEM_occs(ref, ln, max_per_direction=args.cap)
    p_off, p_occ = m.pinned_empty(n_seeds + 1, np.uint64), m.pinned_empty(int(oo[-1]), np.uint64)
    ix.find_MEM_occs(ref, ln, max_per_direction=args.cap, out=(p_off, p_occ))
    t = time.perf_counter()
    oo, occs, trunc = ix.find_MEM_occs(ref, ln, max_per_direction=args.cap, out=(p_off, p_occ))
    dt = time.perf_counter() - t
    cnt = np.diff(oo.astype(np.int64))
    res["occurrences"] = {"seeds": n_seeds, "positions": int(oo[-1]), "per_seed_mean": float(cnt.mean()), "per_seed_max": int(cnt.max()),
                          "cap_per_direction": args.cap, "truncated": bool(trunc), "wall_ms": dt * 1e3, "seeds_per_s": n_seeds / dt,
                          "positions_per_s": int(oo[-1]) / dt, "kernel_launches": int(m.launch_count() - before), "timing": ix.timing()}
    print("[occ_run] occurrences", json.dumps(res["occurrences"]), file=sys.stderr, flush=True)

    # every occurrence spells the seed (three probes per occurrence), own position first, no duplicates inside a seed
    seed_of = np.repeat(np.arange(n_seeds), cnt)
    l_of = ln.astype(np.int64)[seed_of]
    p_of = ref.astype(np.int64)[seed_of]
    o = occs.astype(np.int64)
    ok = bool((o + l_of <= text.shape[0]).all())
    rng = np.random.default_rng(3)
    for probe in (np.zeros_like(l_of), l_of - 1, (rng.random(l_of.shape[0]) * l_of).astype(np.int64)):
        ok = ok and bool((text[o + probe] == text[p_of + probe]).all())
    first = oo[:-1].astype(np.int64)
    ok_first = bool(np.array_equal(occs[first], ref))
    key = seed_of.astype(np.uint64) * np.uint64(1 << 40) + occs
    ok_distinct = bool(np.unique(key).shape[0] == key.shape[0])
    res["checks"] = {"occurrences_spell_the_seed": ok, "own_position_first": ok_first, "distinct_within_seed": ok_distinct}

    from oracle import oracle
    orc = oracle.OracleIndex(args.prefix)
    sel = np.random.default_rng(4).choice(n_seeds, size=min(args.oracle_seeds, n_seeds), replace=False)
    t = time.time()
    woo, wocc = orc.mem_occs(ref[sel], ln[sel])
    same = True
    for j, s in enumerate(sel):
        same = same and np.array_equal(wocc[int(woo[j]):int(woo[j + 1])], occs[int(oo[s]):int(oo[s + 1])])
    res["checks"]["identical_to_oracle"] = {"seeds": int(sel.shape[0]), "identical": bool(same), "oracle_s": round(time.time() - t, 1)}
    print("[occ_run] checks", json.dumps(res["checks"]), file=sys.stderr, flush=True)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(res, open(args.out, "w"), indent=1)
    print(json.dumps(res))
    good = ok and ok_first and ok_distinct and same
    sys.exit(0 if good else 1)


if __name__ == "__main__":
    main()
