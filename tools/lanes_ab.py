#!/usr/bin/env python
"""A/B of the number of chunks in flight per GPU (option "lanes") on one call: wall time of `calls` back-to-back calls per setting,
settings interleaved in rounds so that box noise hits all of them alike.
    python tools/lanes_ab.py [workload] [api] [calls per round] [rounds] [lanes,lanes,...]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import moni_align_b200 as m  # noqa: E402
import workloads  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
api = sys.argv[2] if len(sys.argv) > 2 else "mems_packed16"
calls = int(sys.argv[3]) if len(sys.argv) > 3 else 10
rounds = int(sys.argv[4]) if len(sys.argv) > 4 else 3
lanes = [int(x) for x in (sys.argv[5] if len(sys.argv) > 5 else "4,6,8").split(",")]
pre = workloads.ensure_index(wl)
w = workloads.WORKLOADS[wl]
reads = workloads.reads_of(wl, pairs=w.get("bench_pairs", w["pairs"]))
n, L = reads.shape
total = n * L
hs = m.pinned_empty(total, np.uint8, gpu=0)
hs[:] = reads.reshape(-1)
ho = m.pinned_empty(n + 1, np.uint64, gpu=0)
ho[:] = np.arange(n + 1, dtype=np.uint64) * np.uint64(L)
mo = m.pinned_empty(n + 1, np.uint64, gpu=0)
pa, pb = m.pinned_empty(total, np.uint64, gpu=0), m.pinned_empty(total, np.uint64, gpu=0)
ixs = {}
for l in lanes:
    m.set_option("lanes", l)
    ixs[l] = m.MoniIndex(pre, gpus=[0])
m.set_option("lanes", 0)


def call(ix):
    if api == "mems_packed16":
        return ix.maxrimal_exact_matches_packed16(hs, ho, out=(mo, pa.view(np.uint16)[:total]))
    if api == "mems":
        return ix.maxrimal_exact_matches(hs, ho, out=(mo, pa, pb))
    return ix.matching_statistics_runs(hs, ho, out=(mo, pa.view(np.uint32)[:total], pa.view(np.uint32)[total:2 * total], pb.view(np.uint8)[:total]), packed=True)


res = {l: [] for l in lanes}
for l in lanes:
    for _ in range(3):
        call(ixs[l])
for _ in range(rounds):
    for l in lanes:
        for _ in range(calls):
            t = time.perf_counter()
            call(ixs[l])
            res[l].append((time.perf_counter() - t) * 1e3)
out = {"workload": wl, "api": api, "reads_per_call": int(n),
       "lanes": {str(l): {"calls": len(v), "min_ms": min(v), "median_ms": float(np.median(v)), "mean_ms": float(np.mean(v)), "max_ms": max(v)} for l, v in res.items()}}
print(json.dumps(out))
