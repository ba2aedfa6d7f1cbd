#!/usr/bin/env python
"""Per-chunk timeline (MONI_GPU_TRACE) and per-call wall times of one C-ABI call on a bench workload.
    python tools/trace_run.py [workload] [api] [calls]      api: mems | mems_packed | mems_packed16 | ms | ms_packed | runs | runs_packed"""
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
api = sys.argv[2] if len(sys.argv) > 2 else "mems_packed"
calls = int(sys.argv[3]) if len(sys.argv) > 3 else 8
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
ix = m.MoniIndex(pre, gpus=[0])
u32a, u32b = pa.view(np.uint32)[:total], pa.view(np.uint32)[total:2 * total]
call = {
    "mems": lambda: ix.maxrimal_exact_matches(hs, ho, out=(mo, pa, pb)),
    "mems_packed": lambda: ix.maxrimal_exact_matches_packed(hs, ho, out=(mo, u32a)),
    "mems_packed16": lambda: ix.maxrimal_exact_matches_packed16(hs, ho, out=(mo, pa.view(np.uint16)[:total])),
    "ms": lambda: ix.matching_statistics(hs, ho, out=(pa, pb)),
    "ms_packed": lambda: ix.matching_statistics_packed(hs, ho, out=(u32a, pb.view(np.uint8)[:total], pb.view(np.uint16)[total:2 * total])),
    "runs": lambda: ix.matching_statistics_runs(hs, ho, out=(mo, u32a, u32b, pb)),
    "runs_packed": lambda: ix.matching_statistics_runs(hs, ho, out=(mo, u32a, u32b, pb.view(np.uint8)[:total]), packed=True),
}[api]
for _ in range(3):
    call()
walls = []
for i in range(calls):
    if i == calls - 1:
        os.environ["MONI_GPU_TRACE"] = "1"
    t = time.perf_counter()
    call()
    walls.append((time.perf_counter() - t) * 1e3)
    print(f"[trace_run] {api} call {i}: wall {walls[-1]:.1f} ms, device {ix.timing()['total_ms']:.1f} ms", file=sys.stderr)
print(f"[trace_run] {wl} {api}: wall ms min/median/max {min(walls):.1f} / {float(np.median(walls)):.1f} / {max(walls):.1f}", file=sys.stderr)
