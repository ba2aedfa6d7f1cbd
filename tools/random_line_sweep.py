#!/usr/bin/env python
"""Random-access roofline of the box (SURVEY.md 8(d) denominator): sweeps moni_gpu_probe_random_lines over table sizes,
landing shapes and chains in flight per SM, and prints one JSON document.

    python tools/random_line_sweep.py [--out profiles/r02_random_lines.json] [--tables 1.5,3,12]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

# (shape, chains per thread, 256-thread blocks per SM) -- the combinations probe.cu instantiates
CONFIGS = [(0, 1, 4), (0, 1, 6), (0, 1, 8), (0, 2, 4), (0, 2, 6), (0, 4, 3), (0, 4, 4),
           (1, 1, 8), (1, 2, 6), (1, 2, 8), (1, 4, 4), (1, 4, 6), (1, 4, 8), (1, 8, 4), (1, 8, 6),
           (2, 2, 8), (2, 4, 8), (2, 6, 8), (2, 8, 4), (2, 8, 6), (2, 8, 8), (2, 12, 4), (2, 12, 6), (2, 16, 4)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ap.add_argument("--tables", default="1.5,3,12")
    ap.add_argument("--iters", type=int, default=512)
    ap.add_argument("--gpu", type=int, default=0)
    a = ap.parse_args()
    import moni_align_b200 as m
    res = []
    for gib in [float(x) for x in a.tables.split(",")]:
        tb = int(gib * (1 << 30)) & ~127
        for shape, k, b in CONFIGS:
            try:
                r = m.probe_random_lines(a.gpu, tb, shape, k, b, a.iters)
            except m.MoniGpuError as e:
                r = {"error": str(e), "shape": shape, "chains_per_thread": k, "blocks_per_sm": b}
            r.update({"table_gib": gib, "chains_per_thread": k, "blocks_per_sm_asked": b})
            if "lines_per_s" in r:
                r["g_lines_per_s"] = round(r["lines_per_s"] / 1e9, 2)
                r["line_traffic_gb_per_s"] = round(r["lines_per_s"] * 128 / 1e9, 1)
            print(json.dumps(r), file=sys.stderr, flush=True)
            res.append(r)
        m.probe_release()
    doc = {"what": "dependent random 32-byte accesses, one 128-byte DRAM line each; lines per second", "iters_per_chain": a.iters, "results": res}
    best = {}
    for r in res:
        if "lines_per_s" in r:
            g = r["table_gib"]
            if g not in best or r["lines_per_s"] > best[g]["lines_per_s"]:
                best[g] = r
    doc["best_per_table"] = {str(g): r for g, r in best.items()}
    s = json.dumps(doc, indent=1)
    if a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        open(a.out, "w").write(s)
    print(s)


if __name__ == "__main__":
    main()
