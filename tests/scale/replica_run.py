"""Index replicas on several GPUs of one process: load time with the image copied GPU to GPU after the
first replica, and the same results from 1 and N replicas.     python tests/scale/replica_run.py --gpus 0,1 [--out f.json]"""
import argparse
import hashlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import moni_align_b200 as m  # noqa: E402
import workloads  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg2s")
    ap.add_argument("--gpus", default="0,1")
    ap.add_argument("--reads", type=int, default=2_000_000)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    gpus = [int(x) for x in a.gpus.split(",")]
    pre = workloads.ensure_index(a.workload)
    reads = workloads.reads_of(a.workload, pairs=a.reads // 2)
    flat = np.ascontiguousarray(reads.reshape(-1))
    offs = np.arange(reads.shape[0] + 1, dtype=np.uint64) * np.uint64(reads.shape[1])
    m.set_option("verbose", 1)
    line = {"workload": a.workload, "reads": int(reads.shape[0]), "gpus": gpus}
    for name, g in (("one", gpus[:1]), ("all", gpus)):
        t0 = time.time()
        ix = m.MoniIndex(pre, gpus=g)
        load = time.time() - t0
        ix.maxrimal_exact_matches_packed(flat, offs)                 # warm-up
        t0 = time.time()
        mo, rec = ix.maxrimal_exact_matches_packed(flat, offs)
        dt = time.time() - t0
        h = hashlib.sha256(np.ascontiguousarray(mo).tobytes() + np.ascontiguousarray(rec).tobytes()).hexdigest()
        line[name] = {"gpus": g, "index_load_s": round(load, 3), "mems_packed_pageable_s": round(dt, 4), "sha256": h}
        ix.close()
    line["identical"] = line["one"]["sha256"] == line["all"]["sha256"]
    print(json.dumps(line))
    if a.out:
        open(a.out, "w").write(json.dumps(line) + "\n")
    return 0 if line["identical"] else 1


if __name__ == "__main__":
    sys.exit(main())
