"""Shaped grammar at size: the cfg1 index (12.16 Mbp) loaded once through <prefix>.plain.slp and once through
<prefix>.slp (written by the reference encoder, SlpEncBuild -e SelfShapedSlp_SdSd_Sd); ms and mems of the cfg1 reads must be the
same arrays.  Writes one JSON line.

    python tests/scale/shaped_run.py [--out gpurun_out/shaped_run.json]
"""
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
    ap.add_argument("--workload", default="cfg1")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    pre = workloads.index_prefix(a.workload)
    assert os.path.exists(pre + ".slp") and os.path.exists(pre + ".plain.slp"), "needs both grammars next to " + pre
    reads = workloads.reads_of(a.workload)
    flat = np.ascontiguousarray(reads.reshape(-1))
    offs = (np.arange(reads.shape[0] + 1, dtype=np.uint64) * np.uint64(reads.shape[1]))
    res = {}
    for g in ("plain", "shaped"):
        m.set_option("grammar", g)
        t0 = time.time()
        ix = m.MoniIndex(pre)
        load = time.time() - t0
        ptr, ln = ix.matching_statistics(flat, offs)
        mo, mp, ml = ix.maxrimal_exact_matches(flat, offs)
        h = hashlib.sha256()
        for arr in (ptr, ln, mo, mp, ml):
            h.update(np.ascontiguousarray(arr).tobytes())
        res[g] = {"index_load_s": round(load, 3), "results_sha256": h.hexdigest(), "mems": int(mo[-1]),
                  "grammar_file_bytes": os.path.getsize(pre + (".slp" if g == "shaped" else ".plain.slp"))}
        ix.close()
    m.set_option("grammar", "plain")
    line = {"workload": a.workload, "reads": int(reads.shape[0]), "read_len": int(reads.shape[1]),
            "identical": res["plain"]["results_sha256"] == res["shaped"]["results_sha256"], **res}
    print(json.dumps(line))
    if a.out:
        with open(a.out, "w") as f:
            f.write(json.dumps(line) + "\n")
    return 0 if line["identical"] else 1


if __name__ == "__main__":
    sys.exit(main())
