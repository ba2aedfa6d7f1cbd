"""The drop-in executables at size (the host I/O pipeline): `moni_align_b200/bin/{ms,mems}` on a FASTQ of
the bench workload, wall and phase times with the stage busy times the CLI prints, and a byte comparison with the reference
binaries (oracle/_ref/bin) on a prefix of the same file.  Writes one JSON line.

    python tests/scale/cli_run.py [--workload cfg2s] [--reads 2000000] [--ref-reads 100000] [--threads 16]
"""
import argparse
import hashlib
import json
import os
import re
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import build_index  # noqa: E402
import synth  # noqa: E402
import workloads  # noqa: E402

BIN = os.path.join(ROOT, "moni_align_b200", "bin")


def sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for blk in iter(lambda: f.read(1 << 24), b""):
            h.update(blk)
    return h.hexdigest()


def run_cli(exe, index, fq, out, threads, extra=()):
    t0 = time.time()
    r = subprocess.run([os.path.join(BIN, exe), index, "-p", fq, "-o", out, "-t", str(threads), *extra],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    wall = time.time() - t0
    txt = r.stdout.decode(errors="replace")
    if r.returncode != 0:
        raise RuntimeError(txt[-2000:])
    el = [float(x) for x in re.findall(r"Elapsed time \(s\): ([0-9.eE+-]+)", txt)]
    stages = re.search(r"gpu-call time \(s\) ([0-9.eE+-]+) parse \(s\) ([0-9.eE+-]+) format\+write \(s\) ([0-9.eE+-]+)", txt)
    res = {"wall_s": round(wall, 3), "index_load_s": el[0], "processing_s": el[1]}
    if stages:
        res["busy_s"] = {"gpu_call": float(stages.group(1)), "parse": float(stages.group(2)), "format_write": float(stages.group(3))}
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg2s")
    ap.add_argument("--reads", type=int, default=2_000_000)
    ap.add_argument("--ref-reads", type=int, default=100_000)
    ap.add_argument("--threads", type=int, default=16)
    ap.add_argument("--tmp", default="/tmp/cli_run")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    os.makedirs(a.tmp, exist_ok=True)
    pre = workloads.ensure_index(a.workload)
    reads = workloads.reads_of(a.workload, pairs=a.reads // 2)
    fq, fq_ref = os.path.join(a.tmp, "reads.fq"), os.path.join(a.tmp, "reads_ref.fq")
    synth.write_fastq_matrix(fq, reads)
    synth.write_fastq_matrix(fq_ref, reads[:a.ref_reads])
    line = {"workload": a.workload, "reads": int(reads.shape[0]), "read_len": int(reads.shape[1]), "threads": a.threads,
            "fastq_bytes": os.path.getsize(fq)}
    for exe in ("ms", "mems"):
        out = os.path.join(a.tmp, "ours_" + exe)
        res = run_cli(exe, pre, fq, out, a.threads)
        exts = (".mems",) if exe == "mems" else (".pointers", ".lengths")
        res["text_bytes"] = sum(os.path.getsize(out + e) for e in exts)
        res["reads_per_s_processing"] = round(reads.shape[0] / res["processing_s"])
        line[exe] = res
        # the reference on the first ref_reads reads: our files for the same reads must be the same bytes
        if a.ref_reads:
            o2, r2 = os.path.join(a.tmp, "ours_ref_" + exe), os.path.join(a.tmp, "ref_" + exe)
            run_cli(exe, pre, fq_ref, o2, 3)
            t = build_index.run_reference(exe, pre, fq_ref, r2, threads=a.threads)
            res["reference"] = {"reads": a.ref_reads, "wall_s": round(t["wall"], 3), "processing_s": t["phase"],
                                "identical": all(sha(o2 + e) == sha(r2 + e) for e in exts)}
        if os.path.exists(os.path.join(BIN, exe + "_old")):      # the serial CLI of the previous commit, when built beside it
            h = {e: sha(out + e) for e in exts}
            old = run_cli(exe + "_old", pre, fq, out, a.threads)
            old["same_files"] = all(sha(out + e) == h[e] for e in exts)
            res["serial_cli_before"] = old
        for e in exts:
            os.remove(out + e)
    line["identical"] = all(line[e].get("reference", {}).get("identical", True) for e in ("ms", "mems"))
    print(json.dumps(line))
    if a.out:
        with open(a.out, "w") as f:
            f.write(json.dumps(line) + "\n")
    return 0 if line["identical"] else 1


if __name__ == "__main__":
    sys.exit(main())
