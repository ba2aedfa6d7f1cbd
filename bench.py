#!/usr/bin/env python
"""bench.py -- `moni mems` hot path on B200 (BASELINE.json metric: MS+MEM reads/s, bases/s, % of HBM roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg2|small]

One "step" = one pass of the hot path (MS pointers -> MS lengths -> MEM records) over one batch of synthetic
reads.  Default workload: BASELINE.json configs[1] stand-in (tools/workloads.py: 61.4 Mbp x 7 haplotypes,
1M x 2 x 150 bp paired-end reads per GPU).  N > 1 is launched by torchrun, one rank per GPU; the index is
replicated, every rank processes its own batch (weak scaling), no collective on the data path; torch.distributed
carries only the barrier and the max-over-ranks of the device times.

`value`   whole-job reads/s with the reads already resident in HBM (kernels only, CUDA events on the library's
          launching stream, summed over the K timed steps, max over ranks).
`e2e`     the same metric through the public C-ABI call moni_gpu_mems() with pinned HOST buffers: host->device
          copy of the reads and device->host copy of the MEM records inside the timed region.
`--impl reference` times the UNMODIFIED reference `mems` binary (oracle/_ref/bin, built from /root/reference) on the
          host cores with -t nproc, on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))


_REAL_STDOUT = None


def capture_stdout() -> None:
    """Send everything that writes to fd 1 (NCCL banners, child processes) to stderr, so that stdout carries exactly
    the one JSON line emit() prints."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def emit(line: dict) -> None:
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def log(*a):
    print("[bench]", *a, file=sys.stderr, flush=True)


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, streaming copy)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """Samples nvidia-smi SM clocks and throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu: int):
        super().__init__(daemon=True)
        self.gpu, self.stop_flag, self.rows = gpu, threading.Event(), []

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.1)

    def summary(self) -> dict:
        self.stop_flag.set()
        self.join(timeout=6)
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows for i in range(4) if len(r) > 2 + i and r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


# ------------------------------------------------------------------------------------------ reference arm
def write_sample_fastq(reads: np.ndarray, path: str) -> None:
    import synth
    synth.write_fastq_matrix(path, reads)


def run_reference_mems(prefix: str, fastq: str, threads: int) -> float:
    """Seconds of the reference's own 'Processing patterns' phase (src/mems.cpp dispatcher timer)."""
    import build_index
    with tempfile.TemporaryDirectory() as td:
        r = build_index.run_reference("mems", prefix, fastq, os.path.join(td, "out"), threads=threads)
    if r["phase"] is None:
        raise RuntimeError("could not parse the reference's phase timer:\n" + r["stdout"][-800:])
    return float(r["phase"])


def reference_arm(args) -> None:
    import workloads
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "bin", "mems")):
        emit({"impl": "reference", "unavailable": "oracle/_ref/bin/mems not built (needs /root/reference at build time)"})
        return
    w = workloads.WORKLOADS[args.workload]
    prefix = workloads.ensure_index(args.workload)
    cores = os.cpu_count() or 1
    pairs = max(1000, args.ref_reads // 2)
    reads = workloads.reads_of(args.workload, pairs=pairs)
    with tempfile.TemporaryDirectory() as td:
        fq = os.path.join(td, "sample.fq")
        write_sample_fastq(reads, fq)
        times = []
        for s in range(args.warmup + args.steps):
            t = run_reference_mems(prefix, fq, cores)
            log(f"reference step {s}: {t:.2f} s for {reads.shape[0]} reads on {cores} threads")
            if s >= args.warmup:
                times.append(t)
    total = sum(times)
    value = reads.shape[0] * len(times) / total
    sample = f"{reads.shape[0]} reads ({pairs} pairs x 2 x {w['read_len']} bp) of the same workload per step"
    line = {"metric": "mems_reads_per_s", "value": value, "unit": "reads/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u64", "data": "synthetic", "impl": "reference",
            "bases_per_s": value * w["read_len"],
            "config": {"workload": w["desc"], "reads_per_step": int(reads.shape[0]), "read_len": w["read_len"],
                       "timer": "reference's own 'Processing patterns' phase, index load excluded"},
            "cpu_baseline": {"value": value, "unit": "reads/s", "cores": cores, "kind": "reference", "sample": sample},
            "e2e": {"value": value, "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


_ORIG_AFFINITY: list = []


def bind_near_gpu(local: int) -> str:
    """Pins this rank's host threads to the CPUs NVML reports as local to its GPU, before any pinned buffer is allocated, so
    that the page-locked batch buffers (first touched here) and the PCIe copies stay on the GPU's NUMA node.  Returns a note."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64 + 16)
        near = {64 * w + b for w, x in enumerate(words) for b in range(64) if (x >> b) & 1}
        allowed = os.sched_getaffinity(0)
        _ORIG_AFFINITY.append(allowed)
        use = near & allowed
        if not use or use == allowed:
            return f"no change ({len(near)} local CPUs, {len(allowed)} allowed)"
        os.sched_setaffinity(0, use)
        return f"bound to {len(use)} of {len(allowed)} allowed CPUs"
    except Exception as e:  # best effort: never fail the bench over placement
        return f"skipped: {type(e).__name__}: {e}"


# ------------------------------------------------------------------------------------------ our arm
def ours_arm(args) -> None:
    import torch
    import torch.distributed as dist
    import moni_align_b200 as m
    import workloads
    from moni_align_b200.sharding import max_over_ranks, sum_over_ranks

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa_note = "off" if args.no_numa_bind else bind_near_gpu(local)
    log(f"rank {rank}: CPU placement: {numa_note}")
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    w = workloads.WORKLOADS[args.workload]
    if rank == 0:
        prefix = workloads.ensure_index(args.workload)
    if world > 1:
        dist.barrier()
    prefix = workloads.index_prefix(args.workload)

    t0 = time.time()
    if args.verbose:
        m.set_option("verbose", 1)
    for kv in args.opt:
        k, _, val = kv.partition("=")
        m.set_option(k, val)
    ix = m.MoniIndex(prefix, gpus=[local])
    t_load = time.time() - t0
    log(f"rank {rank}: index n={ix.n} r={ix.r} image {ix.device_bytes / 1e9:.2f} GB loaded in {t_load:.1f} s")

    text = workloads.text_of(args.workload)
    reads = workloads.reads_of(args.workload, text=text, pairs=args.pairs or w["pairs"], seed_shift=1000 * rank)
    del text
    n_reads, read_len = reads.shape
    total = n_reads * read_len
    # pinned host buffers for the end-to-end leg
    h_seqs = m.pinned_empty(total, np.uint8, gpu=local)
    h_seqs[:] = reads.reshape(-1)
    h_offs = m.pinned_empty(n_reads + 1, np.uint64, gpu=local)
    h_offs[:] = np.arange(n_reads + 1, dtype=np.uint64) * np.uint64(read_len)
    h_mo = m.pinned_empty(n_reads + 1, np.uint64, gpu=local)
    h_mp = m.pinned_empty(total, np.uint64, gpu=local)
    h_ml = m.pinned_empty(total, np.uint64, gpu=local)
    del reads

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- event counters for the roofline numerator (one counted run, outside the timed region)
    rb = ix.resident_batch(h_seqs, h_offs)
    m.set_option("count", 1)
    ix.counters(reset=True)
    rb.run("mems")
    cnt = ix.counters(reset=True)
    m.set_option("count", 0)

    # ---- kernel-only leg (inputs resident in HBM)
    for _ in range(args.warmup):
        rb.run("mems")
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    launches0 = m.launch_count()
    k_ms = np.zeros(3)
    for _ in range(args.steps):
        t = rb.run("mems")
        k_ms += np.array([t["pass1_ms"], t["pass2_ms"], t["compact_ms"]])
    barrier()
    launches = m.launch_count() - launches0
    dev_ms = max_over_ranks(float(k_ms.sum()), dev)
    mo, mp, ml = rb.fetch_mems()
    n_mems = int(mo[-1])
    rb.close()

    # ---- end-to-end leg through the C ABI with host buffers: moni_gpu_mems32 (32-bit record fields) is the headline,
    # moni_gpu_mems (64-bit fields, twice the device->host bytes) is reported beside it
    h_mp32 = m.pinned_empty(total, np.uint32, gpu=local)
    h_ml32 = m.pinned_empty(total, np.uint32, gpu=local)

    def e2e_leg(out):
        call = (ix.maxrimal_exact_matches if len(out) == 3 else
                ix.maxrimal_exact_matches_packed16 if out[1].dtype == np.uint16 else ix.maxrimal_exact_matches_packed)
        for _ in range(args.warmup):
            call(h_seqs, h_offs, out=out)
        barrier()
        dev_sum, stage = 0.0, np.zeros(6)
        t_wall = time.perf_counter()
        for _ in range(args.steps):
            call(h_seqs, h_offs, out=out)
            tt = ix.timing()
            dev_sum += tt["total_ms"]
            stage += np.array([tt[k] for k in ("h2d_ms", "pass1_ms", "pass2_ms", "compact_ms", "d2h_ms", "total_ms")])
        barrier()
        wall = (time.perf_counter() - t_wall) * 1e3
        return max_over_ranks(dev_sum, dev), max_over_ranks(wall, dev), stage

    e2e64_dev_ms, e2e64_wall_ms, stage64 = e2e_leg((h_mo, h_mp, h_ml))
    same64 = bool(np.array_equal(h_mo, mo) and np.array_equal(h_mp[:n_mems], mp) and np.array_equal(h_ml[:n_mems], ml))
    h_mo[:] = 0
    e2e32_dev_ms, e2e32_wall_ms, stage32 = e2e_leg((h_mo, h_mp32, h_ml32))
    same32 = bool(np.array_equal(h_mo, mo) and np.array_equal(h_mp32[:n_mems], mp) and np.array_equal(h_ml32[:n_mems], ml))
    # one packed 32-bit word per record (reads are shorter than 65536 bases)
    h_mo[:] = 0
    h_rec = h_mp32
    e2e_dev_ms, e2e_wall_ms, stage = e2e_leg((h_mo, h_rec))
    same = bool(same64 and same32 and np.array_equal(h_mo, mo) and np.array_equal(h_rec[:n_mems] >> 16, mp)
                and np.array_equal(h_rec[:n_mems] & 0xFFFF, ml))
    e2e_api, rec_bytes, e2e_p32 = "moni_gpu_mems_packed", 4, None
    if int(read_len) < 256:
        # headline for short reads: one packed 16-bit word per record (query offset << 8 | length), half the bytes again
        e2e_p32 = (e2e_dev_ms, e2e_wall_ms)
        h_mo[:] = 0
        h_rec16 = h_ml32.view(np.uint16)
        e2e_dev_ms, e2e_wall_ms, stage = e2e_leg((h_mo, h_rec16))
        same = bool(same and np.array_equal(h_mo, mo) and np.array_equal(h_rec16[:n_mems] >> 8, mp) and np.array_equal(h_rec16[:n_mems] & 0xFF, ml))
        e2e_api, rec_bytes = "moni_gpu_mems_packed16", 2
    clocks = sampler.summary() if sampler else None
    all_reads = sum_over_ranks(float(n_reads), dev)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    value = all_reads * args.steps / (dev_ms / 1e3)
    e2e_value = all_reads * args.steps / (e2e_wall_ms / 1e3)
    # roofline of the dominant kernel: B200 fetches one 128-byte DRAM line per random access, so
    #   pass 1 = 137 B per base (row line 128 + pointer store 8 + read byte 1)
    #   pass 2 = 128 B per LCE restart that touches text + 16 B per grammar node visited + 9 B per base + 8 B per MEM record
    # SURVEY.md 8(d)'s layout-independent model is reported beside it.
    alg = {"pass1": 137.0 * cnt["bases"],
           "pass2": 128.0 * cnt["descents"] + 16.0 * cnt["restart_depth"] + 9.0 * cnt["bases"] + 8.0 * n_mems}
    survey = {"pass1": 32.0 * (5 * cnt["bases"] + 6 * cnt["jumps"]), "pass2": 32.0 * (cnt["restart_depth"] + 2 * cnt["chars"])}
    per = {"pass1": k_ms[0] / args.steps, "pass2": k_ms[1] / args.steps}
    dom = "pass1" if per["pass1"] >= per["pass2"] else "pass2"
    peak, peak_src = measured_peak_gbs()
    achieved = alg[dom] / (per[dom] / 1e3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            traffic = json.load(open(tp)).get(dom)
        except Exception:
            traffic = None
    cpu = None
    if not args.no_cpu_baseline and world == 1 and os.path.exists(os.path.join(ROOT, "oracle", "_ref", "bin", "mems")):
        try:
            if _ORIG_AFFINITY:      # the reference gets every host core the process was given, not only the GPU's neighbours
                os.sched_setaffinity(0, _ORIG_AFFINITY[0])
            pairs = max(1000, args.ref_reads // 2)
            sample = workloads.reads_of(args.workload, pairs=pairs)
            cores = os.cpu_count() or 1
            with tempfile.TemporaryDirectory() as td:
                fq = os.path.join(td, "sample.fq")
                write_sample_fastq(sample, fq)
                sec = run_reference_mems(prefix, fq, cores)
            cpu = {"value": sample.shape[0] / sec, "unit": "reads/s", "cores": cores, "kind": "reference",
                   "sample": f"{sample.shape[0]} reads of the same workload, reference `mems -t {cores}`, its own processing-phase timer ({sec:.2f} s)"}
        except Exception as e:  # the baseline is reported, never required
            cpu = {"value": None, "unit": "reads/s", "cores": os.cpu_count(), "kind": "reference", "sample": f"failed: {e}"}
    line = {
        "metric": "mems_reads_per_s", "value": value, "unit": "reads/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
        "data": "synthetic", "bases_per_s": value * read_len,
        "config": {"workload": w["desc"], "reads_per_gpu_per_step": int(n_reads), "read_len": int(read_len), "index_n": ix.n,
                   "index_runs": ix.r, "hbm_image_bytes": ix.device_bytes, "l2": "inputs larger than L2 (index image and read batch)",
                   "parallelism": f"reads sharded x{world}, index replicated, no collective"},
        "stage_ms_per_step": {"pass1": per["pass1"], "pass2": per["pass2"], "compact": k_ms[2] / args.steps},
        "roofline": {"bound": "hbm", "kernel": "k_ms_pointers" if dom == "pass1" else "k_ms_lengths", "achieved": achieved,
                     "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": alg[dom], "algorithmic_bytes_per_base": alg[dom] / max(cnt["bases"], 1),
                     "survey_model_bytes": survey[dom],
                     "other_kernel": {"kernel": "k_ms_lengths" if dom == "pass1" else "k_ms_pointers",
                                      "achieved": alg["pass2" if dom == "pass1" else "pass1"] / (per["pass2" if dom == "pass1" else "pass1"] / 1e3) / 1e9}},
        "counters_per_step": cnt,
        "e2e": {"value": e2e_value, "unit": "reads/s", "api": e2e_api, "h2d_bytes_per_step": int(total + 8 * (n_reads + 1)),
                "d2h_bytes_per_step": int(8 * (n_reads + 1) + rec_bytes * n_mems), "ms_per_step_wall": e2e_wall_ms / args.steps,
                "ms_per_step_device": e2e_dev_ms / args.steps, "same_bits_as_resident_run": same,
                "stage_busy_ms_per_step": dict(zip(["h2d", "pass1", "pass2", "compact", "d2h", "total"], (stage / args.steps).tolist()))},
        **({"e2e_packed32": {"value": all_reads * args.steps / (e2e_p32[1] / 1e3), "unit": "reads/s", "api": "moni_gpu_mems_packed",
                             "d2h_bytes_per_step": int(8 * (n_reads + 1) + 4 * n_mems), "ms_per_step_wall": e2e_p32[1] / args.steps,
                             "ms_per_step_device": e2e_p32[0] / args.steps}} if e2e_p32 else {}),
        "e2e_u32": {"value": all_reads * args.steps / (e2e32_wall_ms / 1e3), "unit": "reads/s", "api": "moni_gpu_mems32",
                    "d2h_bytes_per_step": int(8 * (n_reads + 1) + 8 * n_mems), "ms_per_step_wall": e2e32_wall_ms / args.steps,
                    "stage_busy_ms_per_step": dict(zip(["h2d", "pass1", "pass2", "compact", "d2h", "total"], (stage32 / args.steps).tolist()))},
        "e2e_u64": {"value": all_reads * args.steps / (e2e64_wall_ms / 1e3), "unit": "reads/s", "api": "moni_gpu_mems",
                    "d2h_bytes_per_step": int(8 * (n_reads + 1) + 16 * n_mems), "ms_per_step_wall": e2e64_wall_ms / args.steps,
                    "stage_busy_ms_per_step": dict(zip(["h2d", "pass1", "pass2", "compact", "d2h", "total"], (stage64 / args.steps).tolist()))},
        "mem_records_per_step": n_mems,
        "gpu_launches": int(launches), "options": args.opt, "cpu_placement": numa_note, "host_buffers": f"pinned, NUMA node of the GPU ({m.load_library().moni_gpu_numa_node(local)})", "clocks": clocks, "cpu_baseline": cpu, "index_load_s": t_load,
    }
    emit(line)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg2s", choices=["cfg2", "cfg2s", "small"])
    ap.add_argument("--pairs", type=int, default=0, help="read pairs per GPU per step (default: the workload's)")
    ap.add_argument("--ref-reads", type=int, default=100_000, help="reads in the bounded CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-numa-bind", action="store_true", help="do not pin each rank to the CPUs local to its GPU")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--opt", action="append", default=[], help="library option key=value (moni_gpu_set_option), repeatable")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    capture_stdout()
    if args.impl == "reference":
        reference_arm(args)
    else:
        ours_arm(args)


if __name__ == "__main__":
    main()
