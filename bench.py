#!/usr/bin/env python
"""bench.py -- `moni ms` / `moni mems` hot path on B200 (BASELINE.json metric: MS+MEM reads/s, bases/s, % of the HBM
random-access roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg3|cfg2s|cfg1|cfg4s|small]

One "step" = one pass of the hot path over one batch of synthetic reads: MS pointers -> MS lengths -> MEM records (`mems`,
the headline `value`), and the same two passes with full pointer/length arrays as the result (`ms`, reported under "ms").
Default workload at N = 1: BASELINE.json configs[2], the largest single-GPU configuration (tools/workloads.py "cfg3":
40 Mbp x 32 haplotypes, n = 1.28 G; 2 M reads per step = the first 1 M pairs of the 10 M-read set).  N > 1 is launched by
torchrun, one rank per GPU: the index is replicated, every rank processes its own batch (weak scaling), no collective on the
data path; torch.distributed carries only the barrier and the max-over-ranks of the device times.

`value`        whole-job reads/s with the reads already resident in HBM (kernels only, CUDA events on the library's launching
               stream, summed over the K timed steps, max over ranks).
`e2e`          the same metric through the public C-ABI call with pinned HOST buffers: host->device copy of the reads and
               device->host copy of the results inside the timed region.  `e2e` is the call that returns the reference's own
               record type (u64 pairs, moni_gpu_mems); the packed-record variants are reported beside it.
`parity_sample` the reference `ms` / `mems` binaries run on a sample of this very batch (the `cpu_baseline` leg) and their
               output files are compared number by number with what the GPU returned for those reads.
`roofline`     dominant kernel against the streaming HBM peak (MEASURED_PEAKS.json) and against the random-line rate of the
               box measured in this process (moni_gpu_probe_random_lines).
`--impl reference` times the UNMODIFIED reference `mems` binary (oracle/_ref/bin, built from /root/reference) on the host
               cores with -t nproc, on a bounded sample of the same workload.
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

DEFAULT_WORKLOAD = "cfg3"
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
    """Samples SM clocks and throttle reasons during the timed region: NVML in-process (a query costs microseconds and takes no
    driver-wide lock the way a spawned nvidia-smi does), nvidia-smi only if NVML cannot be loaded."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu: int):
        super().__init__(daemon=True)
        self.gpu, self.stop_flag, self.rows = gpu, threading.Event(), []
        self.paused = False
        self.nv = self.h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(gpu)
            self.max_sm = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _nvml_row(self):
        nv = self.nv
        sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
        mx = self.max_sm
        try:
            r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
        except Exception:
            r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        bits = [getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8), getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20), getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4)]
        return [str(sm), str(mx)] + ["Active" if r & b else "Not Active" for b in bits]

    def sample_once(self):
        try:
            if self.nv is not None:
                self.rows.append(self._nvml_row())
            else:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
        except Exception:
            pass

    def run(self):
        while not self.stop_flag.is_set():
            if not self.paused:
                self.sample_once()
            self.stop_flag.wait(0.25)

    def summary(self) -> dict:
        self.stop_flag.set()
        self.join(timeout=6)
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows for i in range(4) if len(r) > 2 + i and r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows), "source": "nvml" if self.nv is not None else "nvidia-smi"}


# ------------------------------------------------------------------------------------------ reference arm
def run_reference(cmd: str, prefix: str, fastq: str, threads: int, out_prefix: str) -> float:
    """Seconds of the reference's own 'Processing patterns' phase (src/matching_statistics.cpp:500-524, src/mems.cpp dispatcher)."""
    import build_index
    r = build_index.run_reference(cmd, prefix, fastq, out_prefix, threads=threads)
    if r["phase"] is None:
        raise RuntimeError("could not parse the reference's phase timer:\n" + r["stdout"][-800:])
    return float(r["phase"])


def sample_rows(n_reads: int, k: int) -> np.ndarray:
    """Rows of a [R1 block | R2 block] batch that make a k-read sample: the first k/2 pairs, both mates."""
    half = n_reads // 2
    k2 = max(1, min(k // 2, half))
    return np.concatenate([np.arange(k2), half + np.arange(k2)])


def bench_pairs(w: dict, args) -> int:
    return args.pairs or w.get("bench_pairs", w["pairs"])


def reference_arm(args) -> None:
    import synth
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
    pairs = max(500, args.ref_reads // 2)
    reads = workloads.reads_of(args.workload, pairs=pairs)
    with tempfile.TemporaryDirectory() as td:
        fq = os.path.join(td, "sample.fq")
        synth.write_fastq_matrix(fq, reads)
        times = []
        for s in range(args.warmup + args.steps):
            t = run_reference("mems", prefix, fq, cores, os.path.join(td, "out"))
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


# ------------------------------------------------------------------------------------------ checker: reference on a sample
def cpu_baseline_and_parity(args, prefix, reads_flat, n_reads, read_len, gpu_mems, gpu_ms):
    """Runs the unmodified reference `mems` and `ms` on a sample of the batch (all host cores) and compares their output files
    with the GPU results for the same reads.  gpu_mems = (mem_offsets, mem_pos, mem_len) of the whole batch, gpu_ms = (pointers,
    lengths).  Returns (cpu_baseline, parity_sample)."""
    import refparse
    import synth
    if _ORIG_AFFINITY:      # the reference gets every host core the process was given, not only the GPU's neighbours
        os.sched_setaffinity(0, _ORIG_AFFINITY[0])
    cores = os.cpu_count() or 1
    rows = sample_rows(n_reads, args.ref_reads)
    sample = reads_flat.reshape(n_reads, read_len)[rows]
    par = {"reads": int(rows.shape[0]), "against": "output files of the unmodified reference binaries (oracle/_ref/bin/{mems,ms})"}
    with tempfile.TemporaryDirectory() as td:
        fq = os.path.join(td, "sample.fq")
        synth.write_fastq_matrix(fq, sample)
        sec = run_reference("mems", prefix, fq, cores, os.path.join(td, "ref"))
        cpu = {"value": rows.shape[0] / sec, "unit": "reads/s", "cores": cores, "kind": "reference",
               "sample": f"{rows.shape[0]} reads of this batch, reference `mems -t {cores}`, its own processing-phase timer ({sec:.2f} s)"}
        _, cnt, qoff, qlen = refparse.parse_mems(os.path.join(td, "ref.mems"))
        mo, mp, ml = gpu_mems
        a, b = mo[rows].astype(np.int64), mo[rows + 1].astype(np.int64)
        same_counts = bool(np.array_equal(cnt.astype(np.int64), b - a))
        if same_counts:
            idx = np.concatenate([np.arange(x, y) for x, y in zip(a.tolist(), b.tolist())]) if rows.shape[0] else np.zeros(0, np.int64)
            par["mems_identical"] = bool(np.array_equal(mp[idx], qoff) and np.array_equal(ml[idx], qlen))
        else:
            par["mems_identical"] = False
        par["mems_records"] = int(cnt.sum())
        if gpu_ms is not None:
            sec_ms = run_reference("ms", prefix, fq, cores, os.path.join(td, "refms"))
            cpu["ms"] = {"value": rows.shape[0] / sec_ms, "unit": "reads/s", "sample": f"same reads, reference `ms -t {cores}` ({sec_ms:.2f} s)"}
            _, _, pv = refparse.parse_values(os.path.join(td, "refms.pointers"))
            _, _, lv = refparse.parse_values(os.path.join(td, "refms.lengths"))
            ptr, ln = gpu_ms
            g_ptr = ptr.reshape(n_reads, read_len)[rows].reshape(-1)
            g_len = ln.reshape(n_reads, read_len)[rows].reshape(-1)
            par["pointers_identical"] = bool(np.array_equal(g_ptr, pv))
            par["lengths_identical"] = bool(np.array_equal(g_len, lv))
    par["identical"] = bool(par.get("mems_identical") and par.get("pointers_identical", True) and par.get("lengths_identical", True))
    return cpu, par


API_NAMES = {"mems_u64": "moni_gpu_mems", "mems_u64_link": "moni_gpu_mems", "mems_u32": "moni_gpu_mems32", "mems_packed32": "moni_gpu_mems_packed",
             "mems_packed16": "moni_gpu_mems_packed16", "ms_u64": "moni_gpu_ms", "ms_u64_link": "moni_gpu_ms", "ms_packed": "moni_gpu_ms_packed",
             "ms_runs": "moni_gpu_ms_runs", "ms_runs_packed": "moni_gpu_ms_runs_packed"}


def api_name(leg: str) -> str:
    return API_NAMES[leg]


def random_line_peak(m, local: int) -> dict:
    """The random-access roofline of this GPU, measured now (dependent random 32-byte accesses over a 3 GiB table)."""
    out = {"table_bytes": 3 << 30, "unit": "lines/s", "line_bytes": 128, "points": []}
    best = 0.0
    for shape, k, b in ((0, 1, 6), (0, 2, 6), (0, 4, 4), (2, 8, 8)):
        try:
            r = m.probe_random_lines(local, 3 << 30, shape, k, b, 384)
            out["points"].append({"shape": r["shape"], "chains_per_sm": r["chains_per_sm"], "lines_per_s": r["lines_per_s"]})
            best = max(best, r["lines_per_s"])
        except Exception as e:
            out["points"].append({"shape": shape, "chains_per_thread": k, "error": str(e)})
    m.probe_release()
    out["value"] = best
    out["as_line_traffic_gb_per_s"] = best * 128 / 1e9
    return out


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
    cpu_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        # a CPU-side barrier for the one leg where rank 0 drives every GPU itself: ranks waiting in an NCCL barrier would keep a
        # polling kernel on their GPUs
        cpu_group = dist.new_group(backend="gloo")
    w = workloads.WORKLOADS[args.workload]
    if rank == 0:
        workloads.ensure_index(args.workload)
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

    t0 = time.time()
    text = workloads.text_of(args.workload)
    reads = workloads.reads_of(args.workload, text=text, pairs=bench_pairs(w, args), seed_shift=1000 * rank)
    del text
    n_reads, read_len = reads.shape
    total = n_reads * read_len
    log(f"rank {rank}: {n_reads} reads of {read_len} bases prepared in {time.time() - t0:.1f} s")
    # pinned host buffers for the end-to-end legs: two result pools, every leg uses views of them
    h_seqs = m.pinned_empty(total, np.uint8, gpu=local)
    h_seqs[:] = reads.reshape(-1)
    del reads
    h_offs = m.pinned_empty(n_reads + 1, np.uint64, gpu=local)
    h_offs[:] = np.arange(n_reads + 1, dtype=np.uint64) * np.uint64(read_len)
    h_mo = m.pinned_empty(n_reads + 1, np.uint64, gpu=local)
    pool_a = m.pinned_empty(total, np.uint64, gpu=local)
    pool_b = m.pinned_empty(total, np.uint64, gpu=local)

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

    # ---- kernel-only legs (inputs resident in HBM): mems is the headline, ms beside it
    def resident_leg(what):
        for _ in range(args.warmup):
            rb.run(what)
        barrier()
        l0 = m.launch_count()
        k_ms = np.zeros(3)
        for _ in range(args.steps):
            t = rb.run(what)
            k_ms += np.array([t["pass1_ms"], t["pass2_ms"], t["compact_ms"]])
        barrier()
        return k_ms, m.launch_count() - l0

    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    k_ms, launches = resident_leg("mems")
    dev_ms = max_over_ranks(float(k_ms.sum()), dev)
    mo, mp, ml = rb.fetch_mems()
    n_mems = int(mo[-1])
    kms_ms, launches_ms = resident_leg("ms")
    dev_ms_ms = max_over_ranks(float(kms_ms.sum()), dev)
    r_ptr, r_len = rb.fetch_ms()
    rb.close()

    # ---- end-to-end legs through the C ABI with pinned host buffers
    def e2e_leg(call, out):
        for _ in range(args.warmup):
            call(h_seqs, h_offs, out)
        barrier()
        dev_sum, stage, per_step = 0.0, np.zeros(6), []
        t_wall = time.perf_counter()
        for _ in range(args.steps):
            t_s = time.perf_counter()
            call(h_seqs, h_offs, out)
            per_step.append((time.perf_counter() - t_s) * 1e3)
            tt = ix.timing()
            dev_sum += tt["total_ms"]
            stage += np.array([tt[k] for k in ("h2d_ms", "pass1_ms", "pass2_ms", "compact_ms", "d2h_ms", "total_ms")])
        barrier()
        wall = (time.perf_counter() - t_wall) * 1e3
        log(f"rank {rank}: {getattr(call, '__name__', 'call')} per-step wall ms: " + " ".join(f"{x:.1f}" for x in per_step))
        return {"dev": max_over_ranks(dev_sum, dev), "wall": max_over_ranks(wall, dev), "stage": stage, "per_step": per_step}

    # host threads the library uses to widen packed results into u64 columns (moni_gpu_mems, moni_gpu_ms)
    allowed_cpus = len(os.sched_getaffinity(0))
    widen_threads = next((int(o.split("=", 1)[1]) for o in args.opt if o.startswith("widen_threads=")), -1)
    if widen_threads < 0:
        widen_threads = 0 if allowed_cpus < 4 else min(16, allowed_cpus * 2 // 3)     # WidenPool::default_threads (csrc/host_widen.cpp)
        if world > 1:
            # the ranks of one box share its CPUs and the library cannot see its sibling processes: give each rank its share
            widen_threads = max(1, min(8, (allowed_cpus - 2 * world) // world)) if allowed_cpus // world >= 2 else 0     # one CPU per rank stays with its caller thread
            m.set_option("widen_threads", widen_threads)
    legs = {}
    # mems, u64 records: the reference's own record type (std::pair<size_t,size_t>, src/mems.cpp:277-279)
    legs["mems_u64"] = e2e_leg(ix.maxrimal_exact_matches, (h_mo, pool_a, pool_b))
    same = {"mems_u64": bool(np.array_equal(h_mo, mo) and np.array_equal(pool_a[:n_mems], mp) and np.array_equal(pool_b[:n_mems], ml))}
    gpu_mems = (h_mo.copy(), pool_a[:n_mems].copy(), pool_b[:n_mems].copy())      # for the parity sample: what the ABI call returned
    # A/B of the same call with the u64 columns crossing the link as they are (option widen_threads = 0) instead of packed + widened by host threads
    if widen_threads >= 2:
        h_mo[:] = 0
        m.set_option("widen_threads", 0)
        try:
            legs["mems_u64_link"] = e2e_leg(ix.maxrimal_exact_matches, (h_mo, pool_a, pool_b))
        finally:
            m.set_option("widen_threads", widen_threads)
        same["mems_u64_link"] = bool(np.array_equal(h_mo, mo) and np.array_equal(pool_a[:n_mems], mp) and np.array_equal(pool_b[:n_mems], ml))
    a32, b32 = pool_a.view(np.uint32)[:total], pool_b.view(np.uint32)[:total]
    h_mo[:] = 0
    legs["mems_u32"] = e2e_leg(ix.maxrimal_exact_matches, (h_mo, a32, b32))
    same["mems_u32"] = bool(np.array_equal(h_mo, mo) and np.array_equal(a32[:n_mems], mp) and np.array_equal(b32[:n_mems], ml))
    h_mo[:] = 0
    legs["mems_packed32"] = e2e_leg(ix.maxrimal_exact_matches_packed, (h_mo, a32))
    same["mems_packed32"] = bool(np.array_equal(h_mo, mo) and np.array_equal(a32[:n_mems] >> 16, mp) and np.array_equal(a32[:n_mems] & 0xFFFF, ml))
    if int(read_len) < 256:
        h_mo[:] = 0
        a16 = pool_a.view(np.uint16)[:total]
        legs["mems_packed16"] = e2e_leg(ix.maxrimal_exact_matches_packed16, (h_mo, a16))
        same["mems_packed16"] = bool(np.array_equal(h_mo, mo) and np.array_equal(a16[:n_mems] >> 8, mp) and np.array_equal(a16[:n_mems] & 0xFF, ml))
    # ms, u64 pointers and lengths (the reference's std::vector<size_t> pair), then the 7-byte-per-base form
    legs["ms_u64"] = e2e_leg(ix.matching_statistics, (pool_a, pool_b))
    same["ms_u64"] = bool(np.array_equal(pool_a[:total], r_ptr) and np.array_equal(pool_b[:total], r_len))
    gpu_ms = (pool_a[:total].copy(), pool_b[:total].copy()) if (rank == 0 and world == 1 and not args.no_cpu_baseline) else None
    if widen_threads >= 6:
        m.set_option("widen_threads", 0)
        try:
            legs["ms_u64_link"] = e2e_leg(ix.matching_statistics, (pool_a, pool_b))
        finally:
            m.set_option("widen_threads", widen_threads)
        same["ms_u64_link"] = bool(np.array_equal(pool_a[:total], r_ptr) and np.array_equal(pool_b[:total], r_len))
    lo, hi, l16 = pool_a.view(np.uint32)[:total], pool_b.view(np.uint8)[:total], pool_b.view(np.uint16)[total:2 * total]
    legs["ms_packed"] = e2e_leg(ix.matching_statistics_packed, (lo, hi, l16))
    same["ms_packed"] = bool(np.array_equal(m.unpack_pointers(lo, hi), r_ptr) and np.array_equal(l16, r_len))
    # ms as restart runs (moni_gpu_ms_runs): only the offsets where the reference's loop does work leave the device
    q32, n32 = pool_a.view(np.uint32)[:total], pool_a.view(np.uint32)[total:2 * total]
    legs["ms_runs"] = e2e_leg(ix.matching_statistics_runs, (h_mo, q32, n32, pool_b))
    n_runs = int(h_mo[n_reads])
    xp, xl = m.expand_runs(h_offs, h_mo, q32[:n_runs], n32[:n_runs], pool_b[:n_runs])
    same["ms_runs"] = bool(np.array_equal(xp, r_ptr) and np.array_equal(xl, r_len))
    del xp, xl
    rec32, lo32, hi8 = pool_a.view(np.uint32)[:total], pool_a.view(np.uint32)[total:2 * total], pool_b.view(np.uint8)[:total]
    h_mo[:] = 0
    legs["ms_runs_packed"] = e2e_leg(lambda sq, of, out: ix.matching_statistics_runs(sq, of, out=out, packed=True), (h_mo, rec32, lo32, hi8))
    uq, ul, up = m.unpack_runs(rec32[:n_runs], lo32[:n_runs], hi8[:n_runs])
    xp, xl = m.expand_runs(h_offs, h_mo, uq, ul, up)
    same["ms_runs_packed"] = bool(int(h_mo[n_reads]) == n_runs and np.array_equal(xp, r_ptr) and np.array_equal(xl, r_len))
    del xp, xl, uq, ul, up
    clocks = sampler.summary() if sampler else None
    all_reads = sum_over_ranks(float(n_reads), dev)
    n_mems_all = sum_over_ranks(float(n_mems), dev)

    # ---- what the host side of the box can move: every rank copies the byte counts of the packed16 leg, both directions at once
    head_leg = "mems_packed16" if "mems_packed16" in legs else "mems_packed32"
    rec_b = 2 if head_leg == "mems_packed16" else 4
    dma = None
    if world > 1:
        hb, db = int(total), int(8 * (n_reads + 1) + rec_b * n_mems)
        hin, hout = pool_a.view(np.uint8)[:hb], pool_b.view(np.uint8)[:db]
        t_in, t_out = torch.from_numpy(hin), torch.from_numpy(hout)
        d_in = torch.empty(hb, dtype=torch.uint8, device=dev)
        d_out = torch.empty(db, dtype=torch.uint8, device=dev)
        s1, s2 = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)

        def both():
            with torch.cuda.stream(s1):
                d_in.copy_(t_in, non_blocking=True)
            with torch.cuda.stream(s2):
                t_out.copy_(d_out, non_blocking=True)
        for _ in range(2):
            both()
        barrier()
        t_w = time.perf_counter()
        for _ in range(args.steps):
            both()
        barrier()
        dt = max_over_ranks(time.perf_counter() - t_w, dev)
        dma = {"bytes_per_gpu_per_step": hb + db, "aggregate_gb_per_s": world * (hb + db) * args.steps / dt / 1e9,
               "ms_per_step": 1e3 * dt / args.steps, "what": "cudaMemcpyAsync of the packed16 leg's bytes, pinned host memory, host->device and device->host at once on every rank, no kernels"}
        del d_in, d_out

    # ---- the in-library multi-GPU path (north star): ONE host batch of world x n_reads reads through one call on an index that
    # holds a replica on every GPU; the library splits it by bases and lands results in input order.  Rank 0 drives it alone.
    sharded = None
    if world > 1 and not args.no_sharded:
        if rank == 0:
            try:
                ixs = m.MoniIndex(prefix, gpus=list(range(world)))
                big_seqs = m.pinned_empty(total * world, np.uint8, gpu=local)
                for g in range(world):
                    big_seqs[g * total:(g + 1) * total] = h_seqs
                big_offs = m.pinned_empty(n_reads * world + 1, np.uint64, gpu=local)
                big_offs[:] = np.arange(n_reads * world + 1, dtype=np.uint64) * np.uint64(read_len)
                big_mo = m.pinned_empty(n_reads * world + 1, np.uint64, gpu=local)
                big_rec = m.pinned_empty(n_mems * world + 16, np.uint16 if rec_b == 2 else np.uint32, gpu=local)
                call = ixs.maxrimal_exact_matches_packed16 if rec_b == 2 else ixs.maxrimal_exact_matches_packed
                for _ in range(args.warmup):
                    call(big_seqs, big_offs, (big_mo, big_rec))
                t_w = time.perf_counter()
                for _ in range(args.steps):
                    call(big_seqs, big_offs, (big_mo, big_rec))
                dt = time.perf_counter() - t_w
                shift = 8 if rec_b == 2 else 16
                ok = int(big_mo[-1]) == n_mems * world
                for g in range(world):       # every copy of the batch must come back as the resident run's records, in input order
                    seg = big_rec[g * n_mems:(g + 1) * n_mems]
                    ok = ok and bool(np.array_equal(big_mo[g * n_reads:(g + 1) * n_reads + 1] - np.uint64(g * n_mems), mo)
                                     and np.array_equal(seg >> shift, mp) and np.array_equal(seg & ((1 << shift) - 1), ml))
                sharded = {"value": n_reads * world * args.steps / dt, "unit": "reads/s", "api": api_name(head_leg), "gpus_in_one_call": world,
                           "reads_per_call": int(n_reads * world), "ms_per_call_wall": 1e3 * dt / args.steps, "same_bits": bool(ok),
                           "h2d_bytes_per_call": int(world * total), "d2h_bytes_per_call": int(8 * (n_reads * world + 1) + rec_b * n_mems * world)}
                ixs.close()
            except Exception as e:
                import traceback
                traceback.print_exc()
                sharded = {"error": f"{type(e).__name__}: {e}"}
        dist.barrier(group=cpu_group)
    if rank != 0:
        if world > 1:
            dist.barrier(group=cpu_group)          # rank 0 finishes its line
            dist.destroy_process_group()
        return

    value = all_reads * args.steps / (dev_ms / 1e3)
    value_ms = all_reads * args.steps / (dev_ms_ms / 1e3)
    rate = lambda leg: all_reads * args.steps / (legs[leg]["wall"] / 1e3)
    stage_names = ["h2d", "pass1", "pass2", "compact", "d2h", "total"]
    link_rec = 16 if widen_threads < 2 else (2 if int(read_len) < 256 else 4 if int(read_len) < 65536 else 8)   # bytes per record that cross the link
    d2h = {"mems_u64": 8 * (n_reads + 1) + link_rec * n_mems, "mems_u64_link": 8 * (n_reads + 1) + 16 * n_mems, "mems_u32": 8 * (n_reads + 1) + 8 * n_mems,
           "mems_packed32": 8 * (n_reads + 1) + 4 * n_mems, "mems_packed16": 8 * (n_reads + 1) + 2 * n_mems,
           "ms_u64": (7 if widen_threads >= 6 and int(read_len) < 65536 else 16) * total, "ms_u64_link": 16 * total, "ms_packed": 7 * total, "ms_runs": 8 * (n_reads + 1) + 16 * n_runs,
           "ms_runs_packed": 8 * (n_reads + 1) + 9 * n_runs}
    api = API_NAMES
    h2d = int(total)            # reads of one length: the library computes their offsets on the device (k_uniform_offsets)
    if dma:
        dma["e2e_packed_aggregate_gb_per_s"] = world * (h2d + d2h[head_leg]) / (legs[head_leg]["wall"] / args.steps / 1e3) / 1e9
        dma["e2e_packed_fraction_of_ceiling"] = dma["e2e_packed_aggregate_gb_per_s"] / dma["aggregate_gb_per_s"]

    def leg_json(leg):
        return {"value": rate(leg), "unit": "reads/s", "api": api[leg], "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": int(d2h[leg]),
                "ms_per_step_wall": legs[leg]["wall"] / args.steps, "ms_per_step_device": legs[leg]["dev"] / args.steps,
                "same_bits_as_resident_run": same[leg], "ms_per_step_wall_min_median_max": [min(legs[leg]["per_step"]), float(np.median(legs[leg]["per_step"])), max(legs[leg]["per_step"])],
                "stage_busy_ms_per_step": dict(zip(stage_names, (legs[leg]["stage"] / args.steps).tolist()))}

    # roofline of the dominant kernel: B200 fetches one 128-byte DRAM line per random access, so
    #   pass 1 = 137 B per base (row line 128 + pointer store 8 + read byte 1)
    #   pass 2 = 128 B per LCE restart that touches text + 16 B per grammar node visited + 9 B per base + 8 B per MEM record
    # SURVEY.md 8(d)'s layout-independent model is reported beside it.
    alg = {"pass1": 137.0 * cnt["bases"],
           "pass2": 128.0 * cnt["descents"] + 16.0 * cnt["restart_depth"] + 9.0 * cnt["bases"] + 8.0 * n_mems}
    survey = {"pass1": 32.0 * (5 * cnt["bases"] + 6 * cnt["jumps"]), "pass2": 32.0 * (cnt["restart_depth"] + 2 * cnt["chars"])}
    per = {"pass1": k_ms[0] / args.steps, "pass2": k_ms[1] / args.steps}
    dom = "pass1" if per["pass1"] >= per["pass2"] else "pass2"
    oth = "pass2" if dom == "pass1" else "pass1"
    kname = {"pass1": "k_ms_pointers", "pass2": "k_ms_lengths"}
    peak, peak_src = measured_peak_gbs()
    achieved = alg[dom] / (per[dom] / 1e3) / 1e9
    rl = None
    try:
        rl = random_line_peak(m, local)
    except Exception as e:
        log("random-line probe failed:", e)
    # random lines the dominant kernel must fetch: one per LF step in pass 1, one per LCE restart that touches text in pass 2
    lines = {"pass1": float(cnt["bases"]), "pass2": float(cnt["descents"])}
    lines_per_s = lines[dom] / (per[dom] / 1e3)
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            tj = json.load(open(tp))
            if tj.get("workload") == args.workload:
                traffic = tj.get(dom)
        except Exception:
            traffic = None
    cpu = par = None
    if not args.no_cpu_baseline and world == 1 and os.path.exists(os.path.join(ROOT, "oracle", "_ref", "bin", "mems")):
        try:
            cpu, par = cpu_baseline_and_parity(args, prefix, h_seqs, n_reads, read_len, gpu_mems, gpu_ms)
        except Exception as e:  # the baseline is reported, never required
            import traceback
            traceback.print_exc()
            cpu = {"value": None, "unit": "reads/s", "cores": os.cpu_count(), "kind": "reference", "sample": f"failed: {e}"}
    line = {
        "metric": "mems_reads_per_s", "value": value, "unit": "reads/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
        "data": "synthetic", "bases_per_s": value * read_len,
        "config": {"workload": w["desc"], "workload_key": args.workload, "reads_per_gpu_per_step": int(n_reads), "read_len": int(read_len),
                   "index_n": ix.n, "index_runs": ix.r, "hbm_image_bytes": ix.device_bytes,
                   "l2": "inputs larger than L2 (index image and read batch)",
                   "parallelism": f"reads sharded x{world}, index replicated, no collective"},
        "stage_ms_per_step": {"pass1": per["pass1"], "pass2": per["pass2"], "compact": k_ms[2] / args.steps},
        "roofline": {"bound": "hbm", "kernel": kname[dom], "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "traffic_source": "profiles/traffic.json (ncu capture of this workload, not this run)" if traffic else None,
                     "peak_source": peak_src, "algorithmic_bytes_per_launch": alg[dom],
                     "algorithmic_bytes_per_base": alg[dom] / max(cnt["bases"], 1), "survey_model_bytes": survey[dom],
                     "random_line_peak": rl, "random_lines_per_launch": lines[dom], "random_lines_per_s": lines_per_s,
                     "frac_random": (lines_per_s / rl["value"]) if rl and rl.get("value") else None,
                     "other_kernel": {"kernel": kname[oth], "achieved": alg[oth] / (per[oth] / 1e3) / 1e9,
                                      "frac": alg[oth] / (per[oth] / 1e3) / 1e9 / peak}},
        "counters_per_step": cnt,
        "e2e": {**leg_json("mems_u64"), "host_widen_threads": widen_threads, "link_bytes_per_record": link_rec, "host_bytes_per_record": 16,
                "note": "u64 (query offset, length) columns in the caller's host arrays; with host_widen_threads > 0 the records cross the link "
                        "packed and host threads of the library widen them (csrc/host_widen.cpp); e2e_mems_u64_link is the same call with the "
                        "u64 columns copied over the link (option widen_threads=0)"},
        **{"e2e_" + k: leg_json(k) for k in legs if k.startswith("mems_") and k != "mems_u64"},
        "ms": {"value": value_ms, "unit": "reads/s", "bases_per_s": value_ms * read_len, "ms_per_step": dev_ms_ms / args.steps,
               "stage_ms_per_step": {"pass1": kms_ms[0] / args.steps, "pass2": kms_ms[1] / args.steps, "expand": kms_ms[2] / args.steps},
               "e2e": {**leg_json("ms_u64"), "host_widen_threads": widen_threads if widen_threads >= 6 else 0, "host_bytes_per_base": 16,
                       "note": "moni_gpu_ms widens on the host from 7 bytes per base when the pool has 6 threads or more"},
               **({"e2e_u64_link": leg_json("ms_u64_link")} if "ms_u64_link" in legs else {}), "e2e_packed": leg_json("ms_packed"), "e2e_runs": leg_json("ms_runs"),
               "e2e_runs_packed": leg_json("ms_runs_packed"), "runs_per_step": n_runs},
        "mem_records_per_step": n_mems,
        "gpu_launches": int(launches), "gpu_launches_ms_leg": int(launches_ms), "options": args.opt, "cpu_placement": numa_note,
        "host_buffers": f"pinned, NUMA node of the GPU ({m.load_library().moni_gpu_numa_node(local)})", "clocks": clocks,
        "cpu_baseline": cpu, "parity_sample": par, "index_load_s": t_load,
        **({"host_dma_ceiling": dma} if dma else {}), **({"e2e_sharded": sharded} if sharded else {}),
    }
    emit(line)
    if world > 1:
        dist.barrier(group=cpu_group)
        dist.destroy_process_group()


def main():
    import workloads
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(workloads.WORKLOADS))
    ap.add_argument("--pairs", type=int, default=0, help="read pairs per GPU per step (default: the workload's bench_pairs)")
    ap.add_argument("--ref-reads", type=int, default=100_000, help="reads in the bounded CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sharded", action="store_true", help="skip the in-library multi-GPU leg (N > 1)")
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
