#!/usr/bin/env python
"""BASELINE.json configs[2] (and a configs[4]-like ragged case on the same index) on the GPU box:

    python tests/scale/scale_run.py [--workload cfg3] [--out gpurun_out/scale_run.json]

The cfg3 index (40 Mbp x 32 haplotypes, n = 1.28 G) is too large to build inside a gpurun call with the reference's
pfp_thresholds (~10 min) AND to ship complete (the snapshot is limited to 512 MiB): `<prefix>.thrbv.ms` (the reference
builder's output, made in the build container) travels, the text is regenerated from its seed and the grammar file is
made here (tools/mkgrammar + the reference's SlpEncBuild).  If the .thrbv.ms is absent it is built here too.

Stages (each records its outcome in the JSON as soon as it is known):
  load        index load / export / upload time, image size
  kernels     `mems` and `ms` on resident slices of 2 M reads: CUDA-event time of pass 1 / pass 2 / compaction
  e2e         moni_gpu_ms, moni_gpu_mems_packed, moni_gpu_seeds(min_len 25) over ALL reads with pinned host buffers: wall
              time, reads/s, and a 64-bit hash of the complete result
  replicas    the same hash with the batch sharded over 2 and 3 replicas (input order must survive sharding)
  parity      a sample of reads (half R1, half R2) as FASTQ through the UNMODIFIED reference `ms` / `mems` (-t nproc) and
              through the drop-in moni_align_b200/bin/{ms,mems}: the output files must be byte-identical
  seeds       moni_gpu_seeds on a sample against the CPU oracle's restatement of seed_finder::find_mems
  long        ragged ONT-like reads (mean 10 kbp, 15 % errors): e2e rates and file parity against the reference
Checking at scale (test infrastructure: it runs the reference binaries and the CPU oracle as the checker); nothing here is on the product path.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import build_index  # noqa: E402
import synth  # noqa: E402
import workloads  # noqa: E402

RES: dict = {}
OUT = None


def save():
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    with open(OUT, "w") as f:
        json.dump(RES, f, indent=1)


def log(*a):
    print("[scale_run]", *a, file=sys.stderr, flush=True)


def hash64(*arrays) -> str:
    """Order-sensitive 64-bit hash of integer arrays: sum of (value + 1) * (2 * index + 1) mod 2^64, chained."""
    h = np.uint64(0x9E3779B97F4A7C15)
    blk = 1 << 26
    with np.errstate(over="ignore"):
        for a in arrays:
            a = np.asarray(a)
            for s in range(0, a.shape[0], blk):
                v = a[s:s + blk].astype(np.uint64) + np.uint64(1)
                w = (np.arange(s, s + v.shape[0], dtype=np.uint64) << np.uint64(1)) | np.uint64(1)
                h = h * np.uint64(0xD6E8FEB86659FD93) + np.add.reduce(v * w, dtype=np.uint64)
            h = h * np.uint64(0xD6E8FEB86659FD93) + np.uint64(a.shape[0])
    return f"{int(h):016x}"


def stage(name):
    def deco(fn):
        def run(*a, **k):
            t = time.time()
            try:
                RES[name] = fn(*a, **k)
            except Exception as e:  # keep going: every stage is reported on its own
                import traceback
                traceback.print_exc()
                RES[name] = {"error": f"{type(e).__name__}: {e}"}
            RES[name]["stage_s"] = round(time.time() - t, 2)
            log(name, json.dumps(RES[name])[:1500])
            save()
            return RES[name]
        return run
    return deco


def cli_vs_reference(prefix, reads, td, tag, cores):
    """reads -> FASTQ -> reference ms/mems and drop-in ms/mems; returns parity and the timers."""
    fq = os.path.join(td, tag + ".fq")
    synth.write_fastq_list(fq, reads)
    res = {"reads": len(reads), "bases": int(sum(len(r) for r in reads)), "threads": cores}
    for cmd, exts in (("ms", [".pointers", ".lengths"]), ("mems", [".mems"])):
        ref_out, our_out = os.path.join(td, tag + "_ref"), os.path.join(td, tag + "_our")
        r = build_index.run_reference(cmd, prefix, fq, ref_out, threads=cores)
        t = time.time()
        p = subprocess.run([os.path.join(ROOT, "moni_align_b200", "bin", cmd), prefix, "-p", fq, "-o", our_out, "-t", str(cores)],
                           capture_output=True)
        ours_wall = time.time() - t
        if p.returncode != 0:
            raise RuntimeError(f"drop-in {cmd} failed: {p.stderr[-800:]!r}")
        same = True
        for e in exts:
            c = subprocess.run(["cmp", "-s", ref_out + e, our_out + e])
            same = same and c.returncode == 0
        res[cmd] = {"identical_files": bool(same), "reference_phase_s": r["phase"], "reference_wall_s": round(r["wall"], 2),
                    "reference_reads_per_s": len(reads) / r["phase"] if r["phase"] else None,
                    "drop_in_wall_s_incl_index_load": round(ours_wall, 2),
                    "bytes": {e: os.path.getsize(ref_out + e) for e in exts}}
        for e in exts:
            os.remove(ref_out + e)
            os.remove(our_out + e)
    os.remove(fq)
    return res


def main():
    global OUT
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg3")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "scale_run.json"))
    ap.add_argument("--pairs", type=int, default=0)
    ap.add_argument("--slice-reads", type=int, default=2_000_000)
    ap.add_argument("--parity-reads", type=int, default=1_000_000)
    ap.add_argument("--seed-check-reads", type=int, default=20_000)
    ap.add_argument("--long-reads", type=int, default=20_000)
    ap.add_argument("--long-parity-reads", type=int, default=2_000)
    ap.add_argument("--ms-reads", type=int, default=10_000_000, help="reads of the batch that go through moni_gpu_ms (16 B of results per base)")
    ap.add_argument("--max-kernel-slices", type=int, default=8)
    args = ap.parse_args()
    OUT = args.out
    import moni_align_b200 as m
    w = workloads.WORKLOADS[args.workload]
    cores = os.cpu_count() or 1
    RES["workload"] = w["desc"]
    RES["host_cores"] = cores

    t0 = time.time()
    prefix = workloads.ensure_index(args.workload)
    RES["index_files_s"] = round(time.time() - t0, 1)
    RES["index_file_bytes"] = {e: os.path.getsize(prefix + e) for e in (".thrbv.ms", ".plain.slp")}
    save()

    @stage("load")
    def load():
        t = time.time()
        ix = m.MoniIndex(prefix, gpus=[0])
        load.ix = ix
        return {"seconds": round(time.time() - t, 1), "n": ix.n, "r": ix.r, "text_len": ix.text_len, "hbm_image_bytes": ix.device_bytes}
    load()
    ix = load.ix

    t0 = time.time()
    text = workloads.text_of(args.workload)
    reads = workloads.reads_of(args.workload, text=text, pairs=args.pairs or w["pairs"])
    n_reads, L = reads.shape
    total = n_reads * L
    RES["reads"] = {"n": int(n_reads), "len": int(L), "bases": int(total), "prepare_s": round(time.time() - t0, 1)}
    save()
    h_seqs = m.pinned_empty(total, np.uint8)
    h_seqs[:] = reads.reshape(-1)
    h_offs = m.pinned_empty(n_reads + 1, np.uint64)
    h_offs[:] = np.arange(n_reads + 1, dtype=np.uint64) * np.uint64(L)

    @stage("kernels")
    def kernels():
        out = {"slice_reads": args.slice_reads, "slices": []}
        m.set_option("count", 1)
        agg = np.zeros(3)
        agg_ms = np.zeros(2)
        cnt_all: dict = {}
        starts = list(range(0, n_reads, args.slice_reads))
        if len(starts) > args.max_kernel_slices:      # spread the measured slices over the batch (R1 half and R2 half)
            starts = [starts[int(j * (len(starts) - 1) / (args.max_kernel_slices - 1))] for j in range(args.max_kernel_slices)]
        measured = 0
        for s in starts:
            e = min(n_reads, s + args.slice_reads)
            measured += e - s
            offs = (h_offs[s:e + 1] - h_offs[s]).copy()
            rb = ix.resident_batch(h_seqs[s * L:e * L], offs)
            ix.counters(reset=True)
            rb.run("mems")
            cnt = ix.counters(reset=True)
            m.set_option("count", 0)
            rb.run("mems")
            t = [rb.run("mems") for _ in range(3)]
            tm = [rb.run("ms") for _ in range(3)]
            m.set_option("count", 1)
            best = min(t, key=lambda d: d["pass1_ms"] + d["pass2_ms"] + d["compact_ms"])
            best_ms = min(tm, key=lambda d: d["pass1_ms"] + d["pass2_ms"])
            agg += np.array([best["pass1_ms"], best["pass2_ms"], best["compact_ms"]])
            agg_ms += np.array([best_ms["pass1_ms"], best_ms["pass2_ms"]])
            for k, v in cnt.items():
                cnt_all[k] = cnt_all.get(k, 0) + v
            out["slices"].append({"reads": int(e - s), "mems_ms": best, "ms_ms": best_ms})
            rb.close()
        m.set_option("count", 0)
        out["mems_kernel_ms_total"] = dict(zip(["pass1", "pass2", "compact"], agg.tolist()))
        out["ms_kernel_ms_total"] = dict(zip(["pass1", "pass2"], agg_ms.tolist()))
        out["reads_measured"] = int(measured)
        out["mems_reads_per_s_kernels"] = measured / (agg.sum() / 1e3)
        out["ms_reads_per_s_kernels"] = measured / (agg_ms.sum() / 1e3)
        out["counters"] = cnt_all
        out["pass1_GBps_at_137B_per_base"] = 137.0 * measured * L / (agg[0] / 1e3) / 1e9
        return out
    kernels()

    @stage("e2e")
    def e2e():
        out = {}
        big = total > (1 << 32)
        cap = total // 2 if big else total           # records per base stay well below 1/2 on these workloads
        scap = total // 16 if big else total
        mo = m.pinned_empty(n_reads + 1, np.uint64)
        rec = m.pinned_empty(cap, np.uint32)
        ix.maxrimal_exact_matches_packed(h_seqs, h_offs, out=(mo, rec))
        t = time.perf_counter()
        mo_, rec_ = ix.maxrimal_exact_matches_packed(h_seqs, h_offs, out=(mo, rec))
        dt = time.perf_counter() - t
        out["mems_packed"] = {"wall_ms": dt * 1e3, "reads_per_s": n_reads / dt, "bases_per_s": total / dt, "records": int(mo[n_reads]),
                              "timing": ix.timing(), "hash": hash64(mo_, rec_ >> 16, rec_ & 0xFFFF)}
        e2e.mems_hash = out["mems_packed"]["hash"]
        # seeds, min_len 25
        ref = m.pinned_empty(scap, np.uint64)
        ln = m.pinned_empty(scap, np.uint32)
        ix.find_mems(h_seqs, h_offs, 25, out=(mo, ref, rec[:scap], ln))
        t = time.perf_counter()
        so, sr, sq, sl = ix.find_mems(h_seqs, h_offs, 25, out=(mo, ref, rec[:scap], ln))
        dt = time.perf_counter() - t
        out["seeds_min25"] = {"wall_ms": dt * 1e3, "reads_per_s": n_reads / dt, "records": int(so[n_reads]), "timing": ix.timing(),
                              "hash": hash64(so, sr, sq, sl)}
        # ms: 16 bytes of results per base
        k_ms = min(n_reads, args.ms_reads)
        t_ms = k_ms * L
        ptr = m.pinned_empty(t_ms, np.uint64)
        lnb = m.pinned_empty(t_ms, np.uint64)
        ix.matching_statistics(h_seqs[:t_ms], h_offs[:k_ms + 1], out=(ptr, lnb))
        t = time.perf_counter()
        p_, l_ = ix.matching_statistics(h_seqs[:t_ms], h_offs[:k_ms + 1], out=(ptr, lnb))
        dt = time.perf_counter() - t
        out["ms"] = {"reads": int(k_ms), "wall_ms": dt * 1e3, "reads_per_s": k_ms / dt, "bases_per_s": t_ms / dt, "d2h_bytes": 16 * t_ms,
                     "timing": ix.timing(), "hash": hash64(p_, l_)}
        lo, hi, l16 = m.pinned_empty(t_ms, np.uint32), m.pinned_empty(t_ms, np.uint8), m.pinned_empty(t_ms, np.uint16)
        ix.matching_statistics_packed(h_seqs[:t_ms], h_offs[:k_ms + 1], out=(lo, hi, l16))
        t = time.perf_counter()
        ix.matching_statistics_packed(h_seqs[:t_ms], h_offs[:k_ms + 1], out=(lo, hi, l16))
        dt = time.perf_counter() - t
        same = True
        for s0 in range(0, t_ms, 1 << 26):
            sl_ = slice(s0, s0 + (1 << 26))
            same = same and bool(np.array_equal(m.unpack_pointers(lo[sl_], hi[sl_]), p_[sl_]) and np.array_equal(l16[sl_], l_[sl_]))
        out["ms_packed"] = {"reads": int(k_ms), "wall_ms": dt * 1e3, "reads_per_s": k_ms / dt, "bases_per_s": t_ms / dt, "d2h_bytes": 7 * t_ms,
                            "timing": ix.timing(), "same_as_ms": same}
        out["ms"]["max_pointer"] = int(p_[p_ < np.uint64(1 << 62)].max())
        out["ms"]["pointers_at_or_beyond_2^32"] = int((p_[p_ < np.uint64(1 << 62)] >= np.uint64(1 << 32)).sum())
        ro = m.pinned_empty(k_ms + 1, np.uint64)
        rq, rl, rp = m.pinned_empty(t_ms, np.uint32), m.pinned_empty(t_ms, np.uint32), m.pinned_empty(t_ms, np.uint64)
        ix.matching_statistics_runs(h_seqs[:t_ms], h_offs[:k_ms + 1], out=(ro, rq, rl, rp))
        t = time.perf_counter()
        ix.matching_statistics_runs(h_seqs[:t_ms], h_offs[:k_ms + 1], out=(ro, rq, rl, rp))
        dt = time.perf_counter() - t
        nrun = int(ro[k_ms])
        xp, xl = m.expand_runs(h_offs[:k_ms + 1], ro, rq[:nrun], rl[:nrun], rp[:nrun])
        out["ms_runs"] = {"reads": int(k_ms), "wall_ms": dt * 1e3, "reads_per_s": k_ms / dt, "runs": nrun, "d2h_bytes": 16 * nrun + 8 * (k_ms + 1),
                          "timing": ix.timing(), "expands_to_ms": bool(np.array_equal(xp, p_) and np.array_equal(xl, l_))}
        del xp, xl
        # properties of the complete result: lengths are matching statistics of the pointers
        Lm = l_.reshape(-1, L)
        ok1 = ok2 = True
        cap = (L - np.arange(L)).astype(np.uint64)[None, :]
        for s in range(0, k_ms, 1 << 20):
            blk = Lm[s:s + (1 << 20)]
            ok1 = ok1 and bool((blk[:, 1:] + np.uint64(1) >= blk[:, :-1]).all())
            ok2 = ok2 and bool((blk <= cap).all())
        out["ms"]["len_drops_by_at_most_one"] = ok1
        out["ms"]["len_within_read"] = ok2
        sel = np.random.default_rng(1).integers(0, k_ms, size=200_000)
        P = p_.reshape(-1, L)[sel].astype(np.int64)
        Ls = Lm[sel].astype(np.int64)
        rr = reads[sel]
        nz = Ls > 0
        i_idx = np.broadcast_to(np.arange(L)[None, :], Ls.shape)
        p, l, i, r = P[nz], Ls[nz], i_idx[nz], np.nonzero(nz)[0]
        out["ms"]["first_and_last_matched_char_equal"] = bool((text[p] == rr[r, i]).all() and (text[p + l - 1] == rr[r, i + l - 1]).all())
        ext = (i + l < L) & (p + l < text.shape[0])
        out["ms"]["right_maximal"] = bool((text[(p + l)[ext]] != rr[r[ext], (i + l)[ext]]).all())
        return out
    e2e()

    @stage("replicas")
    def replicas():
        out = {}
        k = min(n_reads, args.slice_reads)
        offs = h_offs[:k + 1]
        one = ix.maxrimal_exact_matches_packed(h_seqs[:k * L], offs)
        out["reads"] = int(k)
        out["hash_1"] = hash64(one[0], one[1])
        ndev = int(subprocess.run(["nvidia-smi", "-L"], capture_output=True).stdout.count(b"GPU "))
        for g in (2, 3):
            devs = [d % max(ndev, 1) for d in range(g)]
            ixg = m.MoniIndex(prefix, gpus=devs)
            got = ixg.maxrimal_exact_matches_packed(h_seqs[:k * L], offs)
            so = ixg.find_mems(h_seqs[:k * L], offs, 25)
            so1 = ix.find_mems(h_seqs[:k * L], offs, 25)
            out[f"hash_{g}"] = hash64(got[0], got[1])
            out[f"devices_{g}"] = devs
            out[f"seeds_same_{g}"] = bool(all(np.array_equal(a, b) for a, b in zip(so, so1)))
            ixg.close()
        out["identical"] = bool(out["hash_1"] == out["hash_2"] == out["hash_3"])
        return out
    replicas()

    @stage("seeds")
    def seeds():
        from oracle import oracle
        orc = oracle.OracleIndex(prefix)
        k = args.seed_check_reads
        sel = np.concatenate([np.arange(k // 2), n_reads // 2 + np.arange(k // 2)])
        sub = reads[sel]
        flat, offs = m.pack_reads(sub)
        out = {"reads": int(sub.shape[0])}
        for ml in (0, 25):
            t = time.time()
            want = orc.seeds(flat, offs, ml)
            dt = time.time() - t
            got = ix.find_mems(flat, offs, ml)
            out[f"min_len_{ml}"] = {"identical_to_oracle": bool(all(np.array_equal(a, b) for a, b in zip(got, want))),
                                    "records": int(want[0][-1]), "oracle_s": round(dt, 2)}
        return out
    seeds()

    @stage("parity")
    def parity():
        k = min(args.parity_reads, n_reads)
        sel = np.concatenate([np.arange(k // 2), n_reads // 2 + np.arange(k - k // 2)])
        with tempfile.TemporaryDirectory() as td:
            return cli_vs_reference(prefix, list(reads[sel]), td, "short", cores)
    parity()

    @stage("edge")
    def edge():
        """The edge-case file of SURVEY.md 8(d) on this index: N, lowercase, symbols absent from the text, 1-base reads, reads longer
        than any match, reads that run into the end of the text -- reference binaries vs the drop-in CLIs, byte for byte."""
        rng = np.random.default_rng(4242)
        N = text.shape[0]
        lst = [np.frombuffer(b"A", dtype=np.uint8), np.frombuffer(b"N" * 60, dtype=np.uint8), np.frombuffer(b"acgtacgtacgtacgtnnnn", dtype=np.uint8),
               text[N - 150:N].copy(), text[N - 40:N].copy(), text[0:200].copy(),
               np.frombuffer(b"ACGTNacgtnRYKMxyz0123456789", dtype=np.uint8)[rng.integers(0, 27, size=300)],
               synth.ACGT[rng.integers(0, 4, size=700, dtype=np.uint8)]]
        for _ in range(200):
            s0 = int(rng.integers(0, N - 400))
            r = text[s0:s0 + int(rng.integers(20, 400))].copy()
            k = int(rng.integers(0, 4))
            for q in rng.integers(0, r.shape[0], size=k):
                r[q] = ord("N") if rng.random() < 0.5 else ord("x")
            lst.append(r)
        # far ends of the text: pointers beyond 2^32 when the text is that long
        for s0 in (N - 5000, N // 2, (1 << 32) + 12345 if N > (1 << 32) + 20000 else N // 3):
            lst.append(text[s0:s0 + 250].copy())
        with tempfile.TemporaryDirectory() as td:
            return cli_vs_reference(prefix, lst, td, "edge", cores)
    edge()

    m.release_pinned()

    @stage("long")
    def long_reads():
        lr = synth.long_reads(text, args.long_reads, 501)
        flat, offs = m.pack_reads(lr)
        tot, n = int(offs[-1]), len(lr)
        out = {"reads": n, "bases": tot, "max_len": int(max(len(r) for r in lr))}
        hs = m.pinned_empty(tot, np.uint8)
        hs[:] = flat
        ho = m.pinned_empty(n + 1, np.uint64)
        ho[:] = offs
        mo = m.pinned_empty(n + 1, np.uint64)
        rec = m.pinned_empty(tot, np.uint32)
        ix.maxrimal_exact_matches_packed(hs, ho, out=(mo, rec))
        t = time.perf_counter()
        ix.maxrimal_exact_matches_packed(hs, ho, out=(mo, rec))
        dt = time.perf_counter() - t
        out["mems_packed"] = {"wall_ms": dt * 1e3, "reads_per_s": n / dt, "bases_per_s": tot / dt, "records": int(mo[n]), "timing": ix.timing()}
        a, b = m.pinned_empty(tot, np.uint64), m.pinned_empty(tot, np.uint64)
        ix.matching_statistics(hs, ho, out=(a, b))
        t = time.perf_counter()
        ix.matching_statistics(hs, ho, out=(a, b))
        dt = time.perf_counter() - t
        out["ms"] = {"wall_ms": dt * 1e3, "reads_per_s": n / dt, "bases_per_s": tot / dt, "timing": ix.timing()}
        with tempfile.TemporaryDirectory() as td:
            out["parity"] = cli_vs_reference(prefix, lr[:args.long_parity_reads], td, "long", cores)
        m.release_pinned()
        return out
    long_reads()
    print(json.dumps(RES))
    bad = [k for k, v in RES.items() if isinstance(v, dict) and "error" in v]
    if bad:
        log("stages with errors:", bad)
        sys.exit(1)


if __name__ == "__main__":
    main()
