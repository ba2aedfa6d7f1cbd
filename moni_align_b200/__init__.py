"""moni_align_b200 -- B200 (sm_100a) matching statistics and MEMs for MONI indexes.

Host-side mirror of the reference's two call sites for this path, over the C ABI of
``libmoni_gpu.so`` (include/moni_gpu.h):

    reference (C++)                                            here
    ---------------------------------------------------------  ---------------------------------
    ms_c<slp_t>::ms_c(prefix)      src/matching_statistics.cpp:195   MoniIndex(prefix, gpus=[...])
    ms_c::matching_statistics      src/matching_statistics.cpp:242   MoniIndex.matching_statistics
    mems_c::maxrimal_exact_matches src/mems.cpp:241                  MoniIndex.maxrimal_exact_matches
    seed_finder::find_mems         include/aligner/seed_finder.hpp:126   MoniIndex.find_mems
    seed_finder::find_MEM_occs     include/aligner/seed_finder.hpp:239   MoniIndex.find_MEM_occs

The method names (including the reference's spelling of "maxrimal") and the meaning of the results
are the reference's; the unit of work is a batch of reads instead of one kseq record.  There is no
CPU execution path: constructing a MoniIndex without the built library or without a CUDA device
raises.  Build the library with ``__graft_entry__.build()`` or ``make -C moni_align_b200/csrc``.
"""
from __future__ import annotations

import ctypes
import os
from typing import Iterable, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MONI_GPU_LIB") or os.path.join(_HERE, "libmoni_gpu.so")   # the override is for A/B builds of the same library

_u64p = ctypes.POINTER(ctypes.c_uint64)
NOALLOC_MAX_READS, NOALLOC_MAX_BASES = 65536, 1048576      # MONI_GPU_NOALLOC_MAX_* of include/moni_gpu.h
_lib = None


class MoniGpuError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libmoni_gpu error {code}: {msg}")
        self.code = code


EXPORTS = [
    "moni_gpu_index_load", "moni_gpu_index_free", "moni_gpu_index_info", "moni_gpu_ms", "moni_gpu_mems", "moni_gpu_mems32", "moni_gpu_mems_packed16",
    "moni_gpu_mems_packed", "moni_gpu_seeds", "moni_gpu_seed_occs", "moni_gpu_ms_packed", "moni_gpu_host_alloc_near", "moni_gpu_numa_node",
    "moni_gpu_counters", "moni_gpu_timing", "moni_gpu_set_option", "moni_gpu_host_alloc", "moni_gpu_host_free",
    "moni_gpu_batch_create", "moni_gpu_batch_run", "moni_gpu_batch_fetch_ms", "moni_gpu_batch_fetch_mems",
    "moni_gpu_batch_free", "moni_gpu_launch_count", "moni_gpu_last_error", "moni_gpu_version",
    "moni_gpu_probe_random_lines", "moni_gpu_probe_release", "moni_gpu_ms_runs", "moni_gpu_ms_runs_packed", "moni_gpu_ms_runs_expand",
    "moni_gpu_populate_seeds", "moni_gpu_ms_noalloc",
]


def load_library() -> ctypes.CDLL:
    """dlopen libmoni_gpu.so (in-tree) and declare the prototypes of include/moni_gpu.h."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback)")
    L = ctypes.CDLL(LIB_PATH)
    vp, ci, cc = ctypes.c_void_p, ctypes.c_int, ctypes.c_char_p
    L.moni_gpu_index_load.argtypes = [cc, ctypes.POINTER(ci), ci, ctypes.POINTER(vp)]
    L.moni_gpu_index_free.argtypes = [vp]
    L.moni_gpu_index_free.restype = None
    L.moni_gpu_index_info.argtypes = [vp, _u64p, _u64p, _u64p, _u64p]
    L.moni_gpu_ms.argtypes = [vp, vp, _u64p, ctypes.c_uint64, _u64p, _u64p]
    L.moni_gpu_mems.argtypes = [vp, vp, _u64p, ctypes.c_uint64, _u64p, _u64p, _u64p, ctypes.c_uint64]
    L.moni_gpu_mems32.argtypes = [vp, vp, _u64p, ctypes.c_uint64, _u64p, vp, vp, ctypes.c_uint64]
    L.moni_gpu_mems_packed.argtypes = [vp, vp, _u64p, ctypes.c_uint64, _u64p, vp, ctypes.c_uint64]
    L.moni_gpu_mems_packed16.argtypes = [vp, vp, _u64p, ctypes.c_uint64, _u64p, vp, ctypes.c_uint64]
    L.moni_gpu_seeds.argtypes = [vp, vp, _u64p, ctypes.c_uint64, ctypes.c_uint64, _u64p, _u64p, vp, vp, ctypes.c_uint64]
    L.moni_gpu_seed_occs.argtypes = [vp, _u64p, vp, ctypes.c_uint64, ctypes.c_uint64, _u64p, _u64p, ctypes.c_uint64, ctypes.POINTER(ci)]
    L.moni_gpu_populate_seeds.argtypes = [vp, _u64p, vp, vp, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint64, _u64p, _u64p, ctypes.c_uint64,
                                          _u64p, _u64p, _u64p, vp, vp, _u64p, _u64p, ctypes.c_uint64, ctypes.POINTER(ci)]
    L.moni_gpu_counters.argtypes = [vp, _u64p, ci]
    L.moni_gpu_timing.argtypes = [vp, ctypes.POINTER(ctypes.c_double)]
    L.moni_gpu_set_option.argtypes = [cc, cc]
    L.moni_gpu_host_alloc.argtypes = [ctypes.c_uint64]
    L.moni_gpu_host_alloc.restype = vp
    L.moni_gpu_host_alloc_near.argtypes = [ci, ctypes.c_uint64]
    L.moni_gpu_host_alloc_near.restype = vp
    L.moni_gpu_numa_node.argtypes = [ci]
    L.moni_gpu_ms_packed.argtypes = [vp, vp, _u64p, ctypes.c_uint64, vp, vp, vp]
    L.moni_gpu_host_free.argtypes = [vp]
    L.moni_gpu_host_free.restype = None
    L.moni_gpu_ms_runs.argtypes = [vp, vp, _u64p, ctypes.c_uint64, _u64p, vp, vp, _u64p, ctypes.c_uint64]
    L.moni_gpu_ms_runs_packed.argtypes = [vp, vp, _u64p, ctypes.c_uint64, _u64p, vp, vp, vp, ctypes.c_uint64]
    L.moni_gpu_ms_runs_expand.argtypes = [_u64p, ctypes.c_uint64, _u64p, vp, vp, _u64p, _u64p, _u64p]
    if hasattr(L, "moni_gpu_ms_noalloc"):        # absent from A/B builds of older trees (MONI_GPU_LIB)
        L.moni_gpu_ms_noalloc.argtypes = [vp, ci, vp, vp, ctypes.c_uint64, ctypes.c_uint64, vp, vp, vp]
    L.moni_gpu_batch_create.argtypes = [vp, ci, vp, _u64p, ctypes.c_uint64, ctypes.POINTER(vp)]
    L.moni_gpu_batch_run.argtypes = [vp, ci, ctypes.POINTER(ctypes.c_float)]
    L.moni_gpu_batch_fetch_ms.argtypes = [vp, _u64p, _u64p]
    L.moni_gpu_batch_fetch_mems.argtypes = [vp, _u64p, _u64p, _u64p, ctypes.c_uint64]
    L.moni_gpu_batch_free.argtypes = [vp]
    L.moni_gpu_batch_free.restype = None
    L.moni_gpu_launch_count.restype = ctypes.c_uint64
    L.moni_gpu_probe_random_lines.argtypes = [ci, ctypes.c_uint64, ci, ci, ci, ci, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ci)]
    L.moni_gpu_probe_release.restype = None
    L.moni_gpu_last_error.restype = cc
    L.moni_gpu_version.restype = cc
    _lib = L
    return L


def _check(rc: int) -> None:
    if rc != 0:
        raise MoniGpuError(rc, load_library().moni_gpu_last_error().decode(errors="replace"))


def set_option(key: str, value) -> None:
    _check(load_library().moni_gpu_set_option(key.encode(), str(value).encode()))


def launch_count() -> int:
    return int(load_library().moni_gpu_launch_count())


def pinned_empty(n: int, dtype, gpu: int | None = None) -> np.ndarray:
    """numpy array over cudaHostAlloc memory (on the NUMA node of CUDA device `gpu` when given).  The allocation lives until
    release_pinned() or exit."""
    L = load_library()
    dt = np.dtype(dtype)
    nbytes = max(int(n) * dt.itemsize, 1)
    p = L.moni_gpu_host_alloc(nbytes) if gpu is None else L.moni_gpu_host_alloc_near(int(gpu), nbytes)
    if not p:
        raise MoniGpuError(-4, "pinned allocation failed")
    buf = (ctypes.c_uint8 * nbytes).from_address(p)
    arr = np.frombuffer(buf, dtype=dt, count=int(n))
    _PINNED[p] = nbytes
    return arr


def release_pinned() -> None:
    """Frees every pinned_empty() allocation (the arrays must no longer be used)."""
    L = load_library()
    for p in list(_PINNED):
        L.moni_gpu_host_free(p)
        del _PINNED[p]


_PINNED: dict = {}


def probe_random_lines(gpu: int = 0, table_bytes: int = 3 << 30, shape: int = 0, chains_per_thread: int = 1, blocks_per_sm: int = 6,
                       iters: int = 512) -> dict:
    """Random 128-byte lines per second of dependent chains over a table in HBM (moni_gpu_probe_random_lines): the roofline
    denominator of SURVEY.md 8(d).  Returns {"lines_per_s", "threads_per_sm", "chains_per_sm"}."""
    L = load_library()
    v, t = ctypes.c_double(0), ctypes.c_int(0)
    _check(L.moni_gpu_probe_random_lines(int(gpu), int(table_bytes), int(shape), int(chains_per_thread), int(blocks_per_sm), int(iters),
                                         ctypes.byref(v), ctypes.byref(t)))
    return {"lines_per_s": float(v.value), "threads_per_sm": int(t.value), "chains_per_sm": int(t.value) * int(chains_per_thread),
            "shape": ["ldg256", "ldg128", "cp.async16"][shape], "table_bytes": int(table_bytes)}


def probe_release() -> None:
    load_library().moni_gpu_probe_release()


def pack_reads(reads) -> tuple[np.ndarray, np.ndarray]:
    """(seqs uint8[total], offsets uint64[n+1]) from a 2-D uint8 matrix or a list of bytes/arrays."""
    if isinstance(reads, np.ndarray) and reads.ndim == 2:
        n, m = reads.shape
        return np.ascontiguousarray(reads, dtype=np.uint8).reshape(-1), np.arange(n + 1, dtype=np.uint64) * np.uint64(m)
    lens = np.fromiter((len(r) for r in reads), dtype=np.uint64, count=len(reads))
    offs = np.zeros(len(reads) + 1, dtype=np.uint64)
    np.cumsum(lens, out=offs[1:])
    parts = [np.frombuffer(bytes(r), dtype=np.uint8) if not isinstance(r, np.ndarray) else r.astype(np.uint8, copy=False)
             for r in reads]
    seqs = np.concatenate(parts) if parts and int(offs[-1]) else np.zeros(0, dtype=np.uint8)
    return seqs, offs


def _ptr(a: np.ndarray):
    return ctypes.c_void_p(a.ctypes.data)


def _u64(a: np.ndarray):
    return a.ctypes.data_as(_u64p)


class MoniIndex:
    """An index replicated in the HBM of one or more GPUs (reference: ms_c / mems_c objects)."""

    def __init__(self, prefix: str, gpus: Sequence[int] | None = None):
        L = load_library()
        self._L = L
        gpus = [0] if gpus is None else list(gpus)
        arr = (ctypes.c_int * len(gpus))(*gpus)
        h = ctypes.c_void_p()
        _check(L.moni_gpu_index_load(os.fsencode(prefix), arr, len(gpus), ctypes.byref(h)))
        self._h = h
        self.gpus = gpus
        v = [ctypes.c_uint64() for _ in range(4)]
        _check(L.moni_gpu_index_info(h, *[ctypes.byref(x) for x in v]))
        self.n, self.r, self.text_len, self.device_bytes = (int(x.value) for x in v)

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._L.moni_gpu_index_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- the two reference entry points, batched --------------------------------------------
    def matching_statistics(self, seqs: np.ndarray, offsets: np.ndarray, out=None):
        """(pointers, lengths), uint64[offsets[-1]] each -- ms_c::matching_statistics for every read."""
        seqs = np.ascontiguousarray(seqs, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        total = int(offsets[-1])
        nr = offsets.shape[0] - 1
        if out is None:
            ptr = np.empty(max(total, 1), dtype=np.uint64)
            ln = np.empty(max(total, 1), dtype=np.uint64)
        else:
            ptr, ln = out
        _check(self._L.moni_gpu_ms(self._h, _ptr(seqs), _u64(offsets), nr, _u64(ptr), _u64(ln)))
        return ptr[:total], ln[:total]

    def matching_statistics_noalloc(self, d_seqs: int, d_offsets: int, n_reads: int, n_bases: int, d_pointers: int, d_lengths: int,
                                    stream: int = 0, gpu_slot: int = 0) -> None:
        """moni_gpu_ms_noalloc: matching_statistics on DEVICE buffers (raw addresses on GPU `gpus[gpu_slot]`, e.g. torch
        tensors' data_ptr()), with the library's fixed scratch array instead of allocated scratch -- the call allocates nothing.
        At most NOALLOC_MAX_READS reads / NOALLOC_MAX_BASES bases per call.  No counterpart in the reference."""
        _check(self._L.moni_gpu_ms_noalloc(self._h, int(gpu_slot), ctypes.c_void_p(int(d_seqs)), ctypes.c_void_p(int(d_offsets)), int(n_reads),
                                           int(n_bases), ctypes.c_void_p(int(d_pointers)), ctypes.c_void_p(int(d_lengths)),
                                           ctypes.c_void_p(int(stream))))

    def matching_statistics_packed(self, seqs: np.ndarray, offsets: np.ndarray, out=None):
        """(ptr_lo uint32, ptr_hi uint8, lengths uint16) -- moni_gpu_ms_packed; unpack_pointers() gives the u64 pointers."""
        seqs = np.ascontiguousarray(seqs, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        total = int(offsets[-1])
        nr = offsets.shape[0] - 1
        if out is None:
            lo, hi, ln = np.empty(max(total, 1), np.uint32), np.empty(max(total, 1), np.uint8), np.empty(max(total, 1), np.uint16)
        else:
            lo, hi, ln = out
        _check(self._L.moni_gpu_ms_packed(self._h, _ptr(seqs), _u64(offsets), nr, _ptr(lo), _ptr(hi), _ptr(ln)))
        return lo[:total], hi[:total], ln[:total]

    def matching_statistics_runs(self, seqs: np.ndarray, offsets: np.ndarray, out=None, packed: bool = False):
        """ms_c::matching_statistics as restart runs (moni_gpu_ms_runs): (run_offsets uint64[n+1], run_qpos uint32[k], run_len uint32[k],
        run_ptr uint64[k]); packed=True (moni_gpu_ms_runs_packed): (run_offsets, run_rec uint32[k], ptr_lo uint32[k], ptr_hi uint8[k]).
        expand_runs() rebuilds the reference's two vectors."""
        seqs = np.ascontiguousarray(seqs, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        total = int(offsets[-1])
        nr = offsets.shape[0] - 1
        if out is None:
            ro = np.empty(nr + 1, dtype=np.uint64)
            a = np.empty(max(total, 1), dtype=np.uint32)
            b = np.empty(max(total, 1), dtype=np.uint32)
            c = np.empty(max(total, 1), dtype=np.uint8 if packed else np.uint64)
        else:
            ro, a, b, c = out
        if packed:
            _check(self._L.moni_gpu_ms_runs_packed(self._h, _ptr(seqs), _u64(offsets), nr, _u64(ro), _ptr(a), _ptr(b), _ptr(c), a.shape[0]))
        else:
            _check(self._L.moni_gpu_ms_runs(self._h, _ptr(seqs), _u64(offsets), nr, _u64(ro), _ptr(a), _ptr(b), _u64(c), a.shape[0]))
        k = int(ro[nr])
        return ro, a[:k], b[:k], c[:k]

    def maxrimal_exact_matches(self, seqs: np.ndarray, offsets: np.ndarray, out=None, dtype=np.uint64):
        """(mem_offsets uint64[n+1], mem_pos, mem_len) -- mems_c::maxrimal_exact_matches for every read;
        read r owns records mem_offsets[r]:mem_offsets[r+1], each (query offset, length).  dtype=np.uint32 (or uint32
        `out` arrays) selects moni_gpu_mems32: same records, half the device->host bytes."""
        seqs = np.ascontiguousarray(seqs, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        total = int(offsets[-1])
        nr = offsets.shape[0] - 1
        if out is None:
            mo = np.empty(nr + 1, dtype=np.uint64)
            mp = np.empty(max(total, 1), dtype=dtype)
            ml = np.empty(max(total, 1), dtype=dtype)
        else:
            mo, mp, ml = out
        if mp.dtype == np.uint32:
            assert ml.dtype == np.uint32
            _check(self._L.moni_gpu_mems32(self._h, _ptr(seqs), _u64(offsets), nr, _u64(mo), _ptr(mp), _ptr(ml), mp.shape[0]))
        else:
            _check(self._L.moni_gpu_mems(self._h, _ptr(seqs), _u64(offsets), nr, _u64(mo), _u64(mp), _u64(ml), mp.shape[0]))
        k = int(mo[nr])
        return mo, mp[:k], ml[:k]

    def maxrimal_exact_matches_packed16(self, seqs: np.ndarray, offsets: np.ndarray, out=None):
        """(mem_offsets uint64[n+1], mem_rec uint16[k]) with mem_rec = (query offset << 8) | length -- moni_gpu_mems_packed16;
        every read must be shorter than 256 bases."""
        seqs = np.ascontiguousarray(seqs, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        nr = offsets.shape[0] - 1
        if out is None:
            mo = np.empty(nr + 1, dtype=np.uint64)
            rec = np.empty(max(int(offsets[-1]), 1), dtype=np.uint16)
        else:
            mo, rec = out
        _check(self._L.moni_gpu_mems_packed16(self._h, _ptr(seqs), _u64(offsets), nr, _u64(mo), _ptr(rec), rec.shape[0]))
        return mo, rec[:int(mo[nr])]

    def maxrimal_exact_matches_packed(self, seqs: np.ndarray, offsets: np.ndarray, out=None):
        """(mem_offsets uint64[n+1], mem_rec uint32[k]) with mem_rec = (query offset << 16) | length -- moni_gpu_mems_packed;
        every read must be shorter than 65536 bases."""
        seqs = np.ascontiguousarray(seqs, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        nr = offsets.shape[0] - 1
        if out is None:
            mo = np.empty(nr + 1, dtype=np.uint64)
            rec = np.empty(max(int(offsets[-1]), 1), dtype=np.uint32)
        else:
            mo, rec = out
        _check(self._L.moni_gpu_mems_packed(self._h, _ptr(seqs), _u64(offsets), nr, _u64(mo), _ptr(rec), rec.shape[0]))
        return mo, rec[:int(mo[nr])]

    def find_mems(self, seqs: np.ndarray, offsets: np.ndarray, min_len: int = 0, out=None):
        """(seed_offsets uint64[n+1], seed_ref uint64[k], seed_qpos uint32[k], seed_len uint32[k]) -- seed_finder::find_mems
        (include/aligner/seed_finder.hpp:126-166) for every read: mem_t.pos, mem_t.idx, mem_t.len."""
        seqs = np.ascontiguousarray(seqs, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        total = int(offsets[-1])
        nr = offsets.shape[0] - 1
        if out is None:
            so = np.empty(nr + 1, dtype=np.uint64)
            ref = np.empty(max(total, 1), dtype=np.uint64)
            qp = np.empty(max(total, 1), dtype=np.uint32)
            ln = np.empty(max(total, 1), dtype=np.uint32)
        else:
            so, ref, qp, ln = out
        _check(self._L.moni_gpu_seeds(self._h, _ptr(seqs), _u64(offsets), nr, int(min_len), _u64(so), _u64(ref), _ptr(qp), _ptr(ln),
                                      ref.shape[0]))
        k = int(so[nr])
        return so, ref[:k], qp[:k], ln[:k]

    def find_MEM_occs(self, seed_ref: np.ndarray, seed_len: np.ndarray, max_per_direction: int = 1 << 20, capacity: int | None = None,
                      out=None):
        """(occ_offsets uint64[n+1], occs uint64[k], truncated) -- seed_finder::find_MEM_occs (unfiltered) for every seed:
        its own position, then the occurrences above and below it in suffix-array order."""
        seed_ref = np.ascontiguousarray(seed_ref, dtype=np.uint64)
        seed_len = np.ascontiguousarray(seed_len, dtype=np.uint32)
        n = seed_ref.shape[0]
        tr = ctypes.c_int(0)
        if out is not None:                      # caller's (pinned) arrays: offsets uint64[n+1], positions uint64[capacity]
            off, occs = out
            capacity = cap = occs.shape[0]
        else:
            off = np.empty(n + 1, dtype=np.uint64)
            cap = max(4 * n, 1) if capacity is None else int(capacity)
            occs = np.empty(max(cap, 1), dtype=np.uint64)
        rc = self._L.moni_gpu_seed_occs(self._h, _u64(seed_ref), _ptr(seed_len), n, int(max_per_direction), _u64(off), _u64(occs), cap,
                                        ctypes.byref(tr))
        if rc == -4 and capacity is None and n:      # capacity was a guess: the needed size is in off[n]
            cap = int(off[n])
            occs = np.empty(cap, dtype=np.uint64)
            rc = self._L.moni_gpu_seed_occs(self._h, _u64(seed_ref), _ptr(seed_len), n, int(max_per_direction), _u64(off), _u64(occs), cap,
                                            ctypes.byref(tr))
        _check(rc)
        return off, occs[:int(off[n])], bool(tr.value)

    def populate_seeds(self, seed_offsets, seed_ref, seed_len, seed_qpos, min_len: int, max_per_direction: int = 1 << 20):
        """seed_finder::populate_seeds (seed filtering off) for a batch: returns the reads' final seed vectors in the reference's order
        -- (offsets uint64[n_reads+1], ref uint64, len uint32, qpos uint32, occ_offsets uint64, occs uint64, truncated): per read its
        seeds with their occurrence lists, then the two halves of every seed of at least 2 * min_len bases (moni_gpu_populate_seeds)."""
        seed_offsets = np.ascontiguousarray(seed_offsets, dtype=np.uint64)
        seed_ref = np.ascontiguousarray(seed_ref, dtype=np.uint64)
        seed_len = np.ascontiguousarray(seed_len, dtype=np.uint32)
        seed_qpos = np.ascontiguousarray(seed_qpos, dtype=np.uint32)
        n = seed_ref.shape[0]
        tr, nh = ctypes.c_int(0), ctypes.c_uint64(0)
        off = np.empty(n + 1, dtype=np.uint64)
        hoff = np.empty(2 * n + 1, dtype=np.uint64)
        hpar, href = np.empty(max(2 * n, 1), dtype=np.uint64), np.empty(max(2 * n, 1), dtype=np.uint64)
        hlen, hq = np.empty(max(2 * n, 1), dtype=np.uint32), np.empty(max(2 * n, 1), dtype=np.uint32)
        cap, hcap = max(4 * n, 1), max(8 * n, 1)
        for _ in range(3):
            occs, hoccs = np.empty(cap, dtype=np.uint64), np.empty(hcap, dtype=np.uint64)
            rc = self._L.moni_gpu_populate_seeds(self._h, _u64(seed_ref), _ptr(seed_len), _ptr(seed_qpos), n, int(min_len), int(max_per_direction),
                                                 _u64(off), _u64(occs), cap, ctypes.byref(nh), _u64(hpar), _u64(href), _ptr(hlen), _ptr(hq),
                                                 _u64(hoff), _u64(hoccs), hcap, ctypes.byref(tr))
            if rc == -4 and n:                   # a capacity was a guess: the sizes needed are in off[n] / hoff[n_halves]
                cap = max(cap, int(off[n]))
                hcap = max(hcap, int(hoff[int(nh.value)]) if int(nh.value) else hcap)
                continue
            break
        _check(rc)
        k = int(nh.value)
        # the reference's per-read vectors: the read's seeds, then their halves in seed order
        o_ref, o_len, o_q, lists = [], [], [], []
        offs_out = [0]
        read_of_half = np.searchsorted(seed_offsets, hpar[:k], side="right") - 1 if k else np.zeros(0, dtype=np.int64)
        hstart = np.searchsorted(read_of_half, np.arange(seed_offsets.shape[0] - 1), side="left") if k else np.zeros(seed_offsets.shape[0] - 1, dtype=np.int64)
        hend = np.searchsorted(read_of_half, np.arange(seed_offsets.shape[0] - 1), side="right") if k else np.zeros(seed_offsets.shape[0] - 1, dtype=np.int64)
        for r in range(seed_offsets.shape[0] - 1):
            a, b = int(seed_offsets[r]), int(seed_offsets[r + 1])
            o_ref.append(seed_ref[a:b]); o_len.append(seed_len[a:b]); o_q.append(seed_qpos[a:b])
            lists.append((off[a:b + 1], occs))
            ha, hb = int(hstart[r]), int(hend[r])
            o_ref.append(href[ha:hb]); o_len.append(hlen[ha:hb]); o_q.append(hq[ha:hb])
            lists.append((hoff[ha:hb + 1], hoccs))
            offs_out.append(offs_out[-1] + (b - a) + (hb - ha))
        cnts, segs = [], []
        for lo, src in lists:
            if lo.shape[0] > 1:
                cnts.append(np.diff(lo.astype(np.int64)))
                segs.append(src[int(lo[0]):int(lo[-1])])
        oo = np.zeros(offs_out[-1] + 1, dtype=np.uint64)
        if cnts:
            np.cumsum(np.concatenate(cnts), out=oo[1:])
        cat = lambda xs, dt: np.concatenate(xs).astype(dt) if xs else np.zeros(0, dtype=dt)
        return (np.array(offs_out, dtype=np.uint64), cat(o_ref, np.uint64), cat(o_len, np.uint32), cat(o_q, np.uint32), oo,
                cat(segs, np.uint64), bool(tr.value))

    # -- instrumentation -----------------------------------------------------------------------
    def counters(self, reset: bool = True) -> dict:
        c = (ctypes.c_uint64 * 8)()
        _check(self._L.moni_gpu_counters(self._h, c, 1 if reset else 0))
        names = ["bases", "jumps", "restarts", "restart_depth", "chars", "scan_rows", "general_steps", "descents"]
        return dict(zip(names, (int(x) for x in c)))

    def timing(self) -> dict:
        t = (ctypes.c_double * 6)()
        _check(self._L.moni_gpu_timing(self._h, t))
        return dict(zip(["h2d_ms", "pass1_ms", "pass2_ms", "compact_ms", "d2h_ms", "total_ms"], (float(x) for x in t)))

    def resident_batch(self, seqs: np.ndarray, offsets: np.ndarray, gpu_slot: int = 0) -> "ResidentBatch":
        return ResidentBatch(self, seqs, offsets, gpu_slot)


class ResidentBatch:
    """A read batch kept in one GPU's HBM; run() launches kernels only and returns their CUDA-event times."""

    def __init__(self, index: MoniIndex, seqs: np.ndarray, offsets: np.ndarray, gpu_slot: int = 0):
        self._L = index._L
        self.index = index
        seqs = np.ascontiguousarray(seqs, dtype=np.uint8)
        self.offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        self.n_reads = self.offsets.shape[0] - 1
        self.total = int(self.offsets[-1] - self.offsets[0])
        h = ctypes.c_void_p()
        _check(self._L.moni_gpu_batch_create(index._h, gpu_slot, _ptr(seqs), _u64(self.offsets), self.n_reads, ctypes.byref(h)))
        self._h = h

    def run(self, what: str) -> dict:
        ms = (ctypes.c_float * 3)()
        _check(self._L.moni_gpu_batch_run(self._h, {"ms": 1, "mems": 2}[what], ms))
        return {"pass1_ms": float(ms[0]), "pass2_ms": float(ms[1]), "compact_ms": float(ms[2])}

    def fetch_ms(self):
        ptr = np.empty(max(self.total, 1), dtype=np.uint64)
        ln = np.empty(max(self.total, 1), dtype=np.uint64)
        _check(self._L.moni_gpu_batch_fetch_ms(self._h, _u64(ptr), _u64(ln)))
        return ptr[:self.total], ln[:self.total]

    def fetch_mems(self):
        mo = np.empty(self.n_reads + 1, dtype=np.uint64)
        mp = np.empty(max(self.total, 1), dtype=np.uint64)
        ml = np.empty(max(self.total, 1), dtype=np.uint64)
        _check(self._L.moni_gpu_batch_fetch_mems(self._h, _u64(mo), _u64(mp), _u64(ml), mp.shape[0]))
        k = int(mo[self.n_reads])
        return mo, mp[:k], ml[:k]

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._L.moni_gpu_batch_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def format_ms(names: Iterable[bytes], offsets: np.ndarray, values: np.ndarray) -> bytes:
    """Text of a .pointers / .lengths file (reference src/matching_statistics.cpp:569-598)."""
    out = []
    for r, nm in enumerate(names):
        v = values[int(offsets[r]):int(offsets[r + 1])]
        out.append(b">" + nm + b"\n" + b"".join(b"%d " % int(x) for x in v) + b"\n")
    return b"".join(out)


def format_mems(names: Iterable[bytes], mem_offsets, mem_pos, mem_len) -> bytes:
    """Text of a .mems file (reference src/mems.cpp:585-587)."""
    out = []
    for r, nm in enumerate(names):
        a, b = int(mem_offsets[r]), int(mem_offsets[r + 1])
        out.append(b">" + nm + b"\n" + b"".join(b"(%d,%d) " % (int(mem_pos[t]), int(mem_len[t])) for t in range(a, b)) + b"\n")
    return b"".join(out)


def format_seeds(names: Iterable[bytes], seed_offsets, seed_ref, seed_len, seed_qpos) -> bytes:
    """Seeds as '(' pos ',' len ',' idx ') ' tokens per read (mem_t.pos, .len, .idx; include/aligner/mems.hpp:30-35)."""
    out = []
    for r, nm in enumerate(names):
        a, b = int(seed_offsets[r]), int(seed_offsets[r + 1])
        out.append(b">" + nm + b"\n" + b"".join(b"(%d,%d,%d) " % (int(seed_ref[t]), int(seed_len[t]), int(seed_qpos[t])) for t in range(a, b)) + b"\n")
    return b"".join(out)


def expand_runs(offsets: np.ndarray, run_offsets: np.ndarray, run_qpos: np.ndarray, run_len: np.ndarray, run_ptr: np.ndarray, out=None):
    """(pointers, lengths) uint64[offsets[-1]] from restart runs (moni_gpu_ms_runs_expand): the reference's two vectors."""
    offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
    run_offsets = np.ascontiguousarray(run_offsets, dtype=np.uint64)
    run_qpos = np.ascontiguousarray(run_qpos, dtype=np.uint32)
    run_len = np.ascontiguousarray(run_len, dtype=np.uint32)
    run_ptr = np.ascontiguousarray(run_ptr, dtype=np.uint64)
    total = int(offsets[-1])
    if out is None:
        ptr, ln = np.empty(max(total, 1), dtype=np.uint64), np.empty(max(total, 1), dtype=np.uint64)
    else:
        ptr, ln = out
    _check(load_library().moni_gpu_ms_runs_expand(_u64(offsets), offsets.shape[0] - 1, _u64(run_offsets), _ptr(run_qpos), _ptr(run_len),
                                                  _u64(run_ptr), _u64(ptr), _u64(ln)))
    return ptr[:total], ln[:total]


def unpack_runs(run_rec: np.ndarray, ptr_lo: np.ndarray, ptr_hi: np.ndarray):
    """(run_qpos, run_len, run_ptr) from the packed columns of moni_gpu_ms_runs_packed."""
    return (run_rec >> np.uint32(16)).astype(np.uint32), (run_rec & np.uint32(0xFFFF)).astype(np.uint32), unpack_pointers(ptr_lo, ptr_hi)


def unpack_pointers(ptr_lo: np.ndarray, ptr_hi: np.ndarray) -> np.ndarray:
    """u64 pointers from the two halves moni_gpu_ms_packed returns (values >= 2^40 - 2^31 are the reference's wrapped 2^64 - k)."""
    v = ptr_lo.astype(np.uint64) | (ptr_hi.astype(np.uint64) << np.uint64(32))
    wrapped = v >= np.uint64((1 << 40) - (1 << 31))
    with np.errstate(over="ignore"):
        v[wrapped] = v[wrapped] - np.uint64(1 << 40)
    return v
