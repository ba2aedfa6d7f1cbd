"""TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT.

Python side of the CPU oracle: an independent numpy reader of the reference's index files
(`<idx>.thrbv.ms`, `<idx>.plain.slp`; layouts: SURVEY.md Appendix B, i.e. the serialize() methods at
/root/reference include/ms/ms_pointers.hpp:279-297, thirdparty/r-index/internal/rle_string.hpp:347-359,
thirdparty/sdsl-lite/include/sdsl/{int_vector,sd_vector,select_support_mcl,wt_pc,wt_helper}.hpp,
thirdparty/ShapedSlp/PlainSlp.hpp:322-335) plus a ctypes binding of oracle/ms_oracle.c, the C
restatement of the per-read algorithm.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module.  The product (moni_align_b200) never does.

Pinned: tests/test_oracle_golden.py checks this oracle byte-for-byte against outputs of the UNMODIFIED
reference binaries (oracle/_ref/bin/ms, mems) committed under tests/golden/.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(HERE, "_build", "libms_oracle.so")
_SRC = os.path.join(HERE, "ms_oracle.c")


def build(force: bool = False) -> str:
    """gcc -O2 the C restatement into oracle/_build/libms_oracle.so."""
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(_SRC):
        os.makedirs(os.path.dirname(_SO), exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-std=c99", "-shared", "-fPIC", "-o", _SO, _SRC])
    return _SO


# ----------------------------------------------------------------------------- file decoding
class _Cur:
    def __init__(self, path: str):
        self.buf = np.fromfile(path, dtype=np.uint8)
        self.off = 0
        self.path = path

    def u64(self) -> int:
        v = int.from_bytes(self.buf[self.off:self.off + 8].tobytes(), "little")
        self.off += 8
        return v

    def u32(self) -> int:
        v = int.from_bytes(self.buf[self.off:self.off + 4].tobytes(), "little")
        self.off += 4
        return v

    def u8(self) -> int:
        v = int(self.buf[self.off])
        self.off += 1
        return v

    def raw(self, k: int) -> np.ndarray:
        v = self.buf[self.off:self.off + k]
        if v.shape[0] != k:
            raise ValueError(f"{self.path}: truncated")
        self.off += k
        return v

    def words(self, k: int) -> np.ndarray:
        return np.frombuffer(self.raw(8 * k).tobytes(), dtype="<u8")


def _int_vector(c: _Cur, fixed_width: int = 0):
    bits = c.u64()
    width = fixed_width if fixed_width else c.u8()
    return bits, width, c.words((bits + 63) >> 6)


def _unpack(bits: int, width: int, words: np.ndarray) -> np.ndarray:
    m = bits // width if width else 0
    if m == 0:
        return np.zeros(0, dtype=np.uint64)
    if width == 64:
        return words[:m].astype(np.uint64)
    w = np.concatenate([words, np.zeros(1, dtype=np.uint64)])
    b = np.arange(m, dtype=np.uint64) * np.uint64(width)
    wi = (b >> np.uint64(6)).astype(np.int64)
    sh = b & np.uint64(63)
    lo = w[wi] >> sh
    hi = np.where(sh == 0, np.uint64(0), w[wi + 1] << ((np.uint64(64) - sh) & np.uint64(63)))
    return (lo | hi) & np.uint64((1 << width) - 1)


def _bit_positions(bits: int, words: np.ndarray) -> np.ndarray:
    b = np.unpackbits(words.view(np.uint8), bitorder="little")[:bits]
    return np.flatnonzero(b).astype(np.uint64)


def _skip_select(c: _Cur) -> None:
    arg_cnt = c.u64()
    if not arg_cnt:
        return
    _int_vector(c)
    _int_vector(c, 1)
    for _ in range((arg_cnt + 4095) >> 12):
        _int_vector(c)


def _sd_vector(c: _Cur) -> np.ndarray:
    c.u64()                      # size
    wl = c.u8()
    lb, lw, lwords = _int_vector(c)
    hb, _, hwords = _int_vector(c, 1)
    _skip_select(c)
    _skip_select(c)
    low = _unpack(lb, lw, lwords)
    ones = _bit_positions(hb, hwords)[: low.shape[0]]
    idx = np.arange(low.shape[0], dtype=np.uint64)
    return ((ones - idx) << np.uint64(wl)) | low


def _sparse_sd(c: _Cur):
    u = c.u64()
    if u == 0:
        return 0, np.zeros(0, dtype=np.uint64)
    return u, _sd_vector(c)


def _wt_huff(c: _Cur) -> np.ndarray:
    size = c.u64()
    c.u64()                      # sigma
    bvbits, _, bvwords = _int_vector(c, 1)
    _int_vector(c, 64)
    _skip_select(c)
    _skip_select(c)
    n_nodes = c.u64()
    nodes = []
    for _ in range(n_nodes):
        raw = c.raw(22).tobytes()
        nodes.append((int.from_bytes(raw[0:8], "little"), int.from_bytes(raw[18:20], "little"),
                      int.from_bytes(raw[20:22], "little")))
    c_to_leaf = np.frombuffer(c.raw(512).tobytes(), dtype="<u2")
    c.raw(256 * 8)
    leaf_sym = {int(l): s for s, l in enumerate(c_to_leaf) if l != 0xFFFF}
    bv = np.unpackbits(bvwords.view(np.uint8), bitorder="little")[:bvbits]

    def decode(v: int, count: int) -> np.ndarray:
        pos, c0, c1 = nodes[v]
        if c0 == 0xFFFF:
            return np.full(count, leaf_sym[v], dtype=np.uint8)
        bits = bv[pos:pos + count]
        n1 = int(bits.sum())
        out = np.empty(count, dtype=np.uint8)
        out[bits == 0] = decode(c0, count - n1)
        out[bits == 1] = decode(c1, n1)
        return out

    return decode(0, size) if size else np.zeros(0, dtype=np.uint8)


def read_thrbv_ms(path: str, full: bool = False) -> dict:
    """`full`: the `moni align` index (.thrbv.full.ms, include/ms/moni.hpp:359-382; .thrbv.full.lcp.ms adds a trailing
    slcp vector, include/aligner/moni_lcp.hpp:208-225): same members with the Phi structures in between."""
    c = _Cur(path)
    d: dict = {"terminator_position": c.u64()}
    if c.u64() != 256:
        raise ValueError("F size")
    d["F"] = c.words(256).astype(np.uint64)
    n, R, B = c.u64(), c.u64(), c.u64()
    d.update(n=n, R=R, B=B)
    _, d["run_marks"] = _sparse_sd(c)
    per_letter = []
    count = np.zeros(256, dtype=np.uint64)
    for s in range(256):
        u, ones = _sparse_sd(c)
        count[s] = u
        per_letter.append(ones)
    heads = _wt_huff(c)
    lens = np.zeros(R, dtype=np.uint64)
    for s in range(256):
        if per_letter[s].shape[0]:
            ends = per_letter[s].astype(np.int64)
            lens[heads == s] = np.diff(np.concatenate([[-1], ends])).astype(np.uint64)
    starts = np.zeros(R + 1, dtype=np.uint64)
    np.cumsum(lens, out=starts[1:])
    if int(starts[-1]) != n:
        raise ValueError("run lengths do not sum to n")
    d.update(heads=heads, lens=lens, starts=starts, count=count)
    if full:
        d["pred"] = _sparse_sd(c)[1]
        d["pred_to_run"] = _unpack(*_int_vector(c)).astype(np.uint64)
    d["samples_last"] = _unpack(*_int_vector(c)).astype(np.uint64)
    thr = []
    for s in range(256):
        thr.append(_sparse_sd(c)[1])
    d["thr"] = thr
    if full:
        d["pred_start"] = _sparse_sd(c)[1]
        d["pred_start_to_run"] = _unpack(*_int_vector(c)).astype(np.uint64)
    d["samples_start"] = _unpack(*_int_vector(c)).astype(np.uint64)
    if full and c.off != c.buf.shape[0]:
        d["slcp"] = _unpack(*_int_vector(c)).astype(np.uint64)
    if c.off != c.buf.shape[0]:
        raise ValueError(f"{path}: {c.buf.shape[0] - c.off} trailing bytes")
    return d


def read_plain_slp(path: str) -> dict:
    c = _Cur(path)
    d: dict = {"text_len": c.u64(), "num_rules": c.u64(), "start_var": c.u32(), "alph_size": c.u64()}
    d["alph"] = c.raw(d["alph_size"]).copy()
    d["left"] = _unpack(*_int_vector(c)).astype(np.uint32)
    d["right"] = _unpack(*_int_vector(c)).astype(np.uint32)
    d["explen"] = _unpack(*_int_vector(c)).astype(np.uint64)
    if c.off != c.buf.shape[0]:
        raise ValueError(f"{path}: trailing bytes")
    return d


# ----------------------------------------------------------------------------- ctypes binding
def _p(a: np.ndarray, t):
    return a.ctypes.data_as(ctypes.POINTER(t))


class OracleIndex:
    """The CPU oracle over one index prefix (`prefix.thrbv.ms`, `prefix.plain.slp`)."""

    def __init__(self, prefix: str):
        self.lib = ctypes.CDLL(build())
        L = self.lib
        L.oracle_new.restype = ctypes.c_void_p
        u64p, u8p, u32p = ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint8), ctypes.POINTER(ctypes.c_uint32)
        L.oracle_free.argtypes = [ctypes.c_void_p]
        L.oracle_set_bwt.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64, u64p, u8p, u64p, u64p, u64p, u64p, u64p]
        L.oracle_set_slp.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_uint64, u8p, u32p, u32p, u64p]
        L.oracle_ms_batch.argtypes = [ctypes.c_void_p, u8p, u64p, ctypes.c_uint64, u64p, u64p]
        L.oracle_mems_batch.argtypes = [ctypes.c_void_p, u8p, u64p, ctypes.c_uint64, u64p, u64p, u64p]
        L.oracle_counters.argtypes = [ctypes.c_void_p, u64p, ctypes.c_int]
        L.oracle_char_at.argtypes = [ctypes.c_void_p, ctypes.c_uint64]
        L.oracle_char_at.restype = ctypes.c_uint8
        L.oracle_seeds_batch.argtypes = [ctypes.c_void_p, u8p, u64p, ctypes.c_uint64, ctypes.c_uint64, u64p, u64p, u64p, u64p]
        L.oracle_set_phi.argtypes = [ctypes.c_uint64, u64p, u64p, u64p, u64p]
        L.oracle_find_mem_occs.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64, u64p, ctypes.c_uint64]
        L.oracle_find_mem_occs.restype = ctypes.c_uint64
        L.oracle_find_mem_occs3.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint64, u64p, ctypes.c_uint64, u64p]
        L.oracle_find_mem_occs3.restype = ctypes.c_uint64
        import os as _os
        ext = next((e for e in (".thrbv.ms", ".thrbv.full.ms", ".thrbv.full.lcp.ms") if _os.path.exists(prefix + e)), ".thrbv.ms")
        self.bwt = read_thrbv_ms(prefix + ext, full=ext != ".thrbv.ms")
        self.slp = read_plain_slp(prefix + ".plain.slp")
        b, s = self.bwt, self.slp
        self._keep = dict(
            F=np.ascontiguousarray(b["F"], dtype=np.uint64), heads=np.ascontiguousarray(b["heads"], dtype=np.uint8),
            starts=np.ascontiguousarray(b["starts"], dtype=np.uint64),
            ss=np.ascontiguousarray(b["samples_start"], dtype=np.uint64), sl=np.ascontiguousarray(b["samples_last"], dtype=np.uint64),
            thr_vals=np.ascontiguousarray(np.concatenate(b["thr"] + [np.zeros(1, dtype=np.uint64)]), dtype=np.uint64),
            thr_off=np.zeros(257, dtype=np.uint64),
            alph=np.ascontiguousarray(s["alph"], dtype=np.uint8), left=np.ascontiguousarray(s["left"], dtype=np.uint32),
            right=np.ascontiguousarray(s["right"], dtype=np.uint32), explen=np.ascontiguousarray(s["explen"], dtype=np.uint64))
        k = self._keep
        np.cumsum(np.array([t.shape[0] for t in b["thr"]], dtype=np.uint64), out=k["thr_off"][1:])
        self.h = ctypes.c_void_p(L.oracle_new())
        L.oracle_set_bwt(self.h, b["n"], b["R"], _p(k["F"], ctypes.c_uint64), _p(k["heads"], ctypes.c_uint8),
                         _p(k["starts"], ctypes.c_uint64), _p(k["ss"], ctypes.c_uint64), _p(k["sl"], ctypes.c_uint64),
                         _p(k["thr_vals"], ctypes.c_uint64), _p(k["thr_off"], ctypes.c_uint64))
        L.oracle_set_slp(self.h, s["text_len"], s["num_rules"], s["start_var"], s["alph_size"], _p(k["alph"], ctypes.c_uint8),
                         _p(k["left"], ctypes.c_uint32), _p(k["right"], ctypes.c_uint32), _p(k["explen"], ctypes.c_uint64))

    def __del__(self):
        try:
            self.lib.oracle_free(self.h)
        except Exception:
            pass

    @property
    def n(self) -> int:
        return self.bwt["n"]

    @property
    def text_len(self) -> int:
        return self.slp["text_len"]

    def ms(self, seqs: np.ndarray, offsets: np.ndarray):
        seqs = np.ascontiguousarray(seqs, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        total = int(offsets[-1])
        ptr = np.zeros(max(total, 1), dtype=np.uint64)
        ln = np.zeros(max(total, 1), dtype=np.uint64)
        sq = seqs if seqs.shape[0] else np.zeros(1, dtype=np.uint8)
        self.lib.oracle_ms_batch(self.h, _p(sq, ctypes.c_uint8), _p(offsets, ctypes.c_uint64), offsets.shape[0] - 1,
                                 _p(ptr, ctypes.c_uint64), _p(ln, ctypes.c_uint64))
        return ptr[:total], ln[:total]

    def mems(self, seqs: np.ndarray, offsets: np.ndarray):
        seqs = np.ascontiguousarray(seqs, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        total = int(offsets[-1])
        nr = offsets.shape[0] - 1
        mo = np.zeros(nr + 1, dtype=np.uint64)
        mp = np.zeros(max(total, 1), dtype=np.uint64)
        ml = np.zeros(max(total, 1), dtype=np.uint64)
        sq = seqs if seqs.shape[0] else np.zeros(1, dtype=np.uint8)
        self.lib.oracle_mems_batch(self.h, _p(sq, ctypes.c_uint8), _p(offsets, ctypes.c_uint64), nr, _p(mo, ctypes.c_uint64),
                                   _p(mp, ctypes.c_uint64), _p(ml, ctypes.c_uint64))
        k = int(mo[-1])
        return mo, mp[:k], ml[:k]

    def seeds(self, seqs: np.ndarray, offsets: np.ndarray, min_len: int = 0):
        """(seed_offsets, seed_ref, seed_qpos, seed_len): seed_finder::find_mems for every read."""
        seqs = np.ascontiguousarray(seqs, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        total = int(offsets[-1])
        nr = offsets.shape[0] - 1
        so = np.zeros(nr + 1, dtype=np.uint64)
        cols = [np.zeros(max(total, 1), dtype=np.uint64) for _ in range(3)]
        sq = seqs if seqs.shape[0] else np.zeros(1, dtype=np.uint8)
        self.lib.oracle_seeds_batch(self.h, _p(sq, ctypes.c_uint8), _p(offsets, ctypes.c_uint64), nr, int(min_len),
                                    _p(so, ctypes.c_uint64), *[_p(c, ctypes.c_uint64) for c in cols])
        k = int(so[-1])
        return so, cols[0][:k], cols[1][:k], cols[2][:k]

    def mem_occs(self, seed_ref, seed_len):
        """(occ_offsets, occs): seed_finder::find_MEM_occs (unfiltered) for every seed; needs the `moni align` index."""
        self._set_phi()
        off, out = [0], []
        buf = np.zeros(1 << 16, dtype=np.uint64)
        for p, l in zip(seed_ref, seed_len):
            c = int(self.lib.oracle_find_mem_occs(self.h, int(p), int(l), _p(buf, ctypes.c_uint64), buf.shape[0]))
            if c > buf.shape[0]:
                buf = np.zeros(c, dtype=np.uint64)
                c = int(self.lib.oracle_find_mem_occs(self.h, int(p), int(l), _p(buf, ctypes.c_uint64), buf.shape[0]))
            out.append(buf[:c].copy())
            off.append(off[-1] + c)
        return np.array(off, dtype=np.uint64), (np.concatenate(out) if out else np.zeros(0, dtype=np.uint64))

    def _set_phi(self):
        b = self.bwt
        if "pred" not in b:
            raise ValueError("the oracle index was not loaded from a .thrbv.full.ms file")
        k = self._keep
        for name in ("pred", "pred_to_run", "pred_start", "pred_start_to_run"):
            k[name] = np.ascontiguousarray(b[name], dtype=np.uint64)
        self.lib.oracle_set_phi(b["R"], *[_p(k[nm], ctypes.c_uint64) for nm in ("pred", "pred_to_run", "pred_start", "pred_start_to_run")])

    def _occs3(self, head, up, down, ln):
        buf = np.zeros(1 << 16, dtype=np.uint64)
        na = ctypes.c_uint64(0)
        c = int(self.lib.oracle_find_mem_occs3(self.h, int(head), int(up), int(down), int(ln), _p(buf, ctypes.c_uint64), buf.shape[0], ctypes.byref(na)))
        if c > buf.shape[0]:
            buf = np.zeros(c, dtype=np.uint64)
            c = int(self.lib.oracle_find_mem_occs3(self.h, int(head), int(up), int(down), int(ln), _p(buf, ctypes.c_uint64), buf.shape[0], ctypes.byref(na)))
        return buf[:c].copy(), int(na.value)

    def populate_seeds(self, seed_offsets, seed_ref, seed_len, seed_qpos, min_len: int):
        """seed_finder::populate_seeds with seed filtering off (include/aligner/seed_finder.hpp:311-319 -> populate_seed :258-309), per read:
        every MEM gets its find_MEM_occs list; a MEM with len >= 2 * min_len appends two more seeds to the read's vector -- the left
        half (upper_suffix, len >> 1, idx) whose list is upper_suffix + find_MEM_above(upper_suffix) + find_MEM_below(lower_suffix) at
        the shorter length (:277-291; upper / lower_suffix = occs.back() after the walk above / below, :272-275), and the right half
        (pos + ll, len - ll, idx + ll) with its own find_MEM_occs (:293-301).  Returns (offsets, ref, len, qpos, occ_offsets, occs)
        of the final vectors, in the reference's order (the read's MEMs, then the halves in MEM order)."""
        self._set_phi()
        off, ref, ln, qp, ooff, occs = [0], [], [], [], [0], []
        for r in range(len(seed_offsets) - 1):
            a, b = int(seed_offsets[r]), int(seed_offsets[r + 1])
            tail = []
            for t in range(a, b):
                pos, l, i = int(seed_ref[t]), int(seed_len[t]), int(seed_qpos[t])
                lst, na = self._occs3(pos, pos, pos, l)
                ref.append(pos); ln.append(l); qp.append(i); occs.append(lst); ooff.append(ooff[-1] + len(lst))
                if l >= (min_len << 1):
                    upper, lower = int(lst[na]), int(lst[-1])
                    ll = l >> 1
                    tail.append((upper, ll, i, self._occs3(upper, upper, lower, ll)[0]))
                    tail.append((pos + ll, l - ll, i + ll, self._occs3(pos + ll, pos + ll, pos + ll, l - ll)[0]))
            for p_, l_, i_, lst in tail:
                ref.append(p_); ln.append(l_); qp.append(i_); occs.append(lst); ooff.append(ooff[-1] + len(lst))
            off.append(len(ref))
        u = np.uint64
        return (np.array(off, dtype=u), np.array(ref, dtype=u), np.array(ln, dtype=u), np.array(qp, dtype=u), np.array(ooff, dtype=u),
                np.concatenate(occs) if occs else np.zeros(0, dtype=u))

    def counters(self, reset: bool = True) -> dict:
        out = np.zeros(5, dtype=np.uint64)
        self.lib.oracle_counters(self.h, _p(out, ctypes.c_uint64), 1 if reset else 0)
        return dict(zip(["bases", "jumps", "restarts", "restart_depth", "chars"], (int(x) for x in out)))

    def char_at(self, pos: int) -> int:
        return int(self.lib.oracle_char_at(self.h, pos))


# ----------------------------------------------------------------------------- text formats
def format_ms(names, offsets, values) -> bytes:
    """The reference's .pointers / .lengths text: '>' name '\\n' then each value + ' ', then '\\n'
    (src/matching_statistics.cpp:569-598)."""
    out = []
    for r, nm in enumerate(names):
        v = values[int(offsets[r]):int(offsets[r + 1])]
        out.append(b">" + nm + b"\n" + b"".join(b"%d " % int(x) for x in v) + b"\n")
    return b"".join(out)


def format_mems(names, mem_offsets, mem_pos, mem_len) -> bytes:
    """The reference's .mems text: '(' i ',' len ') ' tokens (src/mems.cpp:585-587)."""
    out = []
    for r, nm in enumerate(names):
        a, b = int(mem_offsets[r]), int(mem_offsets[r + 1])
        out.append(b">" + nm + b"\n" + b"".join(b"(%d,%d) " % (int(mem_pos[t]), int(mem_len[t])) for t in range(a, b)) + b"\n")
    return b"".join(out)


def format_seeds(names, seed_offsets, seed_ref, seed_len, seed_qpos) -> bytes:
    """The text oracle/ref_seeds_harness.cpp prints for the reference's find_mems: '(' pos ',' len ',' idx ') ' tokens."""
    out = []
    for r, nm in enumerate(names):
        a, b = int(seed_offsets[r]), int(seed_offsets[r + 1])
        out.append(b">" + nm + b"\n" + b"".join(b"(%d,%d,%d) " % (int(seed_ref[t]), int(seed_len[t]), int(seed_qpos[t])) for t in range(a, b)) + b"\n")
    return b"".join(out)


def format_occs(names, seed_offsets, seed_ref, seed_len, seed_qpos, occ_offsets, occs) -> bytes:
    """The text `ref_seeds ... occs` prints: '(' pos ',' len ',' idx ':' occ ',' occ ... ') ' tokens."""
    out = []
    for r, nm in enumerate(names):
        a, b = int(seed_offsets[r]), int(seed_offsets[r + 1])
        toks = []
        for t in range(a, b):
            oc = occs[int(occ_offsets[t]):int(occ_offsets[t + 1])]
            toks.append(b"(%d,%d,%d:" % (int(seed_ref[t]), int(seed_len[t]), int(seed_qpos[t])) + b",".join(b"%d" % int(x) for x in oc) + b") ")
        out.append(b">" + nm + b"\n" + b"".join(toks) + b"\n")
    return b"".join(out)
