"""moni_gpu_ms_noalloc (include/moni_gpu.h): the allocation-free variant of moni_gpu_ms -- device buffers in and out, scratch in
the library's fixed file-scope array.  Called through the C ABI and compared bit for bit (unsigned integers: tolerance 0) with
moni_gpu_ms and with the reference binaries' files, on calls in a row that reuse the array.  Collected last on purpose."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, pack, read_fastq

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def m():
    import moni_align_b200 as mod
    mod.load_library()
    return mod


def _noalloc(torch, ix, flat, offs):
    """Run the entry point on device copies of (flat, offs); returns (pointers, lengths) as numpy u64."""
    dev = torch.device("cuda", ix.gpus[0])
    total, nr = int(offs[-1]), offs.shape[0] - 1
    d_seqs = torch.from_numpy(flat.copy() if total else np.zeros(1, np.uint8)).to(dev)
    d_offs = torch.from_numpy(offs.astype(np.int64)).to(dev)
    d_ptr = torch.full((max(total, 1),), -7, dtype=torch.int64, device=dev)
    d_len = torch.full((max(total, 1),), -7, dtype=torch.int64, device=dev)
    torch.cuda.synchronize(dev)
    free0 = torch.cuda.mem_get_info(dev)[0]
    ix.matching_statistics_noalloc(d_seqs.data_ptr(), d_offs.data_ptr(), nr, total, d_ptr.data_ptr(), d_len.data_ptr())
    free1 = torch.cuda.mem_get_info(dev)[0]
    return d_ptr.cpu().numpy().view(np.uint64)[:total], d_len.cpu().numpy().view(np.uint64)[:total], free0 - free1


def test_noalloc_equals_ms_on_calls_in_a_row(m):
    import torch
    import synth
    L = m.load_library()
    assert hasattr(L, "moni_gpu_ms_noalloc")
    for case in ("withN", "pan4"):
        ix = m.MoniIndex(os.path.join(GOLDEN, case + ".fa"))
        names, seqs = read_fastq(os.path.join(GOLDEN, case + ".fq"))
        flat, offs = pack(seqs)
        want_p, want_l = ix.matching_statistics(flat, offs)
        want_p, want_l = want_p.copy(), want_l.copy()
        # a larger ragged batch first (with empty reads and arbitrary bytes), so the array holds other reads' leftovers when
        # the golden batch runs after it -- twice
        rng = np.random.default_rng(11)
        text = np.frombuffer(b"".join(seqs), dtype=np.uint8)
        reads = [r for r in synth.short_reads(text, 2500, 120, 5, err=0.05)]
        reads += [rng.integers(1, 127, size=int(rng.integers(0, 300)), dtype=np.uint8) for _ in range(300)]
        reads.insert(7, np.zeros(0, np.uint8))
        big_flat, big_offs = m.pack_reads(reads)
        bp, bl = ix.matching_statistics(big_flat, big_offs)
        gp, gl, _ = _noalloc(torch, ix, big_flat, big_offs)
        assert np.array_equal(gp, bp) and np.array_equal(gl, bl)
        for rep in range(2):
            gp, gl, grew = _noalloc(torch, ix, flat, offs)
            assert np.array_equal(gp, want_p), (case, rep)
            assert np.array_equal(gl, want_l), (case, rep)
            assert grew <= 0, "device memory in use grew by %d bytes across an allocation-free call" % grew
        # and the reference binaries' own files for that batch
        assert m.format_ms(names, offs, gp) == open(os.path.join(GOLDEN, case + ".pointers"), "rb").read()
        assert m.format_ms(names, offs, gl) == open(os.path.join(GOLDEN, case + ".lengths"), "rb").read()
        ix.close()


def test_noalloc_limits_and_argument_errors(m):
    import torch
    ix = m.MoniIndex(os.path.join(GOLDEN, "pan4.fa"))
    dev = torch.device("cuda", 0)
    one = torch.zeros(16, dtype=torch.int64, device=dev)
    with pytest.raises(m.MoniGpuError) as e:   # beyond the fixed array: refused before anything is read
        ix.matching_statistics_noalloc(one.data_ptr(), one.data_ptr(), 1, m.NOALLOC_MAX_BASES + 1, one.data_ptr(), one.data_ptr())
    assert e.value.code == -1 and "at most" in str(e.value)
    with pytest.raises(m.MoniGpuError) as e:
        ix.matching_statistics_noalloc(one.data_ptr(), one.data_ptr(), m.NOALLOC_MAX_READS + 1, 8, one.data_ptr(), one.data_ptr())
    assert e.value.code == -1 and "at most" in str(e.value)
    host = np.zeros(16, dtype=np.uint64)         # host memory where device memory is expected
    with pytest.raises(m.MoniGpuError) as e:
        ix.matching_statistics_noalloc(host.ctypes.data, host.ctypes.data, 1, 8, host.ctypes.data, host.ctypes.data)
    assert e.value.code == -1 and "device pointers" in str(e.value)
    bad = torch.tensor([0, 9, 4], dtype=torch.int64, device=dev)      # decreasing offsets
    seqs = torch.full((32,), 65, dtype=torch.uint8, device=dev)
    with pytest.raises(m.MoniGpuError) as e:
        ix.matching_statistics_noalloc(seqs.data_ptr(), bad.data_ptr(), 2, 4, one.data_ptr(), one.data_ptr())
    assert e.value.code == -1 and "offsets" in str(e.value)
    # zero reads / zero bases are fine and touch nothing
    z = torch.zeros(1, dtype=torch.int64, device=dev)
    ix.matching_statistics_noalloc(0, z.data_ptr(), 0, 0, 0, 0)
    # at the limit of reads: MAX_READS reads of 16 bases, against moni_gpu_ms
    rng = np.random.default_rng(3)
    flat = rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), size=m.NOALLOC_MAX_READS * 16)
    offs = np.arange(m.NOALLOC_MAX_READS + 1, dtype=np.uint64) * np.uint64(16)
    wp, wl = ix.matching_statistics(flat, offs)
    gp, gl, _ = _noalloc(torch, ix, flat, offs)
    assert np.array_equal(gp, wp) and np.array_equal(gl, wl)
    ix.close()
