"""CPU suite, part 3: the N>1 path.  World-size-2 gloo run of the sharding + in-order gather + max-over-ranks
plumbing that bench.py uses under torchrun; the per-shard compute is done by the oracle here (no GPU)."""
import os
import socket

import numpy as np
import pytest

from conftest import GOLDEN, pack, read_fastq


def test_shard_bounds_cover_and_balance():
    from moni_align_b200.sharding import shard_bounds
    rng = np.random.default_rng(3)
    lens = rng.integers(0, 3000, size=1000).astype(np.uint64)
    offs = np.zeros(1001, dtype=np.uint64)
    np.cumsum(lens, out=offs[1:])
    for parts in (1, 2, 3, 8):
        cut = shard_bounds(offs, parts)
        assert cut[0] == 0 and cut[-1] == 1000 and (np.diff(cut) >= 0).all()
        bases = np.diff(offs[cut].astype(np.int64))
        assert bases.max() - bases.min() <= 2 * 3000 + parts
    z = np.zeros(1, dtype=np.uint64)
    assert shard_bounds(z, 4).tolist() == [0, 0, 0, 0, 0]


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from moni_align_b200.sharding import gather_in_order, max_over_ranks, shard_of, sum_over_ranks
    from oracle import oracle
    pre = os.path.join(GOLDEN, "pan4")
    names, seqs = read_fastq(pre + ".fq")
    flat, offs = pack(seqs)
    s, o, (a, b) = shard_of(flat, offs, rank, world)
    ix = oracle.OracleIndex(pre + ".fa")
    ptr, ln = ix.ms(s, o)
    dist.barrier()
    go, gp = gather_in_order(o, ptr, world, rank)
    _, gl = gather_in_order(o, ln, world, rank)
    t = max_over_ranks(1.0 + rank)
    tot = sum_over_ranks(float(b - a))
    if rank == 0:
        ok = (oracle.format_ms(names, go, gp) == open(pre + ".pointers", "rb").read()
              and oracle.format_ms(names, go, gl) == open(pre + ".lengths", "rb").read())
        q.put((ok, t, tot, len(names)))
    dist.destroy_process_group()


def test_two_rank_gloo_shard_gather_in_input_order():
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok, t, tot, n = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok
    assert t == 2.0
    assert tot == n
