"""world_size-2 gloo test (CPU) of the multi-GPU scheme: contiguous image shards [g*n/G, (g+1)*n/G), parameters replicated, no
collective on the data path; with faults every rank applies the faults scheduled before its shard first, then only the segments
inside it, and the gathered result equals the single-rank run.  The oracle stands in for the device here."""
import os
import sys

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def shard(n, G, g):
    return g * n // G, (g + 1) * n // G


def _worker(rank, world, port, n, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle_py import Oracle, synth_mnist
    data = synth_mnist(n, 77)
    rng = np.random.default_rng(5)
    o = Oracle("lfcW1A1")
    fl = np.array([[int(rng.integers(0, n)), 1 + int(rng.integers(0, 2)), l, 0, 1] for l in (0, 1, 2, 3) for _ in range(10)], np.uint32)
    for row in fl:
        bits = o.weight_layer_bits(int(row[2])) if row[1] == 1 else o.thresh_layer_bits(int(row[2]))
        row[3] = rng.integers(0, bits)
    fl = fl[np.argsort(fl[:, 0], kind="stable")]
    lo, hi = shard(n, world, rank)
    # faults scheduled before the shard are applied up front; the rest keep their time relative to the shard
    for row in fl[fl[:, 0] < lo]:
        o.inject(int(row[1]), int(row[2]), int(row[3]), 1)
    local = fl[(fl[:, 0] >= lo) & (fl[:, 0] < hi)].copy()
    local[:, 0] -= lo
    mine = o.classify_with_faults(data[lo:hi], local) if hi > lo else np.zeros(0, np.int32)
    parts = [None] * world
    dist.all_gather_object(parts, (lo, hi, mine.tolist()))
    t = torch.tensor([float(rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)  # the bench's max-over-ranks reduction
    if rank == 0:
        q.put((parts, float(t.item())))
    dist.destroy_process_group()


def test_two_rank_shards_reproduce_the_single_rank_run():
    from oracle_py import Oracle, synth_mnist
    n, world = 101, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    parts, tmax = q.get(timeout=120)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert tmax == float(world)
    bounds = sorted((lo, hi) for lo, hi, _ in parts)
    assert bounds[0][0] == 0 and bounds[-1][1] == n and all(bounds[i][1] == bounds[i + 1][0] for i in range(world - 1))
    got = np.concatenate([np.array(c, np.int32) for _, _, c in sorted(parts)])
    # single-rank reference with the same fault list
    data = synth_mnist(n, 77)
    rng = np.random.default_rng(5)
    o = Oracle("lfcW1A1")
    fl = np.array([[int(rng.integers(0, n)), 1 + int(rng.integers(0, 2)), l, 0, 1] for l in (0, 1, 2, 3) for _ in range(10)], np.uint32)
    for row in fl:
        bits = o.weight_layer_bits(int(row[2])) if row[1] == 1 else o.thresh_layer_bits(int(row[2]))
        row[3] = rng.integers(0, bits)
    fl = fl[np.argsort(fl[:, 0], kind="stable")]
    want = o.classify_with_faults(data, fl)
    assert np.array_equal(got, want)
