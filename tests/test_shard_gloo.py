"""-m "not gpu": the N>1 path (clip sharding + final gather of logits and class counts) with 2 gloo ranks on CPU.
Each rank runs the oracle on its shard; the gathered result must equal the single-process result."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import common

ROOT = common.ROOT


def test_shard_range_covers_exactly():
    from edison_b200.shard import shard_range
    for n in (0, 1, 7, 8, 1000, 2**20):
        for world in (1, 2, 3, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, n_total, q):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import common as c
    from edison_b200.shard import shard_range, gather_logits, class_counts
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pcm = c.synth_frames(n_total * 31, seed=5).reshape(n_total, 31 * 1024)
    lo, hi = shard_range(n_total, rank, world)
    o = c.Oracle()
    lg, _, _ = o.kws(pcm[lo:hi].reshape(-1), hi - lo)
    lg = torch.from_numpy(lg)
    full = gather_logits(dist, lg, n_total, world)
    counts = class_counts(dist, lg)
    if rank == 0:
        q.put((full.numpy(), counts.numpy()))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gather_matches_single_process(oracle):
    n_total, world = 11, 2   # ragged: 6 + 5
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 300
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_total, q)) for r in range(world)]
    for p in procs: p.start()
    full, counts = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    pcm = common.synth_frames(n_total * 31, seed=5).reshape(-1)
    want, _, _ = oracle.kws(pcm, n_total)
    assert (full == want).all()
    assert (counts == np.bincount(want.astype(np.int64).argmax(axis=1), minlength=10)).all()
