"""CPU, world_size 2, gloo: the host-side sharding / global which.max logic of the N > 1 path."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from diceoptim_b200 import parallel


def test_shard_range_partitions_exactly():
    for N in (0, 1, 7, 64, 10_000_001):
        for world in (1, 2, 3, 8):
            r = [parallel.shard_range(N, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == N
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1


def test_merge_argmax_semantics():
    nan = float("nan")
    assert parallel.merge_argmax([(1.0, 5), (3.0, 9), (3.0, 2), (nan, 0)]) == (3.0, 2)   # first max, NaN skipped
    assert parallel.merge_argmax([(nan, -1), (0.5, -1)])[1] == -1
    assert parallel.merge_topk([(1.0, 4), (2.0, 8), (2.0, 3), (nan, 1), (0.0, -1)], 2) == [(2.0, 3), (2.0, 8)]


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, N, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    v = np.random.default_rng(3).standard_normal(N)
    v[[10, N - 5]] = v.max() + 1.0          # an exact tie across the two shards: the lower index must win
    v[3] = np.nan
    lo, hi = parallel.shard_range(N, rank, world)
    loc = v[lo:hi]
    ok = ~np.isnan(loc)
    li = int(np.flatnonzero(ok & (loc == loc[ok].max()))[0]) if ok.any() else -1
    val, gi = parallel.global_argmax(float(loc[li]) if li >= 0 else float("nan"), li, lo)
    q.put((rank, val, gi))
    dist.barrier()
    dist.destroy_process_group()


def test_global_argmax_world2_gloo():
    N, world = 1001, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, N, q)) for r in range(world)]
    for p in procs: p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs: p.join(timeout=60)
    v = np.random.default_rng(3).standard_normal(N); top = v.max() + 1.0
    assert all(r[2] == 10 and r[1] == top for r in res)
