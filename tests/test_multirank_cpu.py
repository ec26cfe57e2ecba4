"""N > 1 path on the CPU: two gloo ranks each take their contiguous shard (tinyfk_b200.sharding), compute
it, and the gathered result equals the single-rank result; the whole-job rate uses the max over ranks.
The per-shard compute here is the CPU test harness (tests/emu) standing in for each rank's GPU."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tinyfk_b200 import sharding
from tinyfk_b200 import workloads as W

HERE = os.path.dirname(os.path.abspath(__file__))


def test_shard_bounds_cover_without_overlap():
    for n in (0, 1, 7, 128, 1000, 1 << 20):
        for world in (1, 2, 3, 4, 8):
            b = [sharding.shard_bounds(n, world, r) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1)) and all(lo <= hi for lo, hi in b)


def _worker(rank, world, port, n, out_path):
    sys.path.insert(0, HERE)
    from emuapi import EmuModel
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    e = EmuModel(W.urdf_text("pr2.urdf"))
    jids, lids, rot, base = W.setup(e, "cfg2")
    q = W.sample_q([(-1.0, 1.0)] * 7, base, n, seed=0)  # same seed on every rank: one logical batch
    lo, hi = sharding.shard_bounds(n, world, rank)
    P, J = e.solve_fk(q[lo:hi], lids, jids, rot, base, True)
    per = (n + world - 1) // world
    padP = torch.zeros((per, 6), dtype=torch.float64)
    padP[:hi - lo] = torch.from_numpy(np.ascontiguousarray(P))
    gathered = [torch.zeros_like(padP) for _ in range(world)]
    dist.all_gather(gathered, padP)

    def rmax(v):
        t = torch.tensor([v], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    def rsum(v):
        t = torch.tensor([v], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t)

    rate = sharding.aggregate_rate(hi - lo, 10.0 * (rank + 1), world, rmax, rsum)
    if rank == 0:
        full = torch.cat([g[:sharding.shard_bounds(n, world, r)[1] - sharding.shard_bounds(n, world, r)[0]]
                          for r, g in enumerate(gathered)]).numpy()
        Pf, _ = e.solve_fk(q, lids, jids, rot, base, True)
        np.save(out_path, np.array([float(np.array_equal(full, Pf)), rate]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_shards_reassemble(tmp_path):
    out = str(tmp_path / "res.npy")
    n, world = 1001, 2
    mp.spawn(_worker, args=(world, 29500 + os.getpid() % 2000, n, out), nprocs=world, join=True)
    ok, rate = np.load(out)
    assert ok == 1.0
    assert abs(rate - n / (20.0 / 1e3)) < 1e-6  # total items over the slowest rank (20 ms)
