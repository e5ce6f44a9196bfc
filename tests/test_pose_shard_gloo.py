"""N>1 host logic on CPU: world_size-2 gloo processes broadcast a scene and shard a pose list."""
import os

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gaussiansplattingviewer_b200 import pose_shard, scenes


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = scenes.synthetic_gaussians(1000, seed=9) if rank == 0 else None
    ts = pose_shard.broadcast_scene(g, torch.device("cpu"), src=0)
    got = pose_shard.tensors_to_gaussians(ts)
    want = scenes.synthetic_gaussians(1000, seed=9)
    same = all(np.array_equal(getattr(got, f), getattr(want, f)) for f in ("xyz", "rot", "scale", "opacity", "sh"))
    lo, hi = pose_shard.shard_range(1024 + 1, rank, world)
    # max over ranks of a per-rank time, the way bench.py reduces its timings
    t = torch.tensor([float(rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    q.put((rank, same, lo, hi, float(t.item())))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_scene_broadcast_and_pose_sharding():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=120) for _ in ps)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [r[1] for r in res] == [True, True]
    assert (res[0][2], res[0][3], res[1][2], res[1][3]) == (0, 513, 513, 1025)
    assert res[0][4] == res[1][4] == 2.0


def test_shard_helpers_partition_exactly():
    for n in (0, 1, 7, 1024, 1025):
        for world in (1, 2, 3, 8):
            spans = [pose_shard.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
            rr = sorted(i for r in range(world) for i in pose_shard.shard_round_robin(n, r, world))
            assert rr == list(range(n))
