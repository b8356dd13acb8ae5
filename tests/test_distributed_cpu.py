"""world_size-2 gloo test of the only collective on the path: the all-reduce of the episode-statistics partial sums,
plus the env-index sharding that keeps the step path collective-free."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, os.path.join(ROOT, 'gym-ctr-reach_b200'))
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    import torch
    import torch.distributed as dist
    from ctr_reach_envs.distributed import all_reduce_stats, init_process_group, shard_range
    r, lr, w = init_process_group(backend='gloo')
    lo, hi = shard_range(1001, r, w)
    # each rank reports partial sums over its own env shard: [episodes, successes, sum_error, ...]
    v = torch.tensor([hi - lo, (hi - lo) // 2, 0.5 * (hi - lo), 0, 0, 10 * (hi - lo), 0, 0], dtype=torch.float64)
    out = all_reduce_stats(v)
    q.put((r, lo, hi, out.tolist()))
    dist.barrier()
    dist.destroy_process_group()


def test_stats_all_reduce_world2_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, lo0, hi0, s0), (r1, lo1, hi1, s1) = res
    assert (lo0, hi0, lo1, hi1) == (0, 501, 501, 1001)
    assert s0 == s1                                   # every rank holds the global sums
    assert s0[0] == 1001 and s0[1] == 250 + 250 and s0[2] == 500.5 and s0[5] == 10010


def test_all_reduce_is_identity_without_process_group():
    sys.path.insert(0, os.path.join(ROOT, 'gym-ctr-reach_b200'))
    import torch
    from ctr_reach_envs.distributed import all_reduce_stats
    v = torch.arange(8, dtype=torch.float64)
    assert torch.equal(all_reduce_stats(v), v)
