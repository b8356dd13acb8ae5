"""Multi-GPU plumbing: one process per GPU, environments partitioned by contiguous index ranges, NO collective on
the step path.  The only exchange is the all-reduce(SUM) of the 8 episode-statistics partial sums."""
import os

import torch


def shard_range(total_envs, rank, world_size):
    """Contiguous [lo, hi) env-index range of `rank`; sizes differ by at most one."""
    base, rem = divmod(int(total_envs), int(world_size))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def env_info():
    return int(os.environ.get('RANK', 0)), int(os.environ.get('LOCAL_RANK', 0)), int(os.environ.get('WORLD_SIZE', 1))


def init_process_group(backend=None):
    """Initialise torch.distributed from the torchrun environment (nccl on GPU boxes, gloo otherwise)."""
    import torch.distributed as dist
    rank, local_rank, world = env_info()
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = 'nccl' if torch.cuda.is_available() else 'gloo'
        kw = {}
        if backend == 'nccl':
            torch.cuda.set_device(local_rank)
            kw['device_id'] = torch.device('cuda', local_rank)
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    return rank, local_rank, world


def all_reduce_stats(v, device=None):
    """SUM-all-reduce a small float64 vector across ranks (identity when torch.distributed is not initialised)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return v
    if dist.get_backend() == 'nccl':
        w = v.to(device if device is not None else torch.device('cuda', torch.cuda.current_device()))
        dist.all_reduce(w, op=dist.ReduceOp.SUM)
        return w.cpu()
    w = v.clone()
    dist.all_reduce(w, op=dist.ReduceOp.SUM)
    return w


def make_sharded_vec_env(total_envs, rank=None, world_size=None, device=None, **kwargs):
    """CtrReachVecEnv over this rank's shard of a `total_envs`-wide job (default CTR-Reach-v0 kwargs + overrides)."""
    from . import default_kwargs
    from .envs import CtrReachVecEnv
    r, lr, w = env_info()
    rank = r if rank is None else rank
    world_size = w if world_size is None else world_size
    lo, hi = shard_range(total_envs, rank, world_size)
    kw = default_kwargs()
    kw.update(kwargs)
    return CtrReachVecEnv(hi - lo, device=lr if device is None else device, env_offset=lo, **kw)
