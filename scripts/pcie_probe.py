#!/usr/bin/env python
"""Concurrent pinned-memory copy bandwidth per rank (no kernels): explains (or not) the e2e scaling of bench.py.
torchrun --nproc-per-node N scripts/pcie_probe.py [--bind]   -> one line per rank + aggregate."""
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (bind_to_gpu_numa_node)

rank, world, local = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1)), int(os.environ.get('LOCAL_RANK', 0))
bind = '--bind' in sys.argv
torch.cuda.set_device(local)
cpus = bench.bind_to_gpu_numa_node(local) if bind else None
if world > 1:
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
MB = 22                                        # the D2H payload of one bench step (22.3 MB), 3 MB H2D
d = torch.empty(MB << 20, dtype=torch.uint8, device='cuda')
h = torch.empty(MB << 20, dtype=torch.uint8).pin_memory()
d3 = torch.empty(3 << 20, dtype=torch.uint8, device='cuda')
h3 = torch.empty(3 << 20, dtype=torch.uint8).pin_memory()
h.fill_(1), h3.fill_(1)                        # first touch on this rank's CPUs


def timed(fn, reps=60):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps


t_d2h = timed(lambda: h.copy_(d, non_blocking=True))
t_h2d = timed(lambda: d3.copy_(h3, non_blocking=True))
t_both = timed(lambda: (d3.copy_(h3, non_blocking=True), h.copy_(d, non_blocking=True)))
res = torch.tensor([MB * 1.048576e-3 / t_d2h, 3 * 1.048576e-3 / t_h2d, (MB + 3) * 1.048576e-3 / t_both], dtype=torch.float64, device='cuda')
allr = [torch.zeros_like(res) for _ in range(world)]
if world > 1:
    dist.all_gather(allr, res)
else:
    allr = [res]
if rank == 0:
    print("world %d bind %s cpus_rank0 %s" % (world, bind, (len(cpus) if cpus else None)))
    for r, v in enumerate(allr):
        print("  rank %d: D2H 22 MB %.1f GB/s | H2D 3 MB %.1f GB/s | step payload (both) %.1f GB/s" % (r, *v.tolist()))
    tot = torch.stack(allr).sum(0).tolist()
    print("  aggregate: D2H %.1f GB/s, H2D %.1f GB/s, both %.1f GB/s -> PCIe-only ceiling of e2e = %.3e env-steps/s"
          % (tot[0], tot[1], tot[2], tot[2] * 1e9 / (25.4e6 / 131072)))
if world > 1:
    dist.destroy_process_group()
