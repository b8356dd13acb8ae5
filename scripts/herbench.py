#!/usr/bin/env python
"""k_her_sample / k_her_store micro-benchmark for the library named by $CTR_B200_LIB: 131072 envs, 16-slot ring, 2^20 samples,
L2 flushed before every call.  usage: CTR_B200_LIB=... python scripts/herbench.py"""
import os
import sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'gym-ctr-reach_b200'))
import ctr_reach_envs
from ctr_reach_envs.envs import CtrReachVecEnv
from ctr_reach_envs.her import HerReplay

N, cap, m = 131072, 16, 1 << 20
dev = torch.device('cuda', 0)
e = CtrReachVecEnv(N, device=0, **ctr_reach_envs.default_kwargs())
e.reset()
lim = torch.tensor(e.action_space.high, device=dev)
g = torch.Generator(device=dev); g.manual_seed(1)
pool = [((torch.rand((N, 6), device=dev, generator=g) * 2 - 1) * lim).contiguous() for _ in range(4)]
rep = HerReplay(e, cap, n_sampled_goal=4, seed=1)
rep.begin()
for i in range(cap + 2):
    e.step(pool[i % 4]); rep.store(pool[i % 4])
batch = rep.alloc_batch(m)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in ev:
        flush.zero_(); a.record(); fn(); b.record()
    torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in ev)
    return ms[len(ms) // 2] * 1e3


calls = iter(range(10000))
us_s = timed(lambda: rep.sample(m, out=batch, call=next(calls)))
us_t = timed(lambda: rep.store(pool[0]))
print("%-20s k_her_sample %.1f us (2^20 samples) | k_her_store %.1f us (131072 envs) | reward checksum %.1f"
      % (os.path.basename(os.environ.get('CTR_B200_LIB', 'libctr_b200.so')), us_s, us_t, float(batch['reward'].sum()) if isinstance(batch, dict) else 0.0))
