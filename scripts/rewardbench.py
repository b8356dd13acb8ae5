#!/usr/bin/env python
"""k_reward (compute_reward + obs[9:12] rewrite, 80 B / transition) on 2^20 transitions, L2 flushed before every call, for the
library named by $CTR_B200_LIB."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'gym-ctr-reach_b200'))
import ctr_reach_envs
from ctr_reach_envs.envs import CtrReachVecEnv
dev = torch.device('cuda', 0)
e = CtrReachVecEnv(16, device=0, **ctr_reach_envs.default_kwargs())
m = 1 << 20
a = torch.rand((3, m), dtype=torch.float64, device=dev)
d = a[:, torch.randperm(m, device=dev)].contiguous()
out = torch.empty(m, dtype=torch.float64, device=dev)
obs = torch.zeros((13, m), dtype=torch.float64, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for _ in range(3):
    e.compute_reward_soa(a, d, out, obs=obs, tol=0.001)
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(30)]
for s, t in ev:
    flush.zero_()
    s.record(); e.compute_reward_soa(a, d, out, obs=obs, tol=0.001); t.record()
torch.cuda.synchronize()
us = sorted(s.elapsed_time(t) * 1e3 for s, t in ev)
print("%-24s k_reward (relabel) 2^20: median %.2f us (min %.2f) = %.0f GB/s, checksum %.1f" % (
    os.path.basename(os.environ.get('CTR_B200_LIB', 'libctr_b200.so')), us[15], us[0], 80.0 * m / us[15] / 1e3, out.sum().item()))
e.close()
