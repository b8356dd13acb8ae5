#!/usr/bin/env python
"""Times ctr_reset of every env (131072 envs of system 0: goal sampling + 2 FK solves each) for the library named by $CTR_B200_LIB."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'gym-ctr-reach_b200'))
import ctr_reach_envs
from ctr_reach_envs.envs import CtrReachVecEnv
n = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
env = CtrReachVecEnv(n, device=0, **ctr_reach_envs.default_kwargs())
for _ in range(2):
    env.reset()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(5):
    env.reset()
b.record()
torch.cuda.synchronize()
print("%-24s reset %d envs: %.3f ms, desired checksum %.12f" % (os.path.basename(os.environ.get('CTR_B200_LIB', 'libctr_b200.so')), n,
      a.elapsed_time(b) / 5, env.buffer('desired_goal').sum().item()))
w = env.work_counters()
print("   ", {k: v for k, v in w.items()}, env.fault_counters())
env.close()
