#!/usr/bin/env python
"""Times ctr_jacobian (16384 configurations of system 0 at reset joints = 114688 kinematics solves) for the library named by $CTR_B200_LIB."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'gym-ctr-reach_b200'))
import ctr_reach_envs
from ctr_reach_envs.envs import CtrReachVecEnv
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
env = CtrReachVecEnv(n, device=0, **ctr_reach_envs.default_kwargs())
env.reset()
q = env.buffer('joints').t().to(torch.float64).contiguous()
for _ in range(3):
    J, tip = env.jacobian(q)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(20):
    J, tip = env.jacobian(q)
b.record()
torch.cuda.synchronize()
print("%-24s jacobian %d configs: %.1f us, J checksum %.12f tip checksum %.12f" % (
    os.path.basename(os.environ.get('CTR_B200_LIB', 'libctr_b200.so')), n, a.elapsed_time(b) / 20 * 1e3, J.sum().item(), tip.sum().item()))
env.close()
