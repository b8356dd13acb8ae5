#!/usr/bin/env python
"""Times ctr_bvp_pass (32768 configurations of system 0, no load, u_init = 0) for the library named by $CTR_B200_LIB."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'gym-ctr-reach_b200'))
import ctr_reach_envs
from ctr_reach_envs.envs import CtrReachVecEnv
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
env = CtrReachVecEnv(n, device=0, **ctr_reach_envs.default_kwargs())
env.reset()
q = env.buffer('joints').t().to(torch.float64).contiguous()
for _ in range(2):
    out = env.bvp_pass(q)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(5):
    out = env.bvp_pass(q)
b.record()
torch.cuda.synchronize()
ms = a.elapsed_time(b) / 5
st = out['steps'].double()
print("%-28s bvp_pass %d configs: %.3f ms = %.3e passes/s, mean steps %.1f max %d, tip checksum %.12f" % (
    os.path.basename(os.environ.get('CTR_B200_LIB', 'libctr_b200.so')), n, ms, n / ms * 1e3, st.mean().item(), int(st.max().item()),
    out['tip'].sum().item()))
g = st[: n // 10 * 10].view(-1, 10)
print("    lock-step: sum of per-warp max steps x 10 / sum of steps = %.3f" % ((g.max(dim=1).values.sum() * 10) / g.sum()).item())
env.close()
