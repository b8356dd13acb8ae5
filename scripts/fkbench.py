#!/usr/bin/env python
"""ctr_fk on 131072 reset configurations of system 0: time in the given order and in orders sorted by candidate keys (what a
work-sorted hand-out order would be worth for the one-shot FK; the env step sorts by last step's counter instead)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'gym-ctr-reach_b200'))
import ctr_reach_envs
from ctr_reach_envs.envs import CtrReachVecEnv
n = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
env = CtrReachVecEnv(n, device=0, **ctr_reach_envs.default_kwargs())
env.reset()
q = env.buffer('joints').t().to(torch.float64).contiguous()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
def timed(qq):
    for _ in range(2):
        env.forward_kinematics(qq)
    torch.cuda.synchronize()
    a.record()
    for _ in range(10):
        env.forward_kinematics(qq)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / 10 * 1e3
tip, R, info = env.forward_kinematics(q, return_info=True)
nfev = info[:, 0].double()
seg = env.segment_tables(q)
nsec = seg[3].double()
length = (seg[1] * (torch.arange(9, device=q.device)[None, :] < seg[3][:, None])).max(dim=1).values
print("given order            %.1f us (mean nfev %.1f)" % (timed(q), nfev.mean().item()))
coarse = -nsec * 16 - torch.floor(16 * length / 0.431).clamp(0, 15) + torch.rand(n, device=q.device, dtype=torch.float64) * 0.5
for name, key in (('coarse key, random ties', coarse), ('nfev desc', -nfev), ('nsec desc', -nsec), ('nsec desc, length desc', -nsec * 10 - length), ('length desc', -length)):
    print("%-22s %.1f us" % (name, timed(q[torch.argsort(key)].contiguous())))
env.close()
