#!/usr/bin/env python
"""ctr_fk_shape against the C oracle on random configurations after one random step (what the facade test draws): finds corner cases."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'gym-ctr-reach_b200'))
sys.path.insert(0, os.path.join(ROOT, 'oracle'))
import ctr_reach_envs
from ctr_reach_envs.envs import CtrReachVecEnv
import ctr_oracle as O, ctr_oracle_c as C
C.build()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
seed = int(sys.argv[2]) if len(sys.argv) > 2 else 0
tc = C.derive(O.DEFAULT_SYSTEMS)[0]
env = CtrReachVecEnv(n, device=0, seed=seed, **ctr_reach_envs.default_kwargs())
env.reset()
lim = torch.tensor(env.action_space.high, device='cuda:0')
g = torch.Generator(device='cuda:0'); g.manual_seed(seed)
env.step(((torch.rand((n, 6), device='cuda:0', generator=g) * 2 - 1) * lim).contiguous())
q = env.buffer('joints').t().to(torch.float64).contiguous()
out = env.backbone_shape(q)
shape, npts, tidx = out[0].cpu().numpy(), out[1].cpu().numpy(), out[2].cpu().numpy()
qn = q.cpu().numpy()
bad = 0
worst = 0.0
for i in range(n):
    r_ref, t_ref = C.fk_shape(tc, qn[i])
    if npts[i] != len(r_ref) or not np.array_equal(tidx[i], t_ref):
        bad += 1
        print("MISMATCH count/idx at", i, repr(qn[i].tolist()), npts[i], len(r_ref), tidx[i], t_ref)
        continue
    d = np.abs(shape[i, :npts[i]] - r_ref).max()
    worst = max(worst, d)
    if d > 1e-10:
        bad += 1
        print("MISMATCH values at", i, repr(qn[i].tolist()), d)
print("checked", n, "bad", bad, "worst", worst)
