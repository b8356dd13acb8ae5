#!/usr/bin/env python
"""Three ctr_fk_shape calls on the joints of 131072 stepped environments (driver for an ncu capture of k_fk_shape)."""
import os
import sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'gym-ctr-reach_b200'))
import ctr_reach_envs
from ctr_reach_envs.envs import CtrReachVecEnv
N = 131072
e = CtrReachVecEnv(N, device=0, **ctr_reach_envs.default_kwargs())
e.reset()
q = e.buffer('joints').to(torch.float64).t().contiguous()
for _ in range(3):
    r, npts, tidx, tip = e.backbone_shape(q)
torch.cuda.synchronize()
print("samples per config %.1f" % float(npts.double().mean()))
