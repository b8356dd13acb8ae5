#!/usr/bin/env python
"""Times ctr_fk_shape (131072 configurations of system 0 at reset joints, 270 points per env) for the library named by $CTR_B200_LIB."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'gym-ctr-reach_b200'))
import ctr_reach_envs
from ctr_reach_envs import _lib
from ctr_reach_envs.envs import CtrReachVecEnv
n = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
env = CtrReachVecEnv(n, device=0, **ctr_reach_envs.default_kwargs())
env.reset()
dev = env.device
q = env.buffer('joints').to(torch.float64).contiguous()
shape = torch.zeros((n, 270, 3), dtype=torch.float64, device=dev)
npts = torch.empty(n, dtype=torch.int32, device=dev)
tidx = torch.empty((3, n), dtype=torch.int32, device=dev)
tip = torch.empty((3, n), dtype=torch.float64, device=dev)
call = lambda: _lib.check(env.lib.ctr_fk_shape(env.handle, _lib.ptr(q), None, None, n, 270, _lib.ptr(shape), _lib.ptr(npts), _lib.ptr(tidx),
                                               _lib.ptr(tip), env._stream()))
for _ in range(2):
    call()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(10):
    call()
b.record()
torch.cuda.synchronize()
ms = a.elapsed_time(b) / 10
mask = (torch.arange(270, device=dev)[None, :] < npts[:, None])
cs = (shape * mask[:, :, None]).sum().item()
print("%-24s fk_shape %d configs: %.3f ms, mean points %.1f, checksum %.12f, tip checksum %.12f, tip_idx sum %d" % (
    os.path.basename(os.environ.get('CTR_B200_LIB', 'libctr_b200.so')), n, ms, npts.double().mean().item(), cs, tip.sum().item(),
    int(tidx.sum().item())))
env.close()
