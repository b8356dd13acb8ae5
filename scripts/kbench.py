#!/usr/bin/env python
"""Kernel-variant micro-benchmark: times k_step alone (no auto-reset, L2 flushed) and the full step (auto-reset) for the
library named by $CTR_B200_LIB, and checks the tips against the C oracle.  usage: CTR_B200_LIB=... python scripts/kbench.py"""
import os
import sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'gym-ctr-reach_b200'))
sys.path.insert(0, os.path.join(ROOT, 'oracle'))
import ctr_reach_envs
from ctr_reach_envs.envs import CtrReachVecEnv
import ctr_oracle as O
import ctr_oracle_c as C

N = int(os.environ.get('KB_N', 131072))
steps = int(os.environ.get('KB_STEPS', 30))
prec = os.environ.get('KB_PREC', 'f64')
integ = os.environ.get('KB_INTEG', 'dp45')
dev = torch.device('cuda', 0)
kw = ctr_reach_envs.default_kwargs()
lim = None
res = {}
for auto in (False, True):
    env = CtrReachVecEnv(N, device=0, auto_reset=auto, precision=prec, integrator=integ, **kw)
    env.goal_tolerance.current_tol = 0.02 if auto else 1e-9
    env.reset()
    lim = torch.tensor(env.action_space.high, device=dev)
    g = torch.Generator(device=dev); g.manual_seed(1)
    pool = [((torch.rand((N, 6), device=dev, generator=g) * 2 - 1) * lim).contiguous() for _ in range(4)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for i in range(5):
        env.step(pool[i % 4])
    if not auto:
        env.set_state(t=np.zeros(N, dtype=np.int32))
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for i in range(steps):
        flush.zero_()
        ev[i][0].record(); env.step(pool[i % 4]); ev[i][1].record()
        if not auto and i % 20 == 19:
            env.set_state(t=np.zeros(N, dtype=np.int32))
    torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in ev)
    res['auto' if auto else 'kernel'] = (ms[len(ms) // 2], ms[0])
    if not auto:
        n = 4000
        q = env.buffer('joints').t()[:n].cpu().numpy()
        tip = env.forward_kinematics(q).cpu().numpy()
        ref = C.fk_batch(C.derive(O.DEFAULT_SYSTEMS[:1]), q, None, nthreads=8)
        res['err'] = float(np.linalg.norm(tip - ref['tip'], axis=1).max())
    env.close()
print("%-40s k_step median %.4f ms (min %.4f) = %.3e env-steps/s | step+reset %.4f ms = %.3e | max tip err %.2e" % (
    os.path.basename(os.environ.get('CTR_B200_LIB', 'libctr_b200.so')), res['kernel'][0], res['kernel'][1], N / res['kernel'][0] * 1e3,
    res['auto'][0], N / res['auto'][0] * 1e3, res['err']))
