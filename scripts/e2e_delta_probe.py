#!/usr/bin/env python
"""Per-call wall time of step_host in its three forms (full copy, delta, float32) in a steady-state window: looks for outliers."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'gym-ctr-reach_b200'))
import ctr_reach_envs
from ctr_reach_envs.envs import CtrReachVecEnv
N = 131072
env = CtrReachVecEnv(N, device=0, **ctr_reach_envs.default_kwargs())
env.reset()
env.set_state(t=(np.arange(N) % 150).astype(np.int32))
lim = env.action_limits
rng = np.random.RandomState(0)
acts = [torch.tensor(np.ascontiguousarray(rng.uniform(-lim, lim, size=(N, 6)).astype(np.float32).T)).pin_memory().numpy() for _ in range(4)]
out = env.alloc_host_outputs()
for mode in ('full', 'delta', 'full', 'delta'):
    ts = []
    for i in range(64):
        t0 = time.perf_counter()
        env.step_host(acts[i % 4], out, delta=(mode == 'delta'))
        ts.append((time.perf_counter() - t0) * 1e3)
    ts = np.array(ts[8:])
    print("%-6s mean %.3f ms  median %.3f  p90 %.3f  max %.3f  | done/step %.0f" % (mode, ts.mean(), np.median(ts), np.percentile(ts, 90), ts.max(), out['done'].sum()),
          np.round(ts[:16], 2))
env.close()
