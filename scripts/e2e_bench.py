#!/usr/bin/env python
"""End-to-end (host buffers) step time of ctr_step_host for the chunk count in $CTR_HOST_CHUNKS."""
import os, sys, time
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'gym-ctr-reach_b200'))
import ctr_reach_envs
from ctr_reach_envs.envs import CtrReachVecEnv
N = int(os.environ.get('KB_N', 131072))
env = CtrReachVecEnv(N, device=0, **ctr_reach_envs.default_kwargs())
env.reset()
D = env.obs_dim
lim = env.action_space.high
rng = np.random.RandomState(0)
acts = []
for k in range(4):
    t = torch.empty((6, N), dtype=torch.float32).pin_memory()
    t.copy_(torch.tensor(rng.uniform(-lim, lim, size=(N, 6)).astype(np.float32).T.copy()))
    acts.append(t.numpy())
pin = lambda shape, dt: torch.empty(shape, dtype=dt).pin_memory().numpy()
out = env.alloc_host_outputs() if os.environ.get('E2E_PACKED', '1') == '1' else {'observation': pin((D, N), torch.float64), 'achieved_goal': pin((3, N), torch.float64), 'desired_goal': pin((3, N), torch.float64),
       'reward': pin(N, torch.float64), 'done': pin(N, torch.uint8), 'is_success': pin(N, torch.uint8), 'error': pin(N, torch.float64)}
for i in range(5):
    env.step_host(acts[i % 4], out)
ts = []
for i in range(40):
    t0 = time.perf_counter(); env.step_host(acts[i % 4], out); ts.append(time.perf_counter() - t0)
ts.sort()
print("chunks=%s  e2e median %.3f ms (min %.3f) = %.3e env-steps/s  checksum %.0f" % (
    os.environ.get('CTR_HOST_CHUNKS', 'auto'), ts[len(ts) // 2] * 1e3, ts[0] * 1e3, N / ts[len(ts) // 2], out['reward'].sum()))
