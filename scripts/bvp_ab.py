#!/usr/bin/env python
"""Lane-group BVP kernel against the thread-per-pass kernel ($CTR_BVP_THREAD_KERNEL, read once per process) on the same random
configurations (4 systems, tip loads, random u_init).  usage: bvp_ab.py dump <file>   (run once per mode)  |  bvp_ab.py cmp <a> <b>"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if sys.argv[1] == 'cmp':
    a, b = np.load(sys.argv[2]), np.load(sys.argv[3])
    same = a['steps'] == b['steps']
    ok = same & (a['steps'] >= 0)
    print("rows %d, identical step counts %d (%.5f), failed a/b %d/%d" % (len(same), same.sum(), same.mean(), (a['steps'] < 0).sum(), (b['steps'] < 0).sum()))
    for k in ('tip', 'R', 'cost'):
        print("  %-4s max |a - b| over rows with identical steps: %.3e   over all finite rows: %.3e" % (
            k, np.abs(a[k][ok] - b[k][ok]).max(), np.nanmax(np.abs(a[k] - b[k]))))
    sys.exit(0)
import torch
sys.path.insert(0, os.path.join(ROOT, 'gym-ctr-reach_b200'))
sys.path.insert(0, os.path.join(ROOT, 'oracle'))
import ctr_reach_envs
from ctr_reach_envs.envs import CtrReachVecEnv
import ctr_oracle as O
n = 200000
rng = np.random.RandomState(11)
kw = ctr_reach_envs.default_kwargs()
kw['select_systems'] = [0, 1, 2, 3]
env = CtrReachVecEnv(16, device=0, **kw)
sysid = rng.randint(4, size=n).astype(np.int32)
L = np.stack([O.DEFAULT_SYSTEMS[s][:, 0] for s in range(4)])[sysid]
ext = -rng.uniform(0, 1, size=(n, 3)) * (L - 1e-3)
ext = np.sort(ext, axis=1)                                   # beta_1 <= beta_2 <= beta_3 <= 0 (the deployed lengths stay ordered for these systems)
q = np.concatenate([ext, rng.uniform(-np.pi, np.pi, size=(n, 3))], axis=1)
f = rng.uniform(-0.3, 0.3, size=(n, 3)) * (rng.rand(n, 1) < 0.5)
u = rng.uniform(-2, 2, size=(n, 5)) * (rng.rand(n, 1) < 0.5)
out = env.bvp_pass(q, force=f, u_init=u, system=sysid)
np.savez(sys.argv[2], **{k: v.cpu().numpy() for k, v in out.items()})
print("wrote", sys.argv[2], "thread kernel" if os.environ.get('CTR_BVP_THREAD_KERNEL') else "lane-group kernel",
      "mean steps %.1f" % out['steps'].double().mean().item())
