#!/usr/bin/env python
"""Small workload over every lock-free part of the step path, meant to run under compute-sanitizer (scripts/gpu_sanitize.sh):
prefetch slots with episode tags, the three rotating refill queues, per-window arrival counters and the window sort,
the per-warp shared-memory mailbox (racecheck: STS / __syncwarp / LDS ordering), the chunked host step, the replay ring,
the backbone-shape kernel.  Sizes are small (sanitizers slow kernels by 10-100x); tolerance 0.05 so that episodes end and
reset all the time."""
import os
import sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'gym-ctr-reach_b200'))
import ctr_reach_envs
from ctr_reach_envs.envs import CtrReachVecEnv
from ctr_reach_envs.her import HerReplay

N = int(os.environ.get('SAN_N', 3000))
STEPS = int(os.environ.get('SAN_STEPS', 6))
kw = ctr_reach_envs.default_kwargs()
rng = np.random.RandomState(0)
for extra in (dict(), dict(select_systems=[0, 1, 2, 3], domain_rand=0.05, domain_rand_length=0.03)):
    k = dict(kw)
    k.update(extra)
    env = CtrReachVecEnv(N, device=0, max_steps_per_episode=4, **{a: b for a, b in k.items() if a != 'max_steps_per_episode'})
    env.goal_tolerance.current_tol = 0.05
    env.reset()
    lim = torch.tensor(env.action_limits, device='cuda')
    rep = HerReplay(env, 8, n_sampled_goal=4, seed=1)
    rep.begin()
    for i in range(STEPS):
        act = ((torch.rand((N, 6), device='cuda') * 2 - 1) * lim).contiguous()
        env.step(act)
        rep.store(act)
    batch = rep.alloc_batch(4096)
    rep.sample(4096, out=batch, call=0)
    out = env.alloc_host_outputs()
    h_act = np.ascontiguousarray(rng.uniform(-1, 1, size=(6, N)).astype(np.float32) * env.action_limits[:, None])
    for i in range(3):
        env.step_host(h_act, out)
    mask = torch.zeros(N, dtype=torch.uint8, device='cuda')
    mask[::7] = 1
    env.reset(mask=mask)
    q = env.buffer('joints').t().to(torch.float64)[:512].contiguous()
    env.backbone_shape(q, system=env.buffer('system')[:512])
    env.forward_kinematics(q, system=env.buffer('system')[:512].tolist())
    st = env.stats(reduce=False)
    torch.cuda.synchronize()
    print("ok: %d envs, systems %s, episodes %d, resets %d, faults %s" % (N, k['select_systems'], st['episodes'], st['resets'], env.fault_counters()))
    env.close()
