import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'gym-ctr-reach_b200'))
import ctr_reach_envs
from ctr_reach_envs.envs import CtrReachVecEnv
kw = ctr_reach_envs.default_kwargs(); kw['max_steps_per_episode'] = 7
n = 2500
auto = CtrReachVecEnv(n, device=0, auto_reset=True, seed=3, **kw)
man = CtrReachVecEnv(n, device=0, auto_reset=False, seed=3, **kw)
for e in (auto, man):
    e.goal_tolerance.current_tol = 0.05; e.reset()
rng = np.random.RandomState(1); lim = auto.action_space.high
act = torch.tensor(rng.uniform(-lim, lim, size=(n, 6)).astype(np.float32), device='cuda')
_, _, d_a, _ = auto.step(act); _, _, d_m, _ = man.step(act)
man.reset(mask=d_m.to(torch.uint8)); torch.cuda.synchronize()
for k in ('joints', 'desired_joints', 'desired_goal', 'starting_position', 'orientation', 'achieved_goal'):
    a, b = auto.buffer(k).double(), man.buffer(k).double()
    diff = (a - b).abs()
    print(k, "ndiff", int((diff > 0).sum()), "max", float(diff.max()), "n done", int(d_m.sum()))
idx = ((auto.buffer('desired_goal') - man.buffer('desired_goal')).abs().max(0).values > 0).nonzero().flatten()[:5]
print(idx, auto.buffer('desired_joints')[:, idx].t(), auto.buffer('nfev')[idx], man.buffer('nfev')[idx])
