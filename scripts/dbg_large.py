import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ('gym-ctr-reach_b200', 'oracle'):
    sys.path.insert(0, os.path.join(ROOT, p))
import ctr_reach_envs, ctr_oracle as O, ctr_oracle_c as C
from ctr_reach_envs.envs import CtrReachVecEnv
rng = np.random.RandomState(11)
L = C.derive(O.DEFAULT_SYSTEMS[:1])[0][:3]
q = np.stack([O.sample_goal(rng, L) for _ in range(3000)]).astype(np.float64)
big = rng.rand(len(q)) < 0.3
q[big, 3:] += rng.choice([-1.0, 1.0], size=(big.sum(), 3)) * rng.uniform(8.0e4, 4.0e6, size=(big.sum(), 3))
q[::97, 3] = 9.0e4
q[::89, 4] = -89999.5
ref = C.fk_batch(C.derive(O.DEFAULT_SYSTEMS[:1]), q, None)
env = CtrReachVecEnv(8, device=0, **ctr_reach_envs.default_kwargs())
tip, R, info = env.forward_kinematics(q, return_info=True)
e = np.linalg.norm(tip.cpu().numpy() - ref['tip'], axis=1)
info = info.cpu().numpy()
idx = np.argsort(-np.where(big, 0, e))[:12]
for i in idx:
    print(i, big[i], i % 97 == 0, i % 89 == 0, e[i], q[i, 3:], info[i], ref['info'][i])
print("p99", np.percentile(e, 99), "max", e.max(), "max nonbig", e[~big].max())
special = (np.arange(len(q)) % 97 == 0) | (np.arange(len(q)) % 89 == 0)
print("max ordinary", e[~big & ~special].max())
