#!/usr/bin/env python
"""Statistical pins for the two random parts of the reset path, drawn from the REFERENCE's own code
(oracle/ref_loader.py): Obs.sample_goal (obs.py:185-207) and model_utils.sample_parameters (model_utils.py:5-35).
There is no RNG-stream parity to have, so the pin is distributional: the samples below are kept as
sorted arrays, and the GPU tests compare the device's draws against them with a two-sample Kolmogorov-Smirnov bound.

TEST INFRASTRUCTURE; runs only in the build container.  Writes tests/golden/stat_reference.npz:
  goal_joints  f32 [4][4000][6]   accepted joints of sample_goal per tube system
  param_ratio  f64 [10000][3][5]  sampled / nominal for d_i, d_o, E, G, U_x of system 0's tubes at randomization 0.05
  param_fixed  f64 [3][3]         L, L_c, U_y of the sampled tubes (must equal nominal)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_loader  # noqa: E402

pkg, gym = ref_loader.load()
from ctr_reach_envs.envs.model_utils import sample_parameters  # noqa: E402
from ctr_reach_envs.envs.CTR_Python.Tube import Tube  # noqa: E402
import ctr_oracle as O  # noqa: E402


def main():
    np.random.seed(7)
    gym.spaces.seed(7)
    env = gym.make('CTR-Reach-v0', select_systems=[0, 1, 2, 3])
    goals = np.zeros((4, 4000, 6), dtype=np.float32)
    for s in range(4):
        for k in range(4000):
            goals[s, k] = env.trig_obj.sample_goal(s)
    nominal = [Tube(*O.DEFAULT_SYSTEMS[0][i]) for i in range(3)]
    ratio = np.zeros((10000, 3, 5))
    fixed = np.zeros((3, 3))
    for k in range(10000):
        for i, t in enumerate(nominal):
            p = sample_parameters(t, 0.05)
            ratio[k, i] = [p.diameter_inner / t.diameter_inner, p.diameter_outer / t.diameter_outer, p.E / t.E, p.G / t.G,
                           p.U_x / t.U_x]
            if k == 0:
                fixed[i] = [p.L, p.L_c, p.U_y]
                assert p.L == t.L and p.L_c == t.L_c and p.U_y == t.U_y
    path = os.path.join(HERE, '..', 'tests', 'golden', 'stat_reference.npz')
    np.savez_compressed(path, goal_joints=goals, param_ratio=ratio.astype(np.float32), param_fixed=fixed)
    print("wrote", os.path.abspath(path), os.path.getsize(path))


if __name__ == '__main__':
    main()
