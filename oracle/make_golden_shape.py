"""Generate tests/golden/shape_reference.npz from the UNMODIFIED reference (build container only; TEST INFRASTRUCTURE).

Backbone shape of Model.forward_kinematics (ctr_reach_envs/envs/model.py:66-68,137-172): `model.r` = 30 samples per section
(solve_ivp t_eval = np.linspace(span[k], span[k+1] - 1e-6, 30), dense output of scipy RK45) and the split points of
r1 / r2 / r3 (`tip_pos`, model.py:166-172).  Joints: every 3rd configuration of tests/golden/fk_reference.npz (all 4 systems,
sampled + edge sets).  Run: python oracle/make_golden_shape.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_loader  # noqa: E402

OUT = os.path.join(HERE, '..', 'tests', 'golden')
MAXPTS = 270    # 9 sections x 30


def main():
    pkg, gym = ref_loader.load()
    env = gym.make('CTR-Reach-v0', select_systems=[0, 1, 2, 3])
    model = env.model
    cap = {}
    orig = model.ctr_model

    def spy(*a, **k):                      # tip_pos (model.py:166-172) is returned by ctr_model and dropped by the caller
        out = orig(*a, **k)
        cap['tip_pos'] = np.array(out[2], dtype=np.int32)
        return out

    model.ctr_model = spy
    g = np.load(os.path.join(OUT, 'fk_reference.npz'))
    Q, SYS = g['joints'][::3], g['system'][::3]
    R, NP, TIPIDX, TIP = [], [], [], []
    for q, s in zip(Q, SYS):
        tip = model.forward_kinematics(np.asarray(q, dtype=np.float32), int(s))
        r = np.asarray(model.r)
        n = r.shape[0]
        assert n % 30 == 0 and n <= MAXPTS and np.array_equal(tip, r[-1])
        t0, t1, t2 = (int(v) for v in cap['tip_pos'])
        assert np.array_equal(model.r1, r[t1:t0 + 1]) and np.array_equal(model.r2, r[t2:t1 + 1]) and np.array_equal(model.r3, r[:t2 + 1])
        pad = np.zeros((MAXPTS, 3))
        pad[:n] = r
        R.append(pad), NP.append(n), TIPIDX.append([t0, t1, t2]), TIP.append(tip)
    np.savez_compressed(os.path.join(OUT, 'shape_reference.npz'), joints=Q.astype(np.float32), system=SYS.astype(np.int32),
                        r=np.array(R), npts=np.array(NP, dtype=np.int32), tip_idx=np.array(TIPIDX, dtype=np.int32),
                        tip=np.array(TIP))
    print('shape_reference:', len(Q), 'configs; points per config', sorted(set(NP)))


if __name__ == '__main__':
    main()
