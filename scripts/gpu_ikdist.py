import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ('gym-ctr-reach_b200', 'oracle', 'tests'):
    sys.path.insert(0, os.path.join(ROOT, p))
import ctr_oracle as O, ctr_oracle_c as C, ik_oracle as K
from conftest import make_vec
tcs = C.derive(O.DEFAULT_SYSTEMS)
rng = np.random.RandomState(5)
q = np.stack([O.sample_goal(rng, tcs[0][:3]) for _ in range(600)]).astype(np.float64)
_, P = K.jacobian(tcs, q)
target = P + rng.uniform(-1, 1, size=(600, 3)) * 0.004
env = make_vec(8)
for it in (1, 3, 10, 40):
    qg, eg, ig = env.ik_dls(target, q, max_iters=it)
    qo, eo, io = K.dls_ik(tcs, target, q, max_iters=it)
    d = np.abs(qg.cpu().numpy() - qo).max(axis=1)
    de = np.abs(eg.cpu().numpy() - eo)
    print("max_iters %2d: iters equal %.4f | per-config max|dq|: median %.1e p90 %.1e p99 %.1e max %.1e ; frac > 1e-9: %.3f | |derr| max %.1e"
          % (it, (ig.cpu().numpy() == io).mean(), np.median(d), np.percentile(d, 90), np.percentile(d, 99), d.max(), (d > 1e-9).mean(), de.max()))
# how often do GPU and oracle FK take different step decisions on the perturbed joints of ONE Jacobian?
blk = np.repeat(q[None], 7, axis=0)
for c in range(6):
    blk[c + 1, :, c] += 1e-4
blk = blk.reshape(-1, 6)
tip, R, info = env.forward_kinematics(blk, return_info=True)
ref = C.fk_batch(tcs, blk, None)
diff = info.cpu().numpy()[:, 0] != ref['info'][:, 0]
print("solves with a different nfev: %d of %d; max tip diff %.1e (those: %.1e)" % (diff.sum(), len(diff), np.abs(tip.cpu().numpy() - ref['tip']).max(),
      np.abs(tip.cpu().numpy() - ref['tip'])[diff].max() if diff.any() else 0.0))

# ---- locate the deviating configuration and compare the two FK implementations at IDENTICAL inputs
print("---- locating the deviation")
qg_prev = q.copy()
for it in (1, 2, 3):
    qg, _, _ = env.ik_dls(target, q, max_iters=it)
    qo, _, _ = K.dls_ik(tcs, target, q, max_iters=it)
    qg = qg.cpu().numpy()
    d = np.abs(qg - qo).max(axis=1)
    i = int(d.argmax())
    print("after %d iterations: worst config %d, |dq| %.2e" % (it, i, d[i]))
    if d[i] > 1e-9:
        # the Jacobian solves of iteration `it` were made at the joints after it-1 iterations: same input to both
        qin = qg_prev[i:i + 1]
        blk = np.repeat(qin[None], 7, axis=0)
        for c in range(6):
            blk[c + 1, :, c] += 1e-4
        blk = blk.reshape(-1, 6)
        tip, R, info = env.forward_kinematics(blk, return_info=True)
        ref = C.fk_batch(tcs, blk, None)
        print("  joints", np.array2string(qin[0], precision=17))
        print("  gpu    nfev/acc/rej/sec", info.cpu().numpy().tolist())
        print("  oracle nfev/acc/rej/sec", ref['info'].tolist())
        print("  |tip diff| per solve", np.abs(tip.cpu().numpy() - ref['tip']).max(axis=1))
        np.save(os.path.join(ROOT, 'gpurun_out', 'ik_flip_joints.npy'), blk)
        break
    qg_prev = qg
