"""CPU restatement of the finite-difference Jacobian / damped-least-squares IK -- TEST INFRASTRUCTURE ONLY.

Follows ctr_reach_envs/src/jacobian_controller.py:19-73 (dls_ik_position_only: e = Pd - P, dq = J^T inv(J J^T + lam I),
q += dq e, break once ||e|| < 1e-3, at most `num` iterations, returns q and ||e||) and envs/CTR_Python/CTR_Model.py:251-262
(jac: forward differences, eps = 1e-4, one solve per joint).  The reference script imports modules that are not in its tree
(ctr_generic_envs, trajectory_plotter) and differentiates the BVP model CTR_Model, which SURVEY.md 2 puts out of scope, so it
cannot be run here or anywhere: PARITY IS PINNED TO THIS RESTATEMENT ONLY.  The kinematics are the C oracle's
Model.forward_kinematics (pinned to the reference, tests/test_oracle.py).
"""
import numpy as np

import ctr_oracle_c as C


def jacobian(tcs, q, system=None, eps=1e-4, nthreads=8):
    """q [n][6] float64 -> J [n][3][6], tip [n][3]."""
    q = np.asarray(q, dtype=np.float64).reshape(-1, 6)
    n = q.shape[0]
    block = np.repeat(q[None], 7, axis=0)                     # [7][n][6]
    for c in range(6):
        block[c + 1, :, c] = q[:, c] + eps                    # CTR_Model.py:257
    sysb = None if system is None else np.tile(np.asarray(system, dtype=np.int32), 7)
    tips = C.fk_batch(tcs, block.reshape(-1, 6), sysb, nthreads=nthreads)['tip'].reshape(7, n, 3)
    J = np.stack([(tips[c + 1] - tips[0]) / eps for c in range(6)], axis=2)     # CTR_Model.py:259-260
    return J, tips[0]


def dls_ik(tcs, target, q0, system=None, lam=0.25, eps=1e-4, tol=1e-3, max_iters=500, nthreads=8):
    """-> q [n][6], err [n] (last error norm evaluated), iters [n]."""
    q = np.array(q0, dtype=np.float64).reshape(-1, 6)
    target = np.asarray(target, dtype=np.float64).reshape(-1, 3)
    n = q.shape[0]
    err, iters, live = np.zeros(n), np.zeros(n, dtype=np.int32), np.ones(n, dtype=bool)
    for _ in range(max_iters):
        if not live.any():
            break
        idx = np.nonzero(live)[0]
        J, P = jacobian(tcs, q[idx], None if system is None else np.asarray(system)[idx], eps, nthreads)
        e = target[idx] - P                                                          # jacobian_controller.py:44
        A = J @ J.transpose(0, 2, 1) + lam * np.eye(3)
        dq = J.transpose(0, 2, 1) @ np.linalg.inv(A)                                 # :45
        q[idx] += (dq @ e[:, :, None])[:, :, 0]                                      # :47
        en = np.linalg.norm(e, axis=1)
        err[idx], iters[idx] = en, iters[idx] + 1
        live[idx[en < tol]] = False                                                  # :49-50 (after the update)
    return q, err, iters
