"""Finite-difference Jacobian / damped-least-squares IK: ctr_jacobian / ctr_ik_dls against
oracle/ik_oracle.py.  The reference's controller script (src/jacobian_controller.py) cannot be run (missing modules, out-of-scope
BVP model): parity here is against the RESTATEMENT of its algorithm over Model.forward_kinematics, not against reference output.
Tolerances: a Jacobian entry is a tip difference over eps = 1e-4, so the 1e-10 m FK bar becomes 2e-6 per entry (measured ~1e-8);
joints after the whole iteration 1e-6 (rad / m)."""
import numpy as np
import pytest

import ctr_oracle as O
import ik_oracle as K


def _configs(tcs, system, n, seed):
    rng = np.random.RandomState(seed)
    return np.stack([O.sample_goal(rng, tcs[system][:3]) for _ in range(n)]).astype(np.float64), rng


def test_oracle_jacobian_predicts_displacements(oracle_c, tube_constants):
    q, rng = _configs(tube_constants, 0, 200, 1)
    J, P = K.jacobian(tube_constants, q)
    dq = rng.uniform(-1, 1, size=(200, 6)) * np.array([2e-4] * 3 + [2e-3] * 3)
    moved = oracle_c.fk_batch(tube_constants, q + dq, None)['tip']
    lin = P + (J @ dq[:, :, None])[:, :, 0]
    assert np.abs(P - oracle_c.fk_batch(tube_constants, q, None)['tip']).max() == 0.0
    # first-order model.  The algorithm differentiates THROUGH an adaptive solver: the tip carries ~1e-6 m of step-decision
    # noise, which eps = 1e-4 turns into ~1e-2 per Jacobian entry, i.e. ~2e-5 m over |dq| ~ 2e-3 (measured max 2.3e-5, median
    # 5e-6 against a median displacement of 1.6e-4): a property of the reference's scheme, bounded here, not tuned away
    res, disp = np.abs(moved - lin).max(axis=1), np.abs(moved - P).max(axis=1)
    assert res.max() < 5e-5 and np.median(res) < 0.1 * np.median(disp) and disp.max() > 1e-4


def test_oracle_ik_converges_and_stops_after_the_update(tube_constants):
    q, rng = _configs(tube_constants, 0, 48, 2)
    _, P = K.jacobian(tube_constants, q)
    target = P + rng.uniform(-1, 1, size=(48, 3)) * 0.004
    target[:4] = P[:4]                                       # already there: one iteration, then stop (update applied first)
    qq, err, it = K.dls_ik(tube_constants, target, q, max_iters=80)
    assert (it[:4] == 1).all() and (err[:4] < 1e-12).all()
    assert (err < 1e-3).mean() > 0.9 and it.max() <= 80
    # the returned error is the one BEFORE the last update (jacobian_controller.py:44-50); that update uses a noisy
    # finite-difference Jacobian and may not improve it (measured: median ratio after / before 0.88, worst 1.04), so the
    # reached tip is only guaranteed close to the tolerance, not under it
    final = np.linalg.norm(K.jacobian(tube_constants, qq)[1] - target, axis=1)
    assert (final[err < 1e-3] < 1.1e-3).all() and np.median(final[(err < 1e-3) & (err > 0)] / err[(err < 1e-3) & (err > 0)]) < 1.0


@pytest.mark.gpu
@pytest.mark.parametrize('system', [0, 2])
def test_jacobian_matches_oracle(tube_constants, vec_factory, system):
    env = vec_factory(8, select_systems=[0, 1, 2, 3])
    q, _ = _configs(tube_constants, system, 1500, 10 + system)
    sysid = np.full(len(q), system, dtype=np.int32)
    J, tip = env.jacobian(q, system=sysid)
    Jo, Po = K.jacobian(tube_constants, q, sysid)
    assert np.abs(tip.cpu().numpy() - Po).max() < 1e-10
    assert np.abs(J.cpu().numpy() - Jo).max() < 2e-6
    assert np.array_equal(tip.cpu().numpy(), env.forward_kinematics(q, system=sysid).cpu().numpy())   # column block 0 is ctr_fk


@pytest.mark.gpu
def test_ik_dls_matches_oracle(tube_constants, vec_factory):
    env = vec_factory(8)
    q, rng = _configs(tube_constants, 0, 600, 5)
    _, P = K.jacobian(tube_constants, q)
    target = P + rng.uniform(-1, 1, size=(600, 3)) * 0.004
    target[:5] = P[:5]
    for max_iters in (3, 40):                                # a capped run (nobody converged) and a run to convergence
        qg, eg, ig = env.ik_dls(target, q, max_iters=max_iters)
        qo, eo, io = K.dls_ik(tube_constants, target, q, max_iters=max_iters)
        qg, eg, ig = qg.cpu().numpy(), eg.cpu().numpy(), ig.cpu().numpy()
        same = ig == io
        assert same.mean() > 0.99                            # an error within round-off of 1e-3 may stop one iteration apart
        # Measured (profiles/r1m_ik_deviation_gpu_vs_oracle.txt): 99.8 % of the configurations agree to < 1e-9 (median 1e-14), and at
        # IDENTICAL inputs the two FK implementations agree to 4e-16 m with the same step decisions.  The rest is the reference's
        # own kinematics: Segment.py:16 quantises section lengths with 1e-5 * floor(1e5 * diff), every configuration has a breakpoint
        # pair exactly Lc apart, so for joints that are arbitrary doubles the floor can hinge on the LAST BITS of beta -- the two
        # pipelines' joints differ by ~1e-14 after the first 3x3 solve (LU in numpy, adjugate here), one solve lands on the other side,
        # its tip jumps by a quantum (8.4e-6 m) and eps = 1e-4 carries that into the update (7e-6 in the joints).
        dev = np.abs(qg - qo).max(axis=1)[same]
        assert (dev < 1e-9).mean() > 0.99 and dev.max() < 1e-4
        assert (np.abs(eg - eo)[same] < 1e-9).mean() > 0.99 and np.abs(eg - eo)[same].max() < 1e-4
    assert (ig[:5] == 1).all()
    assert ((eg < 1e-3) == (eo < 1e-3)).mean() > 0.99 and (eg < 1e-3).mean() > 0.5     # the same configurations converge in 40 iterations
    reached = env.forward_kinematics(qg).cpu().numpy()
    assert (np.linalg.norm(reached - target, axis=1)[eg < 1e-3] < 1.1e-3).all()      # see the oracle test: error BEFORE the last update


@pytest.mark.gpu
def test_ik_argument_errors(vec_factory):
    from ctr_reach_envs import _lib
    env = vec_factory(8)
    with pytest.raises(_lib.CtrError):
        env.jacobian(np.zeros((4, 6)), eps=0.0)
    with pytest.raises(_lib.CtrError):
        env.ik_dls(np.zeros((4, 3)), np.zeros((4, 6)), max_iters=-1)
