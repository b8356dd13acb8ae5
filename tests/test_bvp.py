"""BVP kinematics back-end (CTR_Python/CTR_Model.py) and the DLS controller on it (row 3).

Golden vectors: tests/golden/bvp_reference.npz, produced by the reference's own CTR_Model class (oracle/make_golden_bvp.py;
the one load-time substitution np.empty -> np.zeros that makes the class reproducible is described there).
CPU part: the C restatement (oracle/bvp_oracle.c) against those vectors.  GPU part: the CUDA kernels through the C-ABI against
the vectors and against the C restatement."""
import numpy as np
import pytest


@pytest.fixture(scope='module')
def bvp(golden):
    return golden('bvp_reference')


def _tcs(oracle_c, oracle_py):
    return oracle_c.derive(oracle_py.DEFAULT_SYSTEMS)


# ------------------------------------------------------------------------------------------------ CPU: oracle pinned
def test_bvp_oracle_pass_vs_reference(bvp, oracle_c, oracle_py):
    tcs = _tcs(oracle_c, oracle_py)
    n = bvp['pass_q'].shape[0]
    assert n >= 500
    worst_tip = worst_cost = worst_R = 0.0
    for i in range(n):
        o = oracle_c.bvp_pass(tcs[bvp['pass_system'][i]], bvp['pass_q'][i], bvp['pass_f'][i], bvp['pass_u'][i],
                              max_step=float(bvp['pass_tol'][i]))
        assert o['steps'] == int(bvp['pass_steps'][i]), i            # the same accept / reject decisions of RK23
        worst_tip = max(worst_tip, np.abs(o['tip'] - bvp['pass_tip'][i]).max())
        worst_cost = max(worst_cost, np.abs(o['cost'] - bvp['pass_cost'][i]).max())
        worst_R = max(worst_R, np.abs(o['R'] - bvp['pass_R'][i]).max())
    assert worst_tip < 1e-12 and worst_R < 1e-11 and worst_cost < 1e-9, (worst_tip, worst_R, worst_cost)


def test_bvp_oracle_jacobian_and_compliance_vs_reference(bvp, oracle_c, oracle_py):
    tcs = _tcs(oracle_c, oracle_py)
    for i in range(bvp['jac_q'].shape[0]):
        J, C, tip = oracle_c.bvp_jac_comp(tcs[bvp['jac_system'][i]], bvp['jac_q'][i], bvp['jac_f'][i], bvp['jac_u'][i])
        assert np.abs(tip - bvp['jac_tip'][i]).max() < 1e-12
        assert np.abs(J - bvp['jac_J'][i]).max() < 1e-7, i             # forward differences / 1e-4 of 1e-12-exact tips
        assert np.abs(C - bvp['jac_C'][i]).max() < 1e-9, i


def test_bvp_reference_minimize_rows_satisfy_their_own_tolerance(bvp, oracle_c, oracle_py):
    """The reference's shooting solutions (scipy.optimize.anderson, f_tol 1e-3), re-evaluated by the restated pass."""
    tcs = _tcs(oracle_c, oracle_py)
    ok = bvp['min_ok'].astype(bool)
    assert ok.sum() >= 12
    for i in np.nonzero(ok)[0]:
        o = oracle_c.bvp_pass(tcs[bvp['min_system'][i]], bvp['min_q'][i], bvp['min_f'][i], bvp['min_u'][i])
        assert np.abs(o['cost'] - bvp['min_cost'][i]).max() < 1e-8
        assert np.abs(o['cost']).max() < 1e-3 and np.abs(o['tip'] - bvp['min_tip'][i]).max() < 1e-11


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
def test_bvp_pass_gpu_vs_reference_and_oracle(bvp, oracle_c, oracle_py, vec_factory):
    env = vec_factory(16, select_systems=[0, 1, 2, 3])
    for tol in (0.01, 0.002):
        sel = np.nonzero(bvp['pass_tol'] == tol)[0]
        out = env.bvp_pass(bvp['pass_q'][sel], force=bvp['pass_f'][sel], u_init=bvp['pass_u'][sel],
                           system=bvp['pass_system'][sel].astype(np.int32), max_step=tol)
        tip, R, cost = out['tip'].cpu().numpy(), out['R'].cpu().numpy(), out['cost'].cpu().numpy()
        assert np.array_equal(out['steps'].cpu().numpy(), bvp['pass_steps'][sel])          # identical RK23 decisions
        assert np.abs(tip - bvp['pass_tip'][sel]).max() < 1e-10                            # bar: 1e-5 m
        assert np.abs(R - bvp['pass_R'][sel]).max() < 1e-9                                 # bar: 1e-5 rad
        assert np.abs(cost - bvp['pass_cost'][sel]).max() < 1e-8
    # random configurations with per-configuration tube constants against the C restatement
    rng = np.random.RandomState(3)
    n = 3000
    tcs = _tcs(oracle_c, oracle_py)
    systems = np.stack([oracle_py.randomize_system(rng, oracle_py.DEFAULT_SYSTEMS[rng.randint(4)], 0.05) for _ in range(n)])
    P = oracle_c.derive(systems)
    q = np.stack([oracle_py.sample_goal(rng, P[i, :3]) for i in range(n)]).astype(np.float64)
    f = rng.uniform(-0.3, 0.3, size=(n, 3))
    u = rng.uniform(-2, 2, size=(n, 5))
    out = env.bvp_pass(q, force=f, u_init=u, params=P)
    tip, cost, steps = out['tip'].cpu().numpy(), out['cost'].cpu().numpy(), out['steps'].cpu().numpy()
    bad = 0
    for i in range(n):
        o = oracle_c.bvp_pass(P[i], q[i], f[i], u[i])
        if o['steps'] != steps[i]:
            bad += 1                                                                     # a borderline decision may flip (1e-16 arithmetic)
            continue
        assert np.abs(o['tip'] - tip[i]).max() < 1e-10 and np.abs(o['cost'] - cost[i]).max() < 1e-7, i
    assert bad <= 3
    assert tcs.shape == (4, 18)


@pytest.mark.gpu
def test_bvp_jacobian_and_compliance_gpu_vs_reference(bvp, vec_factory):
    env = vec_factory(16, select_systems=[0, 1, 2, 3])
    J, C, tip = env.bvp_jacobian(bvp['jac_q'], force=bvp['jac_f'], u_init=bvp['jac_u'], system=bvp['jac_system'].astype(np.int32))
    assert np.abs(tip.cpu().numpy() - bvp['jac_tip']).max() < 1e-10
    assert np.abs(J.cpu().numpy() - bvp['jac_J']).max() < 1e-5       # differences of 1e-10-exact tips over eps = 1e-4
    assert np.abs(C.cpu().numpy() - bvp['jac_C']).max() < 1e-7


@pytest.mark.gpu
def test_bvp_shooting_reaches_the_references_tolerance(bvp, oracle_c, oracle_py, vec_factory):
    """ctr_bvp_solve drives the same residuals below the same f_tol as CTR_Model.minimize; the two root finders stop at
    different points of that tolerance ball (and, for high-twist configurations, possibly at different equilibria)."""
    tcs = _tcs(oracle_c, oracle_py)
    env = vec_factory(16, select_systems=[0, 1, 2, 3])
    ok = bvp['min_ok'].astype(bool)
    q, f, sysid = bvp['min_q'][ok], bvp['min_f'][ok], bvp['min_system'][ok].astype(np.int32)
    out = env.bvp_solve(q, force=f, system=sysid, f_tol=1e-3, max_iters=80)
    res, u, tip = out['residual'].cpu().numpy(), out['u'].cpu().numpy(), out['tip'].cpu().numpy()
    conv = res < 1e-3
    assert conv.sum() >= q.shape[0] - 1, res                       # the reference's own Anderson iteration gives up on 2 of 20 rows
    for i in range(q.shape[0]):                                     # the residual claim, checked by the reference-pinned oracle
        o = oracle_c.bvp_pass(tcs[sysid[i]], q[i], f[i], u[i])
        assert np.abs(np.abs(o['cost']).max() - res[i]) < 1e-6 and np.abs(o['tip'] - tip[i]).max() < 1e-9
    d = np.linalg.norm(tip - bvp['min_tip'][ok], axis=1)[conv]
    # same root as the reference's Anderson iteration wherever the problem has one nearby: median 5e-7 m; the high-twist rows
    # have several equilibria (both solvers end with residual < 1e-3, centimetres apart) -- at most 2 such rows here
    assert np.median(d) < 1e-5 and (d < 1e-4).sum() >= conv.sum() - 2, (np.median(d), np.sort(d)[-4:])
    tight = env.bvp_solve(q, force=f, system=sysid, f_tol=1e-8, max_iters=80)
    rt = tight['residual'].cpu().numpy()
    assert (rt < 1e-8).sum() >= conv.sum() - 1, rt                 # the same rows converge further where anderson stops at 1e-3


@pytest.mark.gpu
def test_dls_controller_on_the_bvp_model_vs_oracle(oracle_c, oracle_py, vec_factory):
    """src/jacobian_controller.py:19-53 as written: CTR_Model(.., f = 0, q, q_0 = 0, 0.01, 1).jac(0) every iteration,
    dq = J^T (J J^T + 0.25 I)^-1 e, stop once the error before the update was < 1e-3.  Oracle: the same loop over the
    reference-pinned C pass (oracle/bvp_oracle.c)."""
    tcs = _tcs(oracle_c, oracle_py)
    rng = np.random.RandomState(12)
    n = 200
    q0 = np.stack([oracle_py.sample_goal(rng, tcs[0, :3]) for _ in range(n)]).astype(np.float64)
    start = np.stack([oracle_c.bvp_pass(tcs[0], q0[i])['tip'] for i in range(n)])
    target = start + rng.uniform(-1, 1, size=(n, 3)) / 200 * 0.6          # jacobian_controller.py:103: ag + rand(3) / 200
    env = vec_factory(16)
    q, err, iters = env.bvp_ik_dls(target, q0, max_iters=200)
    q, err, iters = q.cpu().numpy(), err.cpu().numpy(), iters.cpu().numpy()
    assert (err < 1e-3).mean() > 0.7          # lam = 0.25 against |J J^T| ~ 0.3: the controller as written converges slowly (83 % in 200 iterations)
    agree = 0
    for i in range(0, n, 8):
        qq = q0[i].copy()
        for it in range(200):
            J, _, P = oracle_c.bvp_jac_comp(tcs[0], qq)
            e = target[i] - P
            qq = qq + J.T @ np.linalg.inv(J @ J.T + 0.25 * np.eye(3)) @ e
            if np.linalg.norm(e) < 1e-3:
                break
        if it + 1 == iters[i] and np.abs(qq - q[i]).max() < 1e-6:
            agree += 1
    assert agree >= 0.9 * len(range(0, n, 8)), agree


@pytest.mark.gpu
def test_bvp_pass_is_independent_of_batch_shape_and_isolates_failed_rows(bvp, vec_factory):
    """The lane-group kernel hands passes to whichever group is free: a row's result must not depend on the batch it is in
    (1 row, a partial warp, partial blocks, more rows than resident groups), and a row that cannot be integrated (NaN joint)
    comes back as NaN / steps -1 and is counted, without touching its neighbours."""
    import torch
    env = vec_factory(16, select_systems=[0, 1, 2, 3])
    sel = np.nonzero(bvp['pass_tol'] == 0.01)[0]
    q, f, u, sysid = bvp['pass_q'][sel], bvp['pass_f'][sel], bvp['pass_u'][sel], bvp['pass_system'][sel].astype(np.int32)
    full = env.bvp_pass(q, force=f, u_init=u, system=sysid)
    full = {k: v.cpu().numpy() for k, v in full.items()}
    for n in (1, 7, 11, 41, 333):
        part = env.bvp_pass(q[:n], force=f[:n], u_init=u[:n], system=sysid[:n])
        for k in ('tip', 'R', 'cost', 'steps'):
            assert np.array_equal(part[k].cpu().numpy(), full[k][:n]), (n, k)
    reps = 40                                                       # 40 x 512 rows: more passes than resident lane groups (148 x 3 x 40)
    big = env.bvp_pass(np.tile(q, (reps, 1)), force=np.tile(f, (reps, 1)), u_init=np.tile(u, (reps, 1)), system=np.tile(sysid, reps))
    for k in ('tip', 'cost', 'steps'):
        got = big[k].cpu().numpy().reshape((reps, len(sel)) + full[k].shape[1:])
        assert all(np.array_equal(got[r], full[k]) for r in range(reps)), k
    before = env.fault_counters()['fk_failed']
    qb = q[:64].copy()
    qb[5, 4] = np.nan
    qb[33, 3] = np.nan
    out = env.bvp_pass(qb, force=f[:64], u_init=u[:64], system=sysid[:64])
    steps, tip = out['steps'].cpu().numpy(), out['tip'].cpu().numpy()
    torch.cuda.synchronize()
    bad = np.zeros(64, dtype=bool)
    bad[[5, 33]] = True
    assert np.all(steps[bad] == -1) and np.all(np.isnan(tip[bad]))
    assert np.array_equal(steps[~bad], full['steps'][:64][~bad]) and np.array_equal(tip[~bad], full['tip'][:64][~bad])
    assert env.fault_counters()['fk_failed'] - before == 2


@pytest.mark.gpu
def test_bvp_argument_errors_and_joint_box_ends(oracle_c, oracle_py, vec_factory):
    from ctr_reach_envs import _lib
    env = vec_factory(8)
    q = np.zeros((4, 6))
    for bad in (0.0, -0.01, float('nan')):
        with pytest.raises(_lib.CtrError):
            env.bvp_pass(q, max_step=bad)
        with pytest.raises(_lib.CtrError):
            env.bvp_jacobian(q, max_step=bad)
    rc = env.lib.ctr_bvp_pass(env.handle, None, None, None, None, None, 4, 0.01, None, None, None, None, env._stream())
    assert rc != 0 and b'null' in env.lib.ctr_last_error()
    # both ends of the joint box (fully extended, fully retracted) integrate without faults
    import ctr_oracle as O
    L = O.DEFAULT_SYSTEMS[0][:, 0]
    ends = np.zeros((2, 6))
    ends[1, :3] = -(L - 1e-3)
    res = env.bvp_pass(ends)
    assert (res['steps'].cpu().numpy() > 0).all() and np.isfinite(res['tip'].cpu().numpy()).all()
    tcs = _tcs(oracle_c, oracle_py)
    for i in range(2):                      # retracted: ~1 mm of straight tube along z (0.98 mm after Segment.py:16's 1e-5 quantisation)
        o = oracle_c.bvp_pass(tcs[0], ends[i])
        assert o['steps'] == int(res['steps'][i]) and np.abs(o['tip'] - res['tip'][i].cpu().numpy()).max() < 1e-12
    assert abs(res['tip'].cpu().numpy()[1, 2] - 1e-3) < 5e-5


@pytest.mark.gpu
def test_bvp_scheduling_order_does_not_change_results(oracle_c, oracle_py, vec_factory, monkeypatch):
    """From 16384 passes on, the passes are handed out in a sorted order (few sections first, long robots first;
    k_bvp_order_*).  Same inputs with the ordering switched off ($CTR_BVP_SORT=0, read at ctr_create): bit-identical outputs,
    failed rows (NaN joints) included; Jacobian variants too."""
    rng = np.random.RandomState(21)
    n = 18000
    systems = np.stack([oracle_py.randomize_system(rng, oracle_py.DEFAULT_SYSTEMS[rng.randint(4)], 0.05) for _ in range(n)])
    P = oracle_c.derive(systems)
    q = np.stack([oracle_py.sample_goal(rng, P[i, :3]) for i in range(n)]).astype(np.float64)
    f = rng.uniform(-0.3, 0.3, size=(n, 3))
    q[[17, 4100, 17999], 1] = np.nan
    monkeypatch.setenv('CTR_BVP_SORT', '1')
    env_sorted = vec_factory(16)
    monkeypatch.setenv('CTR_BVP_SORT', '0')
    env_plain = vec_factory(16)
    a = env_sorted.bvp_pass(q, force=f, params=P)
    b = env_plain.bvp_pass(q, force=f, params=P)
    for k in ('tip', 'R', 'cost', 'steps'):
        assert np.array_equal(a[k].cpu().numpy(), b[k].cpu().numpy(), equal_nan=True), k
    assert (a['steps'].cpu().numpy() == -1).sum() == 3
    ok = np.ones(n, dtype=bool)
    ok[[17, 4100, 17999]] = False
    Ja, Ca, ta = env_sorted.bvp_jacobian(q[ok][:4500], force=f[ok][:4500], params=P[ok][:4500])
    Jb, Cb, tb = env_plain.bvp_jacobian(q[ok][:4500], force=f[ok][:4500], params=P[ok][:4500])
    assert np.array_equal(Ja.cpu().numpy(), Jb.cpu().numpy()) and np.array_equal(Ca.cpu().numpy(), Cb.cpu().numpy())
    assert np.array_equal(ta.cpu().numpy(), tb.cpu().numpy())
