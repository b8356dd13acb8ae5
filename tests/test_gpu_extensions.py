"""GPU tests of the extensions `north_star` names beyond the reference: per-env tube LENGTHS with per-env joint limits
(BASELINE.json config 5), normalised actions, and the batch form of the reference's action asserts
(ctr_reach_env.py:131-132).  All through the C-ABI; the checker is the C / numpy oracle."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _per_env_systems(oracle_py, rng, n, r_len):
    """n copies of system 0 whose tube lengths / curved lengths are drawn +-r_len around nominal."""
    systems = np.repeat(oracle_py.DEFAULT_SYSTEMS[:1], n, axis=0).copy()
    systems[:, :, 0] *= rng.uniform(1 - r_len, 1 + r_len, size=(n, 3))
    systems[:, :, 1] *= rng.uniform(1 - r_len, 1 + r_len, size=(n, 3))
    return systems


def test_per_env_lengths_give_per_env_joint_limits(oracle_c, oracle_py, vec_factory):
    """ctr_set_tube_params with per-env L: Obs.set_action's clip box and ordering constants follow each env's own lengths
    (obs.py:59-75,181 evaluated per env), bit-exact against the oracle run with one 'system' per env."""
    n = 6000
    rng = np.random.RandomState(21)
    systems = _per_env_systems(oracle_py, rng, n, 0.08)
    tcs = oracle_c.derive(systems)
    env = vec_factory(n, auto_reset=False, seed=5)
    env.reset()
    env.set_tube_params(tcs)
    # start from joints at / beyond the per-env limits so that every constraint is active somewhere
    L = tcs[:, :3]
    q0 = env.buffer('joints').t().cpu().numpy().copy()
    q0[: n // 2, :3] = (-L[: n // 2] + 1e-3 + rng.uniform(-2e-3, 2e-3, size=(n // 2, 3))).astype(np.float32)
    env.set_state(joints=q0)
    lim = env.action_space.high
    q = q0.copy()
    for _ in range(4):
        act = rng.uniform(-lim, lim, size=(n, 6)).astype(np.float32)
        act[:, :3] = -np.abs(act[:, :3]) if _ % 2 == 0 else act[:, :3]            # push towards the lower limits
        obs, *_rest = env.step(torch.tensor(act, device='cuda'))
        torch.cuda.synchronize()
        q = oracle_c.set_action_batch(q, act, np.arange(n, dtype=np.int32), L, False, 10)
        got = env.buffer('joints').t().cpu().numpy()
        assert np.array_equal(got, q)
        ref = oracle_c.fk_batch(tcs, q, np.arange(n, dtype=np.int32))
        assert np.abs(obs['achieved_goal'].cpu().numpy() - ref['tip']).max() < 1e-10
    low = (-L + 1e-3).astype(np.float32)
    assert (q[:, :3] >= low).all() and (q[:, :3] <= 0).all()
    assert (q[:, :3] == low).any(), "no env reached its own lower limit: the test would not see a wrong box"
    # the nominal box would have been wrong for a good part of the envs
    nominal_low = (-oracle_c.derive(oracle_py.DEFAULT_SYSTEMS[:1])[0][:3] + 1e-3).astype(np.float32)
    assert (q[:, :3] < nominal_low).any()


def test_domain_rand_length_resets_sample_inside_per_env_limits(oracle_c, oracle_py, vec_factory):
    """domain_rand_length: L and L_c drawn at every reset; the goal / start sampling box and the nesting predicates
    (obs.py:59-63,198-200) use the env's own lengths; tips follow the env's own constants."""
    n = 8000
    r = 0.06
    env = vec_factory(n, domain_rand=0.05, domain_rand_length=r, seed=9)
    env.reset()
    torch.cuda.synchronize()
    P = env.buffer('tube_params').t().cpu().numpy()
    nom = oracle_c.derive(oracle_py.DEFAULT_SYSTEMS[:1])[0]
    for k in range(6):                                                   # L[3], Lc[3] inside the +-r box, actually spread
        ratio = P[:, k] / nom[k]
        assert ratio.min() >= 1 - r - 1e-12 and ratio.max() <= 1 + r + 1e-12 and ratio.std() > 0.4 * r
    L = P[:, :3]
    Lf = L.astype(np.float32)
    low = (-L + 1e-3).astype(np.float32)
    for name in ('joints', 'desired_joints'):
        q = env.buffer(name).t().cpu().numpy()
        b = q[:, :3]
        assert (b >= low).all() and (b <= 0).all(), name
        assert (b[:, 0] <= b[:, 1]).all() and (b[:, 1] <= b[:, 2]).all(), name
        assert ((b[:, 0] + Lf[:, 0]) >= (Lf[:, 1] + b[:, 1])).all() and ((b[:, 1] + Lf[:, 1]) >= (Lf[:, 2] + b[:, 2])).all(), name
        assert np.abs(q[:, 3:]).max() <= np.float32(np.pi)
    idx = np.arange(n, dtype=np.int32)
    ref = oracle_c.fk_batch(P, env.buffer('joints').t().cpu().numpy(), idx)
    assert np.abs(env.buffer('achieved_goal').t().cpu().numpy() - ref['tip']).max() < 1e-10
    refg = oracle_c.fk_batch(P, env.buffer('desired_joints').t().cpu().numpy(), idx)
    assert np.abs(env.buffer('desired_goal').t().cpu().numpy() - refg['tip']).max() < 1e-10
    # a step keeps using the env's own box (auto-reset envs excluded: they carry the next episode's draw)
    lim = env.action_space.high
    act = np.tile(np.array([-1e-3, -1e-3, -1e-3, 0, 0, 0], dtype=np.float32), (n, 1))
    q0 = env.buffer('joints').t().cpu().numpy().copy()
    _, _, done, _ = env.step(torch.tensor(act, device='cuda'))
    torch.cuda.synchronize()
    keep = ~done.bool().cpu().numpy()
    q1 = oracle_c.set_action_batch(q0, act, idx, L, False, 10)
    assert keep.sum() > n // 2 and np.array_equal(env.buffer('joints').t().cpu().numpy()[keep], q1[keep])
    # config check: lengths need domain_rand
    with pytest.raises(Exception):
        vec_factory(16, domain_rand=0.0, domain_rand_length=0.05)


def test_action_scale_equals_scaled_raw_actions(vec_factory):
    n = 4096
    a, b = vec_factory(n, seed=3, action_scale=True), vec_factory(n, seed=3)
    assert np.array_equal(a.action_space.high, np.ones(6, dtype=np.float32))
    assert np.array_equal(a.action_limits, b.action_space.high)
    for e in (a, b):
        e.reset()
    rng = np.random.RandomState(4)
    lim = b.action_space.high
    for _ in range(5):
        u = rng.uniform(-1, 1, size=(n, 6)).astype(np.float32)
        raw = (u * lim).astype(np.float32)                               # float32 product, as the kernel forms it
        oa, ra, da, _ia = a.step(torch.tensor(u, device='cuda'))
        ob, rb, db, _ib = b.step(torch.tensor(raw, device='cuda'))
        torch.cuda.synchronize()
        assert torch.equal(a.buffer('joints'), b.buffer('joints'))
        assert torch.equal(oa['observation'], ob['observation']) and torch.equal(ra, rb) and torch.equal(da, db)
    assert a.fault_counters()['invalid_actions'] == 0 and b.fault_counters()['invalid_actions'] == 0


def test_invalid_actions_are_counted_and_optionally_clipped(vec_factory):
    """ctr_reach_env.py:131-132 asserts `not all NaN` and `action_space.contains(action)`; the batch counts such env-steps
    (any NaN or out-of-Box component) and, with action_clip, replaces them by the clipped / zeroed action."""
    n = 2000
    rng = np.random.RandomState(8)
    env, clip, ref = vec_factory(n, seed=2, auto_reset=False), vec_factory(n, seed=2, auto_reset=False, action_clip=True), \
        vec_factory(n, seed=2, auto_reset=False)
    for e in (env, clip, ref):
        e.reset()
        e.stats_reset()
    lim = env.action_space.high
    act = rng.uniform(-lim, lim, size=(n, 6)).astype(np.float32)
    bad_nan = rng.choice(n, 37, replace=False)
    bad_big = np.setdiff1d(rng.choice(n, 80, replace=False), bad_nan)
    act[bad_nan, rng.randint(0, 6, size=bad_nan.size)] = np.nan
    act[bad_big, 0] = 3 * lim[0]
    act[bad_big, 4] = -2 * lim[4]
    fixed = np.clip(np.nan_to_num(act, nan=0.0), -lim, lim)
    env.step(torch.tensor(act, device='cuda'))
    clip.step(torch.tensor(act, device='cuda'))
    ref.step(torch.tensor(fixed, device='cuda'))
    torch.cuda.synchronize()
    want = float(bad_nan.size + bad_big.size)
    assert env.fault_counters()['invalid_actions'] == want and clip.fault_counters()['invalid_actions'] == want
    assert ref.fault_counters()['invalid_actions'] == 0
    assert torch.equal(clip.buffer('joints'), ref.buffer('joints'))
    assert torch.equal(clip.buffer('achieved_goal'), ref.buffer('achieved_goal'))
    good = np.ones(n, dtype=bool)
    good[bad_nan] = good[bad_big] = False
    assert torch.equal(env.buffer('joints')[:, good], ref.buffer('joints')[:, good])     # valid actions are untouched
    env.stats_reset()
    assert env.fault_counters()['invalid_actions'] == 0


def test_scheduling_state_invariants_after_heavy_resets(vec_factory):
    """compute-sanitizer is closed on the build pool, so the lock-free bookkeeping of k_step is checked through its invariants
    after many steps in which most envs finish (4-step episodes, wide tolerance): the slot -> env map is a permutation inside
    every 640-slot window, every window's arrival counter is back at 0, the refill queues hold valid, distinct env ids within
    capacity, every prefetch slot is either ready for the env's next episode or queued for a refill, and no fault was counted."""
    n = 5000                                                  # 7 full windows + a ragged one
    env = vec_factory(n, seed=17, max_steps_per_episode=4)
    env.goal_tolerance.current_tol = 0.05
    env.reset()
    lim = torch.tensor(env.action_limits, device='cuda')
    g = torch.Generator(device='cuda')
    g.manual_seed(3)
    for step in range(25):
        env.step(((torch.rand((n, 6), device='cuda', generator=g) * 2 - 1) * lim).contiguous())
        if step % 6 == 5:
            mask = torch.zeros(n, dtype=torch.uint8, device='cuda')
            mask[step::11] = 1
            env.reset(mask=mask)                              # explicit resets interleaved with the prefetch route
    torch.cuda.synchronize()
    perm = env.debug_buffer(0).cpu().numpy()
    for lo in range(0, n, 640):
        hi = min(lo + 640, n)
        assert np.array_equal(np.sort(perm[lo:hi]), np.arange(lo, hi)), "window %d is not a permutation" % (lo // 640)
    assert int(env.debug_buffer(1).abs().sum()) == 0          # arrival counters reset themselves
    ctr = env.debug_buffer(2).cpu().numpy()
    idx = env.debug_buffer(7)
    cap = env.debug_buffer(3).numel()
    consume = idx % 3                                         # the queue the NEXT launch drains = produced by the last one
    cnt = int(ctr[consume])
    assert 0 < cnt <= cap
    q = env.debug_buffer(3 + consume).cpu().numpy()[:cnt]
    assert q.min() >= 0 and q.max() < n and np.unique(q).size == q.size
    tags = env.debug_buffer(6).cpu().numpy()
    ep = env.buffer('episode').cpu().numpy().astype(np.int64)
    # a slot is tagged with the episode it was drawn for: ep + 1 (ready for the next reset), ep (just consumed: its refill
    # must be waiting in the queue the next launch drains) or 0 (invalidated)
    assert np.all((tags == 0) | (tags == ep) | (tags == ep + 1))
    assert np.isin(np.nonzero(tags != ep + 1)[0], q).all()
    assert (tags == ep + 1).mean() > 0.5                      # most slots are filled ahead of time
    st = env.stats(reduce=False)
    assert st['resets'] > 5 * n and env.fault_counters() == {'fk_failed': 0.0, 'goal_stuck': 0.0, 'invalid_actions': 0.0}


def _spans_from_tables(S, EI, Ux):
    """The reference's tables -> what the solver integrates (model.py:137-151): spans [S_k-1, S_k - 1e-6] (sorted), patterns."""
    t0, tf, pat = [], [], []
    prev = 0.0
    for k in range(len(S)):
        a, b = prev, float(S[k]) - 1e-6
        if b < a:
            a, b = b, a
        prev = float(S[k])
        if a == b:
            continue
        t0.append(a)
        tf.append(b)
        pat.append(sum(1 << i for i in range(3) if EI[i, k] != 0) | sum(8 << i for i in range(3) if Ux[i, k] != 0))
    return t0, tf, pat


def test_segment_tables_direct_bit_exact(golden, oracle_c, tube_constants, vec_factory):
    """The section table the CUDA solver builds (build_sections) against the reference's Segment objects -- not through the
    tips: the fixture's 247 tables (incl. ties, limits, sub-1e-5 sections) and 40 000 random configurations via the C oracle,
    whose tables are themselves bit-exact against the fixture (tests/test_oracle.py)."""
    env = vec_factory(16, select_systems=[0, 1, 2, 3])
    g = golden('segment_reference')
    t0, tf, pat, kept, zero = (x.cpu().numpy() for x in env.segment_tables(g['joints'].astype(np.float64), system=g['system'].astype(np.int32)))
    for i, (q, sysid, tab) in enumerate(zip(g['joints'], g['system'], g['table'])):
        m = int(tab[0])
        T = tab[1:].reshape(13, 6)
        w0, wf, wp = _spans_from_tables(T[0, :m], T[1:4, :m], T[7:10, :m])
        assert kept[i] == len(w0) and kept[i] + zero[i] == m, i
        assert np.array_equal(t0[i, :kept[i]], w0) and np.array_equal(tf[i, :kept[i]], wf) and pat[i, :kept[i]].tolist() == wp, i
    rng = np.random.RandomState(31)
    n = 40000
    sysid = rng.randint(0, 4, size=n).astype(np.int32)
    L = tube_constants[sysid, :3]
    beta = np.sort(-rng.uniform(0, 1, size=(n, 3)) * (L - 1e-3), axis=1)
    beta[::7] = np.round(beta[::7], 3)                         # breakpoints an exact multiple of 1e-5 apart
    beta[::11, 1] = beta[::11, 0]                              # ties
    q = np.concatenate((beta.astype(np.float32).astype(np.float64), np.zeros((n, 3))), axis=1)
    t0, tf, pat, kept, zero = (x.cpu().numpy() for x in env.segment_tables(q, system=sysid))
    for i in range(n):
        S, EI, GJ, Ux, Uy = oracle_c.segment(tube_constants[sysid[i]], q[i, :3])
        w0, wf, wp = _spans_from_tables(S, EI, Ux)
        assert kept[i] == len(w0) and np.array_equal(t0[i, :kept[i]], w0) and np.array_equal(tf[i, :kept[i]], wf) \
            and pat[i, :kept[i]].tolist() == wp, i


def test_random_parts_of_reset_match_reference_distributions(golden, oracle_c, oracle_py, vec_factory):
    """Two-sample Kolmogorov-Smirnov bounds against draws of the REFERENCE's own sample_goal / sample_parameters
    (oracle/make_golden_stats.py).  33 marginals, Bonferroni at 1e-3 overall: D < 2.35 sqrt((n + m) / (n m))."""
    from scipy.stats import ks_2samp
    st = golden('stat_reference')
    crit = lambda a, b: 2.35 * np.sqrt((len(a) + len(b)) / (len(a) * len(b)))
    n = 32000
    env = vec_factory(n, select_systems=[0, 1, 2, 3], seed=77)
    env.reset()
    torch.cuda.synchronize()
    sysid = env.buffer('system').cpu().numpy()
    worst = 0.0
    for name in ('desired_joints', 'joints'):
        q = env.buffer(name).t().cpu().numpy()
        for s in range(4):
            mine = q[sysid == s]
            ref = st['goal_joints'][s]
            assert len(mine) > 6000
            for k in range(6):
                D = ks_2samp(mine[:, k], ref[:, k]).statistic
                worst = max(worst, D / crit(mine[:, k], ref[:, k]))
                assert D < crit(mine[:, k], ref[:, k]), (name, s, k, D)
            for a, b in ((1, 0), (2, 1)):                       # the constrained differences, not only the marginals
                D = ks_2samp(mine[:, a] - mine[:, b], ref[:, a] - ref[:, b]).statistic
                assert D < crit(mine[:, a], ref[:, a]), (name, s, a, b, D)
    # domain randomisation (model_utils.py:5-35): U_x directly, EI / GJ through Tube.py:14-15 from the reference's d_i, d_o, E, G
    env2 = vec_factory(20000, domain_rand=0.05, seed=78)
    env2.reset()
    torch.cuda.synchronize()
    P = env2.buffer('tube_params').t().cpu().numpy()
    nom = oracle_c.derive(oracle_py.DEFAULT_SYSTEMS[:1])[0]
    R = st['param_ratio'].astype(np.float64)                    # [10000][tube][d_i, d_o, E, G, U_x]
    sys0 = oracle_py.DEFAULT_SYSTEMS[0]
    for i in range(3):
        di, do = sys0[i, 2], sys0[i, 3]
        d4 = ((R[:, i, 1] * do) ** 4 - (R[:, i, 0] * di) ** 4) / (do ** 4 - di ** 4)
        for mine, ref, what in ((P[:, 12 + i] / nom[12 + i], R[:, i, 4], 'Ux'), (P[:, 6 + i] / nom[6 + i], R[:, i, 2] * d4, 'EI'),
                                (P[:, 9 + i] / nom[9 + i], R[:, i, 3] * d4, 'GJ')):
            D = ks_2samp(mine, ref).statistic
            worst = max(worst, D / crit(mine, ref))
            assert D < crit(mine, ref), (what, i, D)
    assert np.array_equal(P[:, :6], np.tile(nom[:6], (20000, 1))) and np.array_equal(st['param_fixed'][:, 0], nom[:3])
    assert worst > 0.05                                         # the statistic is being computed on something


@pytest.mark.parametrize('extra', [dict(), dict(domain_rand=0.05, domain_rand_length=0.03, select_systems=[0, 1, 2, 3])])
def test_checkpoint_restore_reproduces_the_run(vec_factory, extra):
    """SURVEY.md 5 (checkpoint / resume): a restored handle continues bit for bit -- through auto-resets, i.e. the RNG position
    (episode counters), the prefetched next episodes, the refill queues and the per-env tube parameters all travel."""
    n = 3000
    kw = dict(seed=21, max_steps_per_episode=6, **extra)
    a = vec_factory(n, **kw)
    a.goal_tolerance.current_tol = 0.04
    a.reset()
    g = torch.Generator(device='cuda')
    g.manual_seed(9)
    lim = torch.tensor(a.action_limits, device='cuda')
    acts = [((torch.rand((n, 6), device='cuda', generator=g) * 2 - 1) * lim).contiguous() for _ in range(20)]
    for k in range(12):
        a.step(acts[k])
    blob = a.checkpoint()
    want = []
    for k in range(12, 20):
        obs, rew, done, info = a.step(acts[k])
        want.append((obs['observation'].clone(), obs['desired_goal'].clone(), rew.clone(), done.clone()))
    b = vec_factory(n, **kw)                 # fresh handle, never reset
    b.restore(blob)
    assert b.goal_tolerance.get_tol() == 0.04
    for k in range(12, 20):
        obs, rew, done, info = b.step(acts[k])
        w = want[k - 12]
        assert torch.equal(obs['observation'], w[0]) and torch.equal(obs['desired_goal'], w[1]), k
        assert torch.equal(rew, w[2]) and torch.equal(done, w[3]), k
    for name in ('joints', 'episode', 't', 'system', 'desired_joints', 'starting_joints', 'tube_params'):
        assert torch.equal(a.buffer(name), b.buffer(name)), name
    sa, sb = a.stats(reduce=False), b.stats(reduce=False)
    for k in ('episodes', 'successes', 'resets', 'env_steps', 'sum_return', 'sum_length'):
        assert sa[k] == sb[k], k                                  # (float sums and the draw counter depend on atomic order / on
    assert abs(sa['sum_error'] - sb['sum_error']) < 1e-9 * sa['sum_error']   #  which launch happened to run a queued refill)
    small = vec_factory(n // 2, **kw)
    with pytest.raises(Exception):
        small.restore(blob)


@pytest.mark.gpu
def test_work_sorted_hand_out_order_does_not_change_results(vec_factory, monkeypatch):
    """Large one-shot launches (ctr_fk from 65536 solves, ctr_fk_shape from 32768, ctr_jacobian from 9363 configurations) are
    solved in a work-sorted order built by a counting-sort pre-pass (build_order).  Same inputs on a handle with the order
    switched off ($CTR_FK_SORT=0, read at ctr_create): every output bit-identical, four systems mixed.  (Finite joints: a warp that
    holds a solve with NaN or |angle difference| >= 9e4 rad runs the library sincos for all its 10 solves, 1 ulp away from the polynomial
    one, so next to such a row the last bits depend on which rows share its warp -- in any order.)"""
    import torch
    n = 70000
    monkeypatch.setenv('CTR_FK_SORT', '1')
    a = vec_factory(n, select_systems=[0, 1, 2, 3], seed=9)
    monkeypatch.setenv('CTR_FK_SORT', '0')
    b = vec_factory(16, select_systems=[0, 1, 2, 3])
    a.reset()
    q = a.buffer('joints').t().to(torch.float64).contiguous()
    sysid = a.buffer('system').clone()
    ta, Ra, ia = a.forward_kinematics(q, system=sysid, return_info=True)
    tb, Rb, ib = b.forward_kinematics(q, system=sysid, return_info=True)
    assert torch.equal(ia, ib) and torch.equal(torch.nan_to_num(ta, nan=7.0), torch.nan_to_num(tb, nan=7.0))
    assert torch.equal(torch.nan_to_num(Ra, nan=7.0), torch.nan_to_num(Rb, nan=7.0))
    qs, ss = q[1000:41000].contiguous(), sysid[1000:41000].contiguous()
    sa, sb = a.backbone_shape(qs, system=ss), b.backbone_shape(qs, system=ss)
    assert torch.equal(sa[1], sb[1]) and torch.equal(sa[2], sb[2]) and torch.equal(sa[3], sb[3])
    mask = (torch.arange(270, device=qs.device)[None, :] < sa[1][:, None])[:, :, None]
    assert torch.equal(sa[0] * mask, sb[0] * mask)                      # rows past npts are unspecified
    qj, sj = q[50000:60000].contiguous(), sysid[50000:60000].contiguous()
    Ja, tja = a.jacobian(qj, system=sj)
    Jb, tjb = b.jacobian(qj, system=sj)
    assert torch.equal(Ja, Jb) and torch.equal(tja, tjb)
