"""Hindsight replay on the device: oracle/her_oracle.py vs ctr_her_* through the C-ABI.

CPU tests pin the oracle: Philox4x32-10 against the Random123 known-answer vectors, the relabelled reward against the
reference fixture (tests/golden/reward_reference.npz), and the 'future' strategy's structural properties.  The strategy
itself is stable_baselines' (third party, absent here): its parity is UNPINNED, see oracle/her_oracle.py.
GPU tests compare ring contents and sampled batches BIT-EXACTLY (they are copies, integer draws and one float64 norm).
"""
import numpy as np
import pytest

import her_oracle as H


# ------------------------------------------------------------------------------------------------ CPU: the oracle
def test_philox_known_answers():
    # Random123 kat_vectors, philox4x32 10 rounds
    kat = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff, 0xffffffff), (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
            (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kat:
        got = H.philox4x32(np.array([ctr], dtype=np.uint32), key)[0]
        assert tuple(int(v) for v in got) == want


def test_relabelled_reward_is_the_reference_compute_reward(golden):
    g = golden('reward_reference')
    for k, tol in enumerate(g['tol']):
        r = H.compute_reward(g['achieved'], g['desired'], tol)
        assert np.array_equal(r, g['reward'][k]) and np.array_equal(np.signbit(r), np.signbit(g['reward'][k]))


def _synthetic_ring(n=64, cap=40, steps=70, D=13, seed=0):
    rng = np.random.RandomState(seed)
    ring = H.Ring(n, cap, D)
    ring.begin(rng.randn(n, D), rng.randn(n, 3) * 0.05, rng.randn(n, 3) * 0.05)
    t = np.zeros(n, dtype=int)
    for _ in range(steps):
        t += 1
        done = ((rng.rand(n) < 0.08) | (t >= 25)).astype(np.uint8)
        t[done != 0] = 0
        ring.store(rng.randn(n, 6).astype(np.float32), rng.randn(n, D), rng.randn(n, 3) * 0.05, rng.randn(n, 3) * 0.05,
                   -rng.randint(0, 2, n).astype(float), done, rng.randn(n, D), rng.randn(n, 3) * 0.05)
    return ring


def test_future_strategy_properties():
    ring = _synthetic_ring()
    n, cap = ring.n, ring.capacity
    b = ring.sample(40000, n_goal=4, tol=0.03, seed=7, call=3)
    slot, env = b['src'] // n, b['src'] % n
    rel = b['goal_src'] >= 0
    k, L = ring.t_in_ep[slot, env], ring.ep_len[slot, env]
    L = np.where(L == 0, ring.cur_t[env], L)
    # the sampled slots are the stored ones, uniformly; the last transition of an episode is never relabelled
    assert ring.stored == cap and len(np.unique(slot)) == cap
    assert not rel[k >= L].any()
    # donors: same env, same episode, at or after the transition
    gslot, genv = b['goal_src'][rel] // n, b['goal_src'][rel] % n
    off = (gslot - slot[rel]) % cap
    assert np.array_equal(genv, env[rel]) and (off >= 0).all() and (k[rel] + off < L[rel]).all()
    assert np.array_equal(ring.t_in_ep[gslot, genv], k[rel] + off)
    assert np.array_equal(b['desired'][rel], ring.next_achieved[gslot, genv])
    # relabelled: reward recomputed from (next achieved, new goal), done cleared; originals untouched
    assert np.array_equal(b['reward'][rel], H.compute_reward(b['next_achieved'][rel], b['desired'][rel], 0.03))
    assert not b['done'][rel].any()
    assert np.array_equal(b['reward'][~rel], ring.reward[slot, env][~rel]) and np.array_equal(b['desired'][~rel], ring.desired[slot, env][~rel])
    # k / (k + 1) of the relabellable samples are relabelled
    frac = rel[k < L].mean()
    assert abs(frac - 0.8) < 0.01
    # a donor equal to the transition itself makes the goal reached: reward -0.0
    same = rel.copy()
    same[rel] = off == 0
    assert (b['reward'][same] == 0).all() and same.any()
    # different call numbers give different batches, the same call the same batch
    assert not np.array_equal(ring.sample(1000, 4, 0.03, 7, 4)['src'], b['src'][:1000])
    assert np.array_equal(ring.sample(1000, 4, 0.03, 7, 3)['src'], b['src'][:1000])


def test_open_and_wrapped_episodes():
    ring = _synthetic_ring(n=8, cap=10, steps=37)          # capacity shorter than an episode (25 steps)
    n = ring.n
    open_len = ring.cur_t
    newest = ring.newest
    for e in range(n):
        if open_len[e] > 0:                                  # the open episode's transitions carry ep_len == 0
            assert ring.ep_len[newest, e] == 0 and ring.t_in_ep[newest, e] == open_len[e]
    b = ring.sample(5000, 4, 0.03, 1, 0)
    rel = b['goal_src'] >= 0
    slot, env = b['src'] // n, b['src'] % n
    gslot = b['goal_src'][rel] // n
    # donors never run past the newest stored slot
    age_t = (newest - slot[rel]) % ring.capacity
    assert (((gslot - slot[rel]) % ring.capacity) <= age_t).all()


# ------------------------------------------------------------------------------------------------ GPU: parity
def _rollout_both(vec_factory, n, cap, steps, tol, **kw):
    import torch
    from ctr_reach_envs.her import HerReplay
    env = vec_factory(n, seed=5, **kw)
    env.goal_tolerance.current_tol = tol
    env.reset()
    rep = HerReplay(env, cap, n_sampled_goal=4, seed=11)
    rep.begin()
    host = lambda name: env.buffer(name).t().cpu().numpy().copy() if env.buffer(name).dim() == 2 else env.buffer(name).cpu().numpy().copy()
    ring = H.Ring(n, cap, env.obs_dim)
    ring.begin(host('observation'), host('achieved_goal'), host('desired_goal'))
    g = torch.Generator(device='cuda')
    g.manual_seed(3)
    lim = torch.tensor(env.action_space.high, device='cuda')
    for _ in range(steps):
        act = ((torch.rand((n, 6), device='cuda', generator=g) * 2 - 1) * lim).contiguous()
        env.step(act)
        rep.store(act)
        ring.store(act.cpu().numpy(), host('observation'), host('achieved_goal'), host('desired_goal'), host('reward'),
                   host('done'), host('terminal_observation'), host('terminal_achieved'), auto_reset=env.auto_reset)
    torch.cuda.synchronize()
    return env, rep, ring


def _ring_dev(rep, key, n, cap):
    return rep.records()[key].cpu().numpy()          # (capacity, N, ...) decoded from the record-major ring


@pytest.mark.gpu
@pytest.mark.parametrize('cap,steps,tol', [(24, 20, 0.02), (16, 45, 0.05), (8, 30, 1e-4)])
def test_store_matches_oracle(vec_factory, cap, steps, tol):
    n = 700
    env, rep, ring = _rollout_both(vec_factory, n, cap, steps, tol)
    assert (rep.stored, rep.newest) == (ring.stored, ring.newest)
    for key in ('obs', 'next_obs', 'achieved', 'next_achieved', 'desired', 'action', 'reward', 'done', 't_in_ep', 'ep_len'):
        dev, ora = _ring_dev(rep, key, n, cap), getattr(ring, key)
        live = np.zeros(cap, dtype=bool)
        live[[(ring.newest - a) % cap for a in range(ring.stored)]] = True
        assert np.array_equal(dev[live], ora[live]), key
    assert np.array_equal(rep.ring['cur_t'].cpu().numpy(), ring.cur_t)
    if tol >= 0.02:
        assert ring.done.any()                               # the finished-episode branch was exercised


@pytest.mark.gpu
@pytest.mark.parametrize('rewrite,reward_mode', [(False, 'sparse'), (True, 'sparse'), (True, 'dense')])
def test_sample_matches_oracle(vec_factory, rewrite, reward_mode):
    n, cap = 700, 16
    env, rep, ring = _rollout_both(vec_factory, n, cap, 40, 0.05, reward_mode=reward_mode)
    rep.rewrite_error = rewrite
    for call, m, tol in ((0, 30000, 0.05), (9, 257, 0.01)):
        got = rep.sample(m, tol=tol, call=call)
        want = ring.sample(m, 4, tol, 11, call, rewrite_error=rewrite, dense=(reward_mode == 'dense'))
        for key, w in want.items():
            d = got[key].cpu().numpy()
            d = d.T if d.ndim == 2 else d
            assert np.array_equal(d, w), key
            if key == 'reward':
                assert np.array_equal(np.signbit(d), np.signbit(w))
        assert (want['goal_src'] >= 0).mean() > 0.5
    flat = rep.flat(got)
    assert flat.shape == (257, env.obs_dim + 6)
    assert np.array_equal(flat[:, env.obs_dim:env.obs_dim + 3].cpu().numpy(), want['achieved'])   # SB order: obs, achieved, desired


@pytest.mark.gpu
def test_her_argument_errors(vec_factory):
    from ctr_reach_envs import _lib
    from ctr_reach_envs.her import HerReplay
    env = vec_factory(64)
    env.reset()
    rep = HerReplay(env, 4)
    with pytest.raises(RuntimeError):
        rep.sample(8)                                        # nothing stored yet
    rep.begin()
    import torch
    act = torch.zeros((64, 6), device='cuda')
    env.step(act)
    rep.store(act)
    with pytest.raises(ValueError):
        rep.store(torch.zeros((6, 7), device='cuda'))
    rep._c.n_envs = 63
    with pytest.raises(_lib.CtrError):
        rep.sample(8)
