"""Offline model of the GPU kernel's lock-step cost (analysis tool, test infrastructure: uses the C oracle).

A warp carries 10 environments and runs  max-over-envs  loop iterations; an iteration costs one step attempt (A) plus the
initial-step block (I) when ANY of its environments starts a section.  This script rolls random-action trajectories with
the C oracle, records each env's per-section attempt counts, and evaluates slot-assignment policies that use the PREVIOUS
step's counts (what k_step can know): identity, sort by nfev, sort by the per-section signature.

    python oracle/lockstep_sim.py [n_envs] [steps]
"""
import ctypes
import sys

import numpy as np

import ctr_oracle as O
import ctr_oracle_c as C

A_COST, I_COST = 867.0, 435.0   # warp instructions per attempt pass / per initial-step block (profiles/r1s ncu source page)


def attempts(tc, q, nthreads=8):
    q = np.ascontiguousarray(q, dtype=np.float64)
    att = np.zeros((q.shape[0], 9), dtype=np.int32)
    C.lib().ctr_oracle_fk_attempts_batch(C._p(np.ascontiguousarray(tc)), C._p(q), ctypes.c_int64(q.shape[0]), C._p(att),
                                         ctypes.c_int(nthreads))
    return att


def warp_cost(att):
    """att [n][9] in slot order (n multiple of 10) -> total cost, ideal cost."""
    n = att.shape[0]
    tot = att.sum(1)
    nsec = (att > 0).sum(1)
    # iteration index at which each section starts: cumulative attempts before it
    start = np.cumsum(att, 1) - att                           # [n][9]
    T = int(tot.max()) + 1
    init = np.zeros((n, T), dtype=bool)
    rows = np.repeat(np.arange(n), 9).reshape(n, 9)
    m = att > 0
    init[rows[m], start[m]] = True
    w_iters = tot.reshape(-1, 10).max(1)
    w_init = init.reshape(-1, 10, T).any(1).sum(1)
    cost = (w_iters * A_COST + w_init * I_COST).sum() * 10    # per-lane-group units: a warp charges all 10 groups
    ideal = (tot * A_COST + nsec * I_COST).sum()
    return cost, ideal


def signature(att):
    key = np.zeros(att.shape[0], dtype=np.int64)
    for k in range(9):
        key = key * 8 + np.minimum(att[:, k], 7)
    return key


def windowed_perm(key, window):
    n = key.shape[0]
    perm = np.arange(n)
    for b in range(0, n, window):
        perm[b:b + window] = b + np.argsort(key[b:b + window], kind='stable')
    return perm


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 20480
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
    tc = C.derive(O.DEFAULT_SYSTEMS[:1])[0]
    L = tc[:3]
    rng = np.random.RandomState(0)
    q = np.stack([O.sample_goal(rng, L) for _ in range(n)]).astype(np.float32)
    lim = np.array([1e-3] * 3 + [np.deg2rad(5)] * 3, dtype=np.float32)
    prev = attempts(tc, q)
    res = {}
    for s in range(steps):
        act = rng.uniform(-lim, lim, size=(n, 6)).astype(np.float32)
        q = C.set_action_batch(q, act, None, [L], False, 10)
        cur = attempts(tc, q)
        pol = {'identity': np.arange(n)}
        nfev_prev = 6 * prev.sum(1) + 2 * (prev > 0).sum(1)
        for w in (640, 2560, 10240):
            pol['nfev/%d' % w] = windowed_perm(nfev_prev, w)
            pol['sig/%d' % w] = windowed_perm(signature(prev), w)
            pol['nsec,nfev/%d' % w] = windowed_perm((prev > 0).sum(1) * 1000 + nfev_prev, w)
        pol['oracle-sig (this step)/640'] = windowed_perm(signature(cur), 640)
        for k, p in pol.items():
            c, ideal = warp_cost(cur[p])
            res.setdefault(k, []).append(c / ideal)
        prev = cur
    print("mean attempts %.2f sections %.2f" % (cur.sum(1).mean(), (cur > 0).sum(1).mean()))
    for k, v in res.items():
        print("%-28s cost / ideal = %.3f" % (k, np.mean(v)))


if __name__ == '__main__':
    main()
