"""ctypes front-end of oracle/libctr_oracle.so (the C oracle) -- TEST INFRASTRUCTURE ONLY."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build():
    subprocess.check_call(['make', '-s', '-C', _HERE])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, 'libctr_oracle.so')
        if not os.path.exists(path):
            build()
        _LIB = ctypes.CDLL(path)
        _LIB.ctr_oracle_fk_batch.restype = ctypes.c_int
        _LIB.ctr_oracle_step_batch.restype = ctypes.c_int
        _LIB.ctr_oracle_segment.restype = ctypes.c_int
    return _LIB


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def derive(systems):
    """systems f64[nsys][3][8] -> tc f64[nsys][18] = L, Lc, EI, GJ, Ux, Uy."""
    systems = np.ascontiguousarray(systems, dtype=np.float64).reshape(-1, 3, 8)
    tc = np.empty((systems.shape[0], 18))
    for s in range(systems.shape[0]):
        lib().ctr_oracle_derive(_p(systems[s]), _p(tc[s]))
    return tc


def fk_batch(tcs, joints, system=None, rtol=1e-3, atol=1e-6, nthreads=0):
    """joints [n][6] -> dict(tip[n][3], R[n][9], state[n][18], info[n][4] = nfev, accepted, rejected, sections)."""
    tcs = np.ascontiguousarray(tcs, dtype=np.float64)
    q = np.ascontiguousarray(joints, dtype=np.float64).reshape(-1, 6)
    n = q.shape[0]
    sysa = None if system is None else np.ascontiguousarray(np.broadcast_to(system, (n,)), dtype=np.int32)
    state = np.empty((n, 18))
    info = np.empty((n, 4), dtype=np.int32)
    bad = lib().ctr_oracle_fk_batch(_p(tcs), _p(sysa), _p(q), ctypes.c_int64(n), ctypes.c_double(rtol),
                                    ctypes.c_double(atol), _p(state), _p(info), ctypes.c_int(nthreads))
    return dict(tip=state[:, 6:9].copy(), R=state[:, 9:18].copy(), state=state, info=info, failed=bad)


def fk_shape(tc, q, rtol=1e-3, atol=1e-6):
    """One configuration -> (r [npts][3], tip_idx [3]): Model.r and tip_pos of the reference (model.py:66-68,166-172)."""
    r = np.zeros((270, 3))
    tip = np.zeros(3, dtype=np.int32)
    fn = lib().ctr_oracle_fk_shape
    fn.restype = ctypes.c_int
    n = fn(_p(np.ascontiguousarray(tc, dtype=np.float64)), _p(np.ascontiguousarray(q, dtype=np.float64)),
           ctypes.c_double(rtol), ctypes.c_double(atol), _p(r), _p(tip))
    if n < 0:
        raise RuntimeError("ctr_oracle_fk_shape failed")
    return r[:n].copy(), tip


def segment(tc, beta):
    S = np.zeros(9)
    tab = np.zeros((4, 3, 9))
    m = lib().ctr_oracle_segment(_p(np.ascontiguousarray(tc, dtype=np.float64)),
                                 _p(np.ascontiguousarray(beta, dtype=np.float64)), _p(S), _p(tab))
    return S[:m], tab[0, :, :m], tab[1, :, :m], tab[2, :, :m], tab[3, :, :m]


def limits(L, constrain_alpha):
    L = np.asarray(L, dtype=np.float64)
    lim = np.pi if constrain_alpha else np.inf
    low = np.concatenate((-L + 1e-3, np.full(3, -lim))).astype(np.float32)
    high = np.concatenate((np.zeros(3), np.full(3, lim))).astype(np.float32)
    return np.concatenate((low, high)), np.float32(float(L[1]) - float(L[0])), np.float32(float(L[2]) - float(L[1]))


def set_action_batch(joints, action, system, tube_lengths, constrain_alpha, n_substeps):
    j = np.ascontiguousarray(joints, dtype=np.float32).reshape(-1, 6).copy()
    a = np.ascontiguousarray(action, dtype=np.float32).reshape(-1, 6)
    n = j.shape[0]
    tl = np.asarray(tube_lengths, dtype=np.float64).reshape(-1, 3)
    lims = np.empty((tl.shape[0], 12), dtype=np.float32)
    dl = np.empty((tl.shape[0], 2), dtype=np.float32)
    for s in range(tl.shape[0]):
        lims[s], dl[s, 0], dl[s, 1] = limits(tl[s], constrain_alpha)
    sysa = None if system is None else np.ascontiguousarray(np.broadcast_to(system, (n,)), dtype=np.int32)
    lib().ctr_oracle_set_action_batch(_p(j), _p(a), _p(sysa), _p(lims), _p(dl), ctypes.c_int(n_substeps), ctypes.c_int64(n))
    return j


def step_batch(tc, L, joints, action, desired, t, tol, n_substeps=10, max_steps=150, constrain_alpha=False, nthreads=0):
    """In-place batched env step (constraints -> FK -> reward/done); returns achieved, reward, done."""
    lim, d10, d21 = limits(L, constrain_alpha)
    n = joints.shape[0]
    achieved = np.empty((n, 3))
    reward = np.empty(n)
    done = np.empty(n, dtype=np.uint8)
    bad = lib().ctr_oracle_step_batch(_p(np.ascontiguousarray(tc, dtype=np.float64)), _p(lim), ctypes.c_float(d10),
                                      ctypes.c_float(d21), ctypes.c_int(n_substeps), ctypes.c_int(max_steps),
                                      ctypes.c_double(tol), _p(joints), _p(action), _p(desired), _p(t), _p(achieved),
                                      _p(reward), _p(done), ctypes.c_int64(n), ctypes.c_int(nthreads))
    assert bad == 0
    return achieved, reward, done


def bvp_pass(tc, q, f=None, u_init=None, max_step=0.01, rtol=1e-3, atol=1e-6):
    """One shooting pass of the BVP model (CTR_Model.ode_solver, method 1): dict(tip[3], R[9], cost[5], y[20], steps)."""
    out = np.zeros(38)
    f = np.zeros(3) if f is None else np.ascontiguousarray(f, dtype=np.float64).reshape(3)
    u = np.zeros(5) if u_init is None else np.ascontiguousarray(u_init, dtype=np.float64).reshape(5)
    fn = lib().ctr_oracle_bvp_pass
    fn.restype = ctypes.c_int
    rc = fn(_p(np.ascontiguousarray(tc, dtype=np.float64)), _p(np.ascontiguousarray(q, dtype=np.float64)), _p(f), _p(u),
            ctypes.c_double(max_step), ctypes.c_double(rtol), ctypes.c_double(atol), _p(out))
    if rc:
        raise RuntimeError("ctr_oracle_bvp_pass failed")
    return dict(tip=out[:3].copy(), R=out[3:12].copy(), cost=out[12:17].copy(), y=out[17:37].copy(), steps=int(out[37]))


def bvp_jac_comp(tc, q, f=None, u_init=None, max_step=0.01, rtol=1e-3, atol=1e-6):
    """CTR_Model.jac / comp (forward differences over passes): J (3, 6), C (3, 3), tip (3,)."""
    J, Cm, tip = np.zeros((3, 6)), np.zeros((3, 3)), np.zeros(3)
    f = np.zeros(3) if f is None else np.ascontiguousarray(f, dtype=np.float64).reshape(3)
    u = np.zeros(5) if u_init is None else np.ascontiguousarray(u_init, dtype=np.float64).reshape(5)
    fn = lib().ctr_oracle_bvp_jac_comp
    fn.restype = ctypes.c_int
    rc = fn(_p(np.ascontiguousarray(tc, dtype=np.float64)), _p(np.ascontiguousarray(q, dtype=np.float64)), _p(f), _p(u),
            ctypes.c_double(max_step), ctypes.c_double(rtol), ctypes.c_double(atol), _p(J), _p(Cm), _p(tip))
    if rc:
        raise RuntimeError("ctr_oracle_bvp_jac_comp failed")
    return J, Cm, tip
