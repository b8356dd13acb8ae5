"""CPU oracle for the CTR-Reach env-step path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

A numpy / pure-Python restatement of the reference algorithm (keshaviyengar/gym-ctr-reach).
Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference`
leg may import this module, and only as the checker / the timed CPU baseline -- never as the
thing shipped.  The product path (gym-ctr-reach_b200/) never imports it and has no CPU fallback.

Parity status: PINNED.  `oracle/make_golden.py` runs the unmodified reference in the build
container (through oracle/ref_loader.py) and writes tests/golden/*.npz; tests/test_oracle.py
checks this restatement against those fixtures and against a subset of the reference's own
shipped (joint -> tip) vectors (saved_policies/**/evaluations*.csv).

Citations are relative to /root/reference/ctr_reach_envs/ ; the adaptive integrator is the
third-party scipy.integrate.solve_ivp(method='RK45') (un-pinned by the reference, scipy 1.18.1
in this image: scipy/integrate/_ivp/{rk,common,ivp}.py), restated in `_dp45_section`.
"""
import math

import numpy as np

NUM_TUBES = 3
EXT_TOL = 1e-3  # envs/obs.py:14

# ----------------------------------------------------------------------------------------------
# default registration kwargs (ctr_reach_envs/__init__.py:6-94)
# columns: length, length_curved, diameter_inner, diameter_outer, stiffness, torsional_stiffness,
#          x_curvature, y_curvature  (argument order of Tube, envs/CTR_Python/Tube.py:7-8)
# ----------------------------------------------------------------------------------------------
DEFAULT_SYSTEMS = np.array([
    [[431e-3, 103e-3, 0.7e-3, 1.10e-3, 10.25e+10, 18.79e+10, 21.3, 0.0],
     [332e-3, 113e-3, 1.4e-3, 1.8e-3, 68.6e+10, 11.53e+10, 13.1, 0.0],
     [174e-3, 134e-3, 2e-3, 2.4e-3, 16.96e+10, 14.25e+10, 3.5, 0.0]],
    [[370e-3, 45e-3, 0.3e-3, 0.4e-3, 50e+10, 2.3e+10, 15.8, 0.0],
     [305e-3, 100e-3, 0.7e-3, 0.9e-3, 50e+10, 2.3e+10, 9.27, 0.0],
     [170e-3, 100e-3, 1.2e-3, 1.5e-3, 50e+10, 2.3e+10, 4.37, 0.0]],
    [[309e-3, 145e-3, 0.7e-3, 1.1e-3, 75e+9, 25e+9, 13.52, 0.0],
     [275e-3, 114e-3, 1.4e-3, 1.8e-3, 75e+9, 25e+9, 11.68, 0.0],
     [173e-3, 173e-3, 1.83e-3, 2.39e-3, 75e+9, 25e+9, 10.8, 0.0]],
    [[150e-3, 100e-3, 1.0e-3, 2.4e-3, 5e+10, 2.3e+10, 15.82, 0.0],
     [100e-3, 21.6e-3, 3.0e-3, 3.8e-3, 5e+10, 2.3e+10, 11.8, 0.0],
     [70e-3, 8.8e-3, 4.4e-3, 5.4e-3, 5.0e+10, 2.3e+10, 20.04, 0.0]],
], dtype=np.float64)


def derive_tube_constants(system):
    """system f64[3][8] -> dict(L, Lc, EI, GJ, Ux, Uy) each f64[3].

    I = pi (do^4 - di^4) / 64, J = pi (do^4 - di^4) / 32 (Tube.py:14-15, math.pow);
    EI = E * I, GJ = G * J (Segment.py:57-59)."""
    system = np.asarray(system, dtype=np.float64)
    out = dict(L=system[:, 0].copy(), Lc=system[:, 1].copy(), Ux=system[:, 6].copy(), Uy=system[:, 7].copy())
    EI, GJ = np.empty(3), np.empty(3)
    for i in range(3):
        di, do, E, G = (float(v) for v in system[i, 2:6])
        J = (math.pi * (math.pow(do, 4) - math.pow(di, 4))) / 32
        I = (math.pi * (math.pow(do, 4) - math.pow(di, 4))) / 64
        EI[i] = E * I
        GJ[i] = G * J
    out['EI'], out['GJ'] = EI, GJ
    return out


# ----------------------------------------------------------------------------------------------
# segmentation (envs/CTR_Python/Segment.py:6-61), value-based restatement
# ----------------------------------------------------------------------------------------------
def segment(tc, beta):
    """Transition points -> (S[m], EI[3][m], GJ[3][m], Ux[3][m], Uy[3][m]).

    Segment.py:12-16: points = [0, beta, d_c, d_tip]; len = 1e-5 * floor(1e5 * diff(sort(points))).
    Segment.py:23-43: tube i is present on sorted interval k iff beta_i <= lo_k and hi_k <= d_tip_i,
    curved iff additionally d_c_i <= lo_k (the a/b/c += 1 adjustments only move across zero-length
    intervals, which :46-50 delete).  Segment.py:52-61: cumulative sum, shift by min(beta), keep
    the columns that end after the template (s > 0).  All arithmetic IEEE float64, this order."""
    beta = np.asarray(beta, dtype=np.float64)
    d_tip = tc['L'] + beta
    d_c = d_tip - tc['Lc']
    pts = np.sort(np.concatenate(([0.0], beta, d_c, d_tip)))
    length = 1e-5 * np.floor(1e5 * np.diff(pts))
    bmin = beta.min()
    S, cols = [], []
    cum = 0.0
    for k in range(9):
        if length[k] == 0:
            continue
        cum = cum + length[k]
        if cum + bmin > 0:
            S.append(cum + bmin)
            cols.append(k)
    m = len(S)
    EI, GJ, Ux, Uy = (np.zeros((3, m)) for _ in range(4))
    for j, k in enumerate(cols):
        lo, hi = pts[k], pts[k + 1]
        for i in range(3):
            if beta[i] <= lo and hi <= d_tip[i]:
                EI[i, j] = tc['EI'][i]
                GJ[i, j] = tc['GJ'][i]
                if d_c[i] <= lo:
                    Ux[i, j] = tc['Ux'][i]
                    Uy[i, j] = tc['Uy'][i]
    return np.array(S), EI, GJ, Ux, Uy


# ----------------------------------------------------------------------------------------------
# ODE right-hand side (envs/model.py:72-117)
# ----------------------------------------------------------------------------------------------
def ode_rhs(y, ux0, uy0, ei, gj):
    """y = [uz(3), alpha(3), r(3), R(9 row-major)] -> dy/ds (model.py:83-117)."""
    dy = np.empty(18)
    a = y[3:6]
    inv = 1 / (ei[0] + ei[1] + ei[2])
    ux = np.empty(3)
    uy = np.empty(3)
    for i in range(3):
        sx = 0.0
        sy = 0.0
        for j in range(3):
            c = math.cos(a[i] - a[j])
            s = math.sin(a[i] - a[j])
            sx = sx + (ei[j] * ux0[j] * c + ei[j] * uy0[j] * s)          # model.py:87-90
            sy = sy + (-ei[j] * ux0[j] * s + ei[j] * uy0[j] * c)         # model.py:91-94
        ux[i] = inv * sx
        uy[i] = inv * sy
    for j in range(3):
        if ei[j] == 0:                                                    # model.py:97-99
            dy[j] = 0.0
            dy[3 + j] = 0.0
        else:                                                             # model.py:101-102
            dy[j] = (ei[j] / gj[j]) * (ux[j] * uy0[j] - uy[j] * ux0[j])
            dy[3 + j] = y[j]
    R = y[9:18]
    u0, u1, u2 = ux[0], uy[0], y[0]
    dy[6], dy[7], dy[8] = R[2], R[5], R[8]                                # r' = R e3, model.py:108
    for r in range(3):                                                    # R' = R u_hat, model.py:107,109
        r0, r1, r2 = R[3 * r], R[3 * r + 1], R[3 * r + 2]
        dy[9 + 3 * r] = r1 * u2 - r2 * u1
        dy[10 + 3 * r] = -r0 * u2 + r2 * u0
        dy[11 + 3 * r] = r0 * u1 - r1 * u0
    return dy


# Dormand-Prince 5(4) tableau as scipy states it (rk.py: class RK45)
_C = (0.0, 1 / 5, 3 / 10, 4 / 5, 8 / 9, 1.0)
_A = ((),
      (1 / 5,),
      (3 / 40, 9 / 40),
      (44 / 45, -56 / 15, 32 / 9),
      (19372 / 6561, -25360 / 2187, 64448 / 6561, -212 / 729),
      (9017 / 3168, -355 / 33, 46732 / 5247, 49 / 176, -5103 / 18656))
_B = (35 / 384, 0.0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84)
_E = (-71 / 57600, 0.0, 71 / 16695, -71 / 1920, 17253 / 339200, -22 / 525, 1 / 40)
_P = np.array([
    [1, -8048581381 / 2820520608, 8663915743 / 2820520608, -12715105075 / 11282082432],
    [0, 0, 0, 0],
    [0, 131558114200 / 32700410799, -68118460800 / 10900136933, 87487479700 / 32700410799],
    [0, -1754552775 / 470086768, 14199869525 / 1410260304, -10690763975 / 1880347072],
    [0, 127303824393 / 49829197408, -318862633887 / 49829197408, 701980252875 / 199316789632],
    [0, -282668133 / 205662961, 2019193451 / 616988883, -1453857185 / 822651844],
    [0, 40617522 / 29380423, -110615467 / 29380423, 69997945 / 29380423]])
SAFETY, MIN_FACTOR, MAX_FACTOR = 0.9, 0.2, 10.0  # rk.py:8-11
RTOL, ATOL = 1e-3, 1e-6                          # solve_ivp defaults (model.py:149-151 passes none)


def _rms(x):
    return math.sqrt(float(np.dot(x, x))) / math.sqrt(x.size)  # common.py: norm


def _dp45_section(f, t0, tf, y0, rtol=RTOL, atol=ATOL, dense_final=True):
    """solve_ivp(RK45) from t0 to tf (tf >= t0), returning y(tf), nfev, accepted, rejected.

    Restates scipy: RungeKutta.__init__ (f0 + select_initial_step, common.py:68-134),
    RungeKutta._step_impl + rk_step (rk.py:14-178) and the t_eval path of ivp.py:700-728,
    whose last sample is the dense-output polynomial (rk.py: RkDenseOutput) at x = 1."""
    y = np.array(y0, dtype=np.float64)
    nfev = acc = rej = 0
    if tf == t0:  # base.py OdeSolver.step(): t == t_bound -> finished, constant output
        return y, 1, 0, 0  # __init__ still evaluates f0 once
    f0 = f(y)
    nfev += 1
    # ---- select_initial_step
    interval = abs(tf - t0)
    scale = atol + np.abs(y) * rtol
    d0 = _rms(y / scale)
    d1 = _rms(f0 / scale)
    h0 = 1e-6 if (d0 < 1e-5 or d1 < 1e-5) else 0.01 * d0 / d1
    h0 = min(h0, interval)
    f1 = f(y + h0 * f0)
    nfev += 1
    d2 = _rms((f1 - f0) / scale) / h0
    if d1 <= 1e-15 and d2 <= 1e-15:
        h1 = max(1e-6, h0 * 1e-3)
    else:
        h1 = (0.01 / max(d1, d2)) ** (1 / 5)
    h_abs = min(100 * h0, h1, interval)
    # ---- step loop
    t = t0
    fcur = f0
    K = np.empty((7, 18))
    while t != tf:
        min_step = 10 * abs(np.nextafter(t, np.inf) - t)
        if h_abs < min_step:
            h_abs = min_step
        rejected = False
        while True:
            if h_abs < min_step:
                raise RuntimeError("step size too small")
            t_new = t + h_abs
            if t_new - tf > 0:
                t_new = tf
            h = t_new - t
            h_abs = abs(h)
            K[0] = fcur
            for s in range(1, 6):
                dy = np.zeros(18)
                for j in range(s):
                    dy = dy + K[j] * _A[s][j]
                K[s] = f(y + dy * h)
            acc_b = np.zeros(18)
            for j in range(6):
                acc_b = acc_b + K[j] * _B[j]
            y_new = y + h * acc_b
            f_new = f(y_new)
            K[6] = f_new
            nfev += 6
            scale = atol + np.maximum(np.abs(y), np.abs(y_new)) * rtol
            err = np.zeros(18)
            for j in range(7):
                err = err + K[j] * _E[j]
            error_norm = _rms(err * h / scale)
            if error_norm < 1:
                if error_norm == 0:
                    factor = MAX_FACTOR
                else:
                    factor = min(MAX_FACTOR, SAFETY * error_norm ** -0.2)
                if rejected:
                    factor = min(1.0, factor)
                h_abs *= factor
                acc += 1
                break
            h_abs *= max(MIN_FACTOR, SAFETY * error_norm ** -0.2)
            rejected = True
            rej += 1
        if t_new == tf and dense_final:
            # t_eval[-1] == t_bound: value stored by solve_ivp is RkDenseOutput at x == 1
            Q = K.T.dot(_P)
            y_out = h * Q.dot(np.ones(4)) + y
        else:
            y_out = y_new
        t, y, fcur = t_new, (y_out if t_new == tf else y_new), f_new
    return y, nfev, acc, rej


def forward_kinematics(tc, joints, integrator='dp45', full=False, rtol=RTOL, atol=ATOL):
    """joints [b0,b1,b2,a0,a1,a2] -> tip xyz (model.py:30-70, 119-174).

    integrator 'dp45'  : in-file restatement of scipy RK45;
               'scipy' : call scipy.integrate.solve_ivp exactly as model.py:149-151 does (the
                         third-party dependency itself; used for the timed CPU baseline)."""
    q = np.asarray(joints)
    beta = q[0:3].astype(np.float64)                 # model.py:50-51 (float32 + int64 -> float64)
    alpha = q[3:6].astype(np.float64)                # model.py:61
    S, EI, GJ, Ux, Uy = segment(tc, beta)
    a1 = float(alpha[0])                             # model.py:57-60
    y = np.zeros(18)
    y[3:6] = alpha
    y[9:18] = [math.cos(a1), -math.sin(a1), 0, math.sin(a1), math.cos(a1), 0, 0, 0, 1]
    span = np.append([0.0], S)                       # model.py:137
    nfev = acc = rej = 0
    for k in range(len(S)):
        t0, tf = span[k], span[k + 1] - 1e-6         # model.py:141 (linspace end points)
        if tf < t0:                                  # model.py:145-151: re-sorted, (min, max)
            t0, tf = tf, t0
        rhs = lambda yy, k=k: ode_rhs(yy, Ux[:, k], Uy[:, k], EI[:, k], GJ[:, k])
        if integrator == 'scipy':
            from scipy.integrate import solve_ivp
            sol = solve_ivp(lambda s, yy: rhs(yy), (t0, tf), y, t_eval=np.linspace(t0, tf, 30))
            y = sol.y[:, -1].copy()
            nfev += sol.nfev
        else:
            y, n, a_, r_ = _dp45_section(rhs, t0, tf, y, rtol, atol)
            nfev += n
            acc += a_
            rej += r_
    if full:
        return dict(tip=y[6:9].copy(), R=y[9:18].copy(), uz=y[0:3].copy(), alpha=y[3:6].copy(),
                    nfev=nfev, accepted=acc, rejected=rej, sections=len(S))
    return y[6:9].copy()


# ----------------------------------------------------------------------------------------------
# joint limits / action application / goal sampling (envs/obs.py)
# ----------------------------------------------------------------------------------------------
def joint_limits(L, constrain_alpha):
    """float32 low/high of the joint Box (obs.py:50-76): beta in [-L + 1e-3, 0], alpha in +-inf or +-pi."""
    L = np.asarray(L, dtype=np.float64)
    lim = np.pi if constrain_alpha else np.inf
    low = np.concatenate((-L + EXT_TOL, np.full(3, -lim))).astype(np.float32)
    high = np.concatenate((np.zeros(3), np.full(3, lim))).astype(np.float32)
    return low, high


def set_action(joints, action, L, low, high, n_substeps=1):
    """obs.py:166-183 applied n_substeps times (ctr_reach_env.py:134-135); all arithmetic float32
    (NEP-50: python-float operands are weak)."""
    j = np.asarray(joints, dtype=np.float32).copy()
    a = np.asarray(action, dtype=np.float32)
    d10 = np.float32(float(L[1]) - float(L[0]))
    d21 = np.float32(float(L[2]) - float(L[1]))
    for _ in range(n_substeps):
        j = np.clip(j + a, low, high)
        j[0] = min(j[0], j[1])
        j[0] = max(j[0], np.float32(d10 + j[1]))
        j[1] = min(j[1], j[2])
        j[1] = max(j[1], np.float32(d21 + j[2]))
    return j


def goal_is_valid(joints, L):
    """Acceptance predicate of the rejection sampler (obs.py:198-200), float32 arithmetic."""
    b = np.asarray(joints, dtype=np.float32)
    Lf = [np.float32(float(v)) for v in L]
    ok = True
    for i in (1, 2):
        ok = ok and bool(b[i - 1] <= b[i]) and bool(np.float32(b[i - 1] + Lf[i - 1]) >= np.float32(Lf[i] + b[i]))
    return ok


def sample_goal(rng, L):
    """Rejection sampling of valid joints (obs.py:185-207). rng: np.random.RandomState."""
    low = np.concatenate((-np.asarray(L, dtype=np.float64) + EXT_TOL, np.full(3, -np.pi))).astype(np.float32)
    high = np.concatenate((np.zeros(3), np.full(3, np.pi))).astype(np.float32)
    for _ in range(1001):
        q = rng.uniform(low, high).astype(np.float32)
        if goal_is_valid(q, L):
            return q
    raise ValueError("Stuck sampling goals...")


# ----------------------------------------------------------------------------------------------
# observation / reward / flags / tolerance (envs/obs.py, obs_utils.py, ctr_reach_env.py, goal_tolerance.py)
# ----------------------------------------------------------------------------------------------
def joint_rep(joints, representation='egocentric'):
    """[cos a, sin a, b] x 3 (obs_utils.py:41-53, 69-80).

    proprioceptive: float32 joints stay float32, so the trig is float32 arithmetic.
    egocentric: prop2ego's np.diff(..., prepend=0) concatenates an int64 zero with the float32 joints, which promotes
    to float64 -- the first differences and the trig are float64 arithmetic on float32-valued joints."""
    j = np.asarray(joints, dtype=np.float32)
    b, a = j[:3], j[3:]
    if representation == 'egocentric':
        b = np.diff(b.astype(np.float64), prepend=0.0)
        a = np.diff(a.astype(np.float64), prepend=0.0)
    rep = []
    for bi, ai in zip(b, a):
        rep += [np.cos(ai), np.sin(ai), bi]
    return np.array(rep).astype(np.float64)


def get_obs(joints, desired, achieved, tol, system, num_systems=1, representation='egocentric'):
    """obs.py:136-164."""
    tail = [tol, system] if num_systems > 1 else [tol]
    return np.concatenate([joint_rep(joints, representation),
                           np.asarray(desired, dtype=np.float64) - np.asarray(achieved, dtype=np.float64),
                           np.array(tail, dtype=np.float64)])


def compute_reward(achieved, desired, tol):
    """ctr_reach_env.py:160-170: -(||a - d|| > tol) as float64 (values -0.0 / -1.0)."""
    d = np.linalg.norm(np.asarray(achieved) - np.asarray(desired), axis=-1)
    return -(d > tol).astype(np.float64)


def step_flags(achieved, desired, tol, t, max_steps):
    """reward, done, is_success, error (ctr_reach_env.py:139-140, 155-156); t is the post-increment count.

    The reference takes the reward distance with norm(axis=-1) (unfused (x^2 + y^2) + z^2) but is_success / error with
    the 1-D norm (a BLAS dot, whose fused/unfused summation is implementation defined): for 1-D inputs both calls are
    repeated here verbatim; batches use the axis form throughout (the two can differ in the last ulp of `error`)."""
    achieved, desired = np.asarray(achieved), np.asarray(desired)
    if achieved.ndim == 1:
        err = np.linalg.norm(desired - achieved)
    else:
        err = np.linalg.norm(desired - achieved, axis=-1)
    reward = compute_reward(achieved, desired, tol)
    done = (reward == 0) | (np.asarray(t) >= max_steps)
    return reward, done, err < tol, err


class GoalTolerance(object):
    """goal_tolerance.py:8-55 (constant / linear / decay schedule; `set_tol` pins the value)."""

    def __init__(self, p):
        self.init_tol, self.final_tol, self.N_ts = p['initial_tol'], p['final_tol'], p['N_ts']
        self.function, self.set_tol = p['function'], p['set_tol']
        assert self.function in ('constant', 'linear', 'decay')
        if self.function == 'linear':
            self.a = (self.final_tol - self.init_tol) / self.N_ts
            self.b = self.init_tol
        if self.function == 'decay':
            self.a = self.init_tol
            self.r = 1 - np.power((self.final_tol / self.init_tol), 1 / self.N_ts)
        self.current_tol = self.init_tol if self.set_tol == 0 else self.set_tol

    def update(self, timestep):
        if self.set_tol == 0:
            if self.function == 'linear' and timestep <= self.N_ts:
                self.current_tol = self.a * timestep + self.b
            elif self.function == 'decay' and timestep <= self.N_ts:
                self.current_tol = self.a * np.power(1 - self.r, timestep)
            else:
                self.current_tol = self.final_tol
        else:
            self.current_tol = self.set_tol

    def get_tol(self):
        return self.current_tol


def randomize_system(rng, system, r):
    """model_utils.py:5-35: uniform +-r on d_i, d_o, E, G, Ux; L, L_c, Uy untouched (drawn with r = 0)."""
    out = np.array(system, dtype=np.float64)
    for i in range(3):
        for c, rr in ((0, 0), (1, 0), (2, r), (3, r), (4, r), (5, r), (6, r), (7, 0)):
            v = system[i][c]
            out[i, c] = rng.uniform(v - v * rr, v + v * rr)
    return out


# ----------------------------------------------------------------------------------------------
# a single-env port of the step path, used only as the timed CPU baseline (bench.py)
# ----------------------------------------------------------------------------------------------
class OracleEnv(object):
    """reset/step of ctr_reach_env.py:70-158 on top of the functions above (default kwargs)."""

    def __init__(self, system=DEFAULT_SYSTEMS[0], tol=0.02, n_substeps=10, max_steps=150, seed=0,
                 integrator='scipy', constrain_alpha=False):
        self.tc = derive_tube_constants(system)
        self.L = self.tc['L']
        self.low, self.high = joint_limits(self.L, constrain_alpha)
        self.tol, self.n_substeps, self.max_steps = tol, n_substeps, max_steps
        self.rng = np.random.RandomState(seed)
        self.integrator = integrator
        lim = np.concatenate((np.full(3, 0.001), np.full(3, np.deg2rad(5)))).astype(np.float32)
        self.act_low, self.act_high = -lim, lim

    def sample_action(self):
        return self.rng.uniform(self.act_low, self.act_high).astype(np.float32)

    def reset(self):
        self.t = 0
        self.desired_joints = sample_goal(self.rng, self.L)
        self.desired_goal = forward_kinematics(self.tc, self.desired_joints, self.integrator)
        self.joints = sample_goal(self.rng, self.L)
        self.achieved = forward_kinematics(self.tc, self.joints, self.integrator)
        return get_obs(self.joints, self.desired_goal, self.achieved, self.tol, 0)

    def step(self, action):
        self.joints = set_action(self.joints, action, self.L, self.low, self.high, self.n_substeps)
        self.achieved = forward_kinematics(self.tc, self.joints, self.integrator)
        self.t += 1
        reward, done, success, err = step_flags(self.achieved, self.desired_goal, self.tol, self.t, self.max_steps)
        obs = get_obs(self.joints, self.desired_goal, self.achieved, self.tol, 0)
        return obs, float(reward), bool(done), {'is_success': bool(success), 'error': float(err)}
