"""CtrReachVecEnv -- N CTR-Reach environments stepped by one CUDA launch (new batched front-end).

Host side of the drop-in: argument handling mirrors CtrReachEnv.__init__ (reference envs/ctr_reach_env.py:14-68);
all per-step work (Obs.set_action x n_substeps, Model.forward_kinematics, compute_reward, done/info, Obs.get_obs,
auto-reset with goal sampling) happens in libctr_b200.so.  torch is used for device memory views and streams only.
"""
import ctypes

import numpy as np
import torch

from .. import _lib
from ..spaces import Box, Dict
from .goal_tolerance import GoalTolerance

NUM_TUBES = 3
ZERO_TOL = 1e-4   # obs.py:16
EXT_TOL = 1e-3    # obs.py:14

_TUBE_ORDER = ('length', 'length_curved', 'diameter_inner', 'diameter_outer', 'stiffness', 'torsional_stiffness',
               'x_curvature', 'y_curvature')


class _DeviceBuffer(object):
    """Zero-copy view of a handle-owned device buffer through __cuda_array_interface__."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {'shape': tuple(shape), 'typestr': typestr, 'data': (int(ptr), False),
                                         'version': 2, 'strides': None}


def systems_table(ctr_systems_parameters, select_systems):
    """dict-of-dicts kwargs (reference __init__.py:7-70) -> float64 [n_selected][3][8] (ctr_reach_env.py:19-29)."""
    rows = []
    for system in ctr_systems_parameters:
        tubes = ctr_systems_parameters[system]
        rows.append([[float(tubes[t][k]) for k in _TUBE_ORDER] for t in tubes])
    table = np.array(rows, dtype=np.float64)
    if table.ndim != 3 or table.shape[1:] != (NUM_TUBES, 8):
        raise ValueError("each CTR system needs exactly 3 tubes with 8 parameters")
    return np.ascontiguousarray(table[list(select_systems)])


def build_spaces(tube_lengths, goal_tolerance, num_systems, extension_action_limit, rotation_action_limit):
    """Declarative spaces of the reference (obs.py:78-134, ctr_reach_env.py:55-62)."""
    rep_low, rep_high = [], []
    for length in np.amax(np.asarray(tube_lengths), axis=0):
        rep_low += [-1, -1, -length + ZERO_TOL]
        rep_high += [1, 1, 0]
    tail_low = [-0.2, -0.2, -0.2, goal_tolerance.final_tol]
    tail_high = [0.2, 0.2, 0.2, goal_tolerance.init_tol]
    if num_systems > 1:
        tail_low.append(0)
        tail_high.append(num_systems - 1)
    goal_box = lambda: Box(low=np.array([-0.1, -0.1, 0]), high=np.array([0.1, 0.1, 0.2]), dtype="float32")
    observation_space = Dict(dict(
        desired_goal=goal_box(), achieved_goal=goal_box(),
        observation=Box(low=np.array(rep_low + tail_low), high=np.array(rep_high + tail_high), dtype="float32")))
    lim = np.concatenate((np.full(NUM_TUBES, extension_action_limit), np.full(NUM_TUBES, np.deg2rad(rotation_action_limit))))
    action_space = Box(low=-lim, high=lim, dtype="float32")
    return observation_space, action_space


class CtrReachVecEnv(object):
    """Batched CTR-Reach.  kwargs = those of the reference's CtrReachEnv plus

    num_envs, device, integrator ('dp45' = scipy-RK45 parity mode | 'rk4'), precision ('f64' | 'f32'), rk4_steps,
    reward_mode ('sparse' = reference | 'dense' = -distance), auto_reset, seed, env_offset (global index of env 0,
    for multi-GPU shards: results do not depend on how the env range is partitioned), and three extensions that default to
    the reference's behaviour: domain_rand_length (+-fraction on the tube lengths at every reset, with per-env joint limits;
    needs domain_rand > 0), action_scale (actions in [-1, 1] x the action limits), action_clip (clip / zero invalid action
    components instead of passing them through).  Where the reference asserts on an action (ctr_reach_env.py:131-132) the
    batch counts: fault_counters()['invalid_actions'].

    Tensors returned by reset()/step() are views of device buffers that the next step()/reset() overwrites; layouts are
    (N, D) views of SoA [D][N] storage.
    """

    def __init__(self, num_envs, ctr_systems_parameters, goal_tolerance_parameters, noise_parameters,
                 joint_representation, initial_joints, constrain_alpha, extension_action_limit, rotation_action_limit,
                 max_steps_per_episode, n_substeps, evaluation, select_systems, resample_joints=True,
                 length_based_sample=False, domain_rand=0.0, device=0, integrator='dp45', precision='f64', rk4_steps=8,
                 reward_mode='sparse', auto_reset=True, seed=0, env_offset=0, rtol=1e-3, atol=1e-6, domain_rand_length=0.0,
                 action_scale=False, action_clip=False):
        assert joint_representation in ['egocentric', 'proprioceptive']
        if length_based_sample:
            # the reference branch reads a non-existent attribute and raises (ctr_reach_env.py:84-92)
            raise AttributeError("'CtrReachEnv' object has no attribute 'system_parameters' (length_based_sample)")
        if not torch.cuda.is_available():
            raise RuntimeError("CtrReachVecEnv needs a CUDA device: the step path has no CPU fallback")
        self.lib = _lib.load()
        self.num_envs = int(num_envs)
        self.device = torch.device('cuda', int(device) if not isinstance(device, torch.device) else device.index or 0)
        self.select_systems = list(select_systems)
        self.systems = systems_table(ctr_systems_parameters, self.select_systems)
        self.num_systems = self.systems.shape[0]
        self.tube_lengths = self.systems[:, :, 0].tolist()
        self.noise_parameters = noise_parameters          # stored, never applied (reference: obs.py:35-38)
        self.joint_representation = joint_representation
        self.max_steps_per_episode = max_steps_per_episode
        self.n_substeps = n_substeps
        self.constrain_alpha = constrain_alpha
        self.evaluation = evaluation
        self.resample_joints = resample_joints
        self.length_based_sample = length_based_sample
        self.domain_rand = domain_rand
        self.domain_rand_length = domain_rand_length
        self.action_scale, self.action_clip = bool(action_scale), bool(action_clip)
        self.initial_joints = np.asarray(initial_joints)
        self.extension_action_limit = extension_action_limit
        self.rotation_action_limit = rotation_action_limit
        self.reward_mode = reward_mode
        self.auto_reset = auto_reset
        self.goal_tolerance = GoalTolerance(goal_tolerance_parameters)
        self.observation_space, self.action_space = build_spaces(
            self.tube_lengths, self.goal_tolerance, self.num_systems, extension_action_limit, rotation_action_limit)
        self.action_limits = self.action_space.high.copy()      # joint deltas per step (float32), whatever the action convention
        if self.action_scale:                                   # a policy's normalised output
            self.action_space = Box(low=-np.ones(6), high=np.ones(6), dtype="float32")

        cfg = _lib.make_config(self.systems, n_substeps=n_substeps, max_steps_per_episode=max_steps_per_episode,
                               joint_representation=joint_representation, constrain_alpha=constrain_alpha,
                               extension_action_limit=extension_action_limit, rotation_action_limit=rotation_action_limit,
                               resample_joints=resample_joints, auto_reset=auto_reset, integrator=integrator,
                               precision=precision, rk4_steps=rk4_steps, reward_mode=reward_mode, rtol=rtol, atol=atol,
                               initial_tol=self.goal_tolerance.get_tol(), domain_rand=domain_rand, seed=seed,
                               env_offset=env_offset, domain_rand_length=domain_rand_length, action_scale=action_scale,
                               action_clip=action_clip)
        h = ctypes.c_void_p()
        _lib.check(self.lib.ctr_create(ctypes.byref(cfg), self.num_envs, self.device.index, ctypes.byref(h)))
        self.handle = h
        self.obs_dim = int(self.lib.ctr_obs_dim(h))
        N, D = self.num_envs, self.obs_dim
        self._buf = {}
        for name, field, rows, typestr in (
                ('joints', _lib.F_JOINTS, 6, '<f4'), ('desired_goal', _lib.F_DESIRED_GOAL, 3, '<f8'),
                ('desired_joints', _lib.F_DESIRED_JOINTS, 6, '<f4'), ('starting_joints', _lib.F_STARTING_JOINTS, 6, '<f4'),
                ('starting_position', _lib.F_STARTING_POSITION, 3, '<f8'), ('achieved_goal', _lib.F_ACHIEVED_GOAL, 3, '<f8'),
                ('orientation', _lib.F_ORIENTATION, 9, '<f8'), ('t', _lib.F_T, 0, '<i4'), ('system', _lib.F_SYSTEM, 0, '<i4'),
                ('episode', _lib.F_EPISODE, 0, '<u4'), ('observation', _lib.F_OBS, D, '<f8'), ('reward', _lib.F_REWARD, 0, '<f8'),
                ('done', _lib.F_DONE, 0, '|u1'), ('is_success', _lib.F_SUCCESS, 0, '|u1'), ('error', _lib.F_ERROR, 0, '<f8'),
                ('terminal_observation', _lib.F_TERMINAL_OBS, D, '<f8'), ('terminal_achieved', _lib.F_TERMINAL_ACHIEVED, 3, '<f8'),
                ('tube_params', _lib.F_TUBE_PARAMS, 18, '<f8'), ('nfev', _lib.F_NFEV, 0, '<i4')):
            p = ctypes.c_void_p()
            _lib.check(self.lib.ctr_buffer(h, field, ctypes.byref(p), None))
            shape = (rows, N) if rows else (N,)
            if typestr == '<u4':   # torch has no uint32 arithmetic; view the counter as int32
                typestr = '<i4'
            self._buf[name] = torch.as_tensor(_DeviceBuffer(p.value, shape, typestr), device=self.device)
        self._closed = False
        self._pending = None

    # ------------------------------------------------------------------------------------------------ helpers
    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def buffer(self, name):
        """SoA device tensor ([rows][N] or [N]) of a state / output field (zero copy)."""
        return self._buf[name]

    def _obs_dict(self):
        return {'observation': self._buf['observation'].t(), 'achieved_goal': self._buf['achieved_goal'].t(),
                'desired_goal': self._buf['desired_goal'].t()}

    # ------------------------------------------------------------------------------------------------ gym-like API
    def seed(self, seed=None):
        if seed is not None:
            _lib.check(self.lib.ctr_seed(self.handle, int(seed) & 0xFFFFFFFFFFFFFFFF))

    def reset(self, goal=None, system=None, mask=None):
        """Reset all envs (or those with mask != 0).  goal: (N, 3) float64 or None; system: (N,) int or None
        (reference reset(goal, system), ctr_reach_env.py:70-114)."""
        g = s = m = None
        if goal is not None:
            g = torch.as_tensor(goal, dtype=torch.float64, device=self.device).reshape(self.num_envs, 3).t().contiguous()
        if system is not None:
            s = torch.as_tensor(system, dtype=torch.int32, device=self.device).expand(self.num_envs).contiguous()
        if mask is not None:
            m = torch.as_tensor(mask, device=self.device).to(torch.uint8).contiguous()
        _lib.check(self.lib.ctr_set_tolerance(self.handle, float(self.goal_tolerance.get_tol())))
        _lib.check(self.lib.ctr_reset(self.handle, _lib.ptr(m), _lib.ptr(g), _lib.ptr(s), self._stream()))
        return self._obs_dict()

    def step_async(self, actions):
        a = actions
        if not (torch.is_tensor(a) and a.is_cuda and a.dtype == torch.float32):
            a = torch.as_tensor(np.asarray(actions), dtype=torch.float32).to(self.device, non_blocking=True)
        if a.shape == (self.num_envs, 6):
            sf, se = a.stride(1), a.stride(0)
        elif a.shape == (6, self.num_envs):
            sf, se = a.stride(0), a.stride(1)
        else:
            raise ValueError("actions must have shape (num_envs, 6) or (6, num_envs)")
        self._pending = a   # keep alive until the kernel has consumed it
        _lib.check(self.lib.ctr_set_tolerance(self.handle, float(self.goal_tolerance.get_tol())))
        _lib.check(self.lib.ctr_step_strided(self.handle, _lib.ptr(a), sf, se, self._stream()))

    def step_wait(self):
        b = self._buf
        info = {'is_success': b['is_success'], 'error': b['error'],
                'terminal_observation': b['terminal_observation'].t(), 'terminal_achieved_goal': b['terminal_achieved'].t()}
        if self.evaluation:   # ctr_reach_env.py:144-153
            info.update(errors_pos=b['error'], errors_orient=0, system_idx=b['system'],
                        position_tolerance=self.goal_tolerance.get_tol(), orientation_tolerance=0,
                        achieved_goal=b['achieved_goal'].t(), desired_goal=b['desired_goal'].t(),
                        starting_position=b['starting_position'].t(), q_desired=b['desired_joints'].t(),
                        q_achieved=b['joints'].t(), q_starting=b['starting_joints'].t())
        return self._obs_dict(), b['reward'], b['done'], info

    def step(self, actions):
        """actions: float32 CUDA tensor (N, 6) (any strides) or (6, N).  Returns (obs dict, reward (N,) f64, done (N,) u8,
        info dict of tensors).  With auto_reset, finished envs are already reset in the returned obs and their last
        observation is in info['terminal_observation']."""
        self.step_async(actions)
        return self.step_wait()

    def compute_reward(self, achieved_goal, desired_goal, info=None, out=None):
        """HER-style batched reward (ctr_reach_env.py:160-170); inputs (..., 3) float64 CUDA tensors."""
        a = torch.as_tensor(achieved_goal, dtype=torch.float64, device=self.device)
        d = torch.as_tensor(desired_goal, dtype=torch.float64, device=self.device)
        assert a.shape == d.shape
        lead = a.shape[:-1]
        a2 = a.reshape(-1, 3).t().contiguous()
        d2 = d.reshape(-1, 3).t().contiguous()
        m = a2.shape[1]
        if out is None:
            out = torch.empty(m, dtype=torch.float64, device=self.device)
        _lib.check(self.lib.ctr_compute_reward(self.handle, _lib.ptr(a2), _lib.ptr(d2), m,
                                               float(self.goal_tolerance.get_tol()), -1, _lib.ptr(out), self._stream()))
        return out.reshape(lead)

    def compute_reward_soa(self, achieved, desired, out, obs=None, tol=None):
        """Zero-copy variant on SoA [3][M] float64 tensors; optionally rewrites obs[9:12] (HER relabel)."""
        m = achieved.shape[1]
        tol = float(self.goal_tolerance.get_tol() if tol is None else tol)
        _lib.check(self.lib.ctr_relabel(self.handle, _lib.ptr(achieved), _lib.ptr(desired), m, tol, -1, _lib.ptr(out),
                                        _lib.ptr(obs), self._stream()))
        return out

    def forward_kinematics(self, joints, system=None, integrator=None, precision=None, return_info=False):
        """Batched Model.forward_kinematics (model.py:30-70): joints (n, 6) -> tip (n, 3) [, R (n, 9), info (n, 4)]."""
        q = torch.as_tensor(joints, dtype=torch.float64, device=self.device).reshape(-1, 6).t().contiguous()
        n = q.shape[1]
        s = None if system is None else torch.as_tensor(system, dtype=torch.int32, device=self.device).expand(n).contiguous()
        tip = torch.empty((3, n), dtype=torch.float64, device=self.device)
        R = torch.empty((9, n), dtype=torch.float64, device=self.device)
        info = torch.empty((4, n), dtype=torch.int32, device=self.device)
        _lib.check(self.lib.ctr_fk(self.handle, _lib.ptr(q), _lib.ptr(s), None, n,
                                   -1 if integrator is None else _lib.INTEGRATOR[integrator],
                                   -1 if precision is None else _lib.PRECISION[precision],
                                   _lib.ptr(tip), _lib.ptr(R), _lib.ptr(info), self._stream()))
        return (tip.t(), R.t(), info.t()) if return_info else tip.t()

    def segment_tables(self, joints, system=None):
        """Segment(...) of the reference (CTR_Python/Segment.py) as the solver uses it: joints (n, 6) -> t0 (n, 9), tf (n, 9),
        pattern (n, 9) int32, kept (n,), zero_length (n,)."""
        q = torch.as_tensor(joints, dtype=torch.float64, device=self.device).reshape(-1, 6).t().contiguous()
        n = q.shape[1]
        s = None if system is None else torch.as_tensor(system, dtype=torch.int32, device=self.device).expand(n).contiguous()
        t0 = torch.empty((9, n), dtype=torch.float64, device=self.device)
        tf = torch.empty((9, n), dtype=torch.float64, device=self.device)
        pat = torch.empty((9, n), dtype=torch.int32, device=self.device)
        cnt = torch.empty(n, dtype=torch.int32, device=self.device)
        _lib.check(self.lib.ctr_segment_tables(self.handle, _lib.ptr(q), _lib.ptr(s), None, n, _lib.ptr(t0), _lib.ptr(tf), _lib.ptr(pat),
                                               _lib.ptr(cnt), self._stream()))
        return t0.t(), tf.t(), pat.t(), cnt & 0xff, cnt >> 8

    def backbone_shape(self, joints, system=None, max_pts=270):
        """Batched Model.r / tip_pos (model.py:66-68,137-172): joints (n, 6) -> r (n, max_pts, 3) float64 (30 samples per
        section, rows past npts[i] are unspecified), npts (n,), tip_idx (n, 3), tip (n, 3).  split_backbone() gives r1, r2, r3."""
        q = torch.as_tensor(joints, dtype=torch.float64, device=self.device).reshape(-1, 6).t().contiguous()
        n = q.shape[1]
        s = None if system is None else torch.as_tensor(system, dtype=torch.int32, device=self.device).expand(n).contiguous()
        shape = torch.zeros((n, max_pts, 3), dtype=torch.float64, device=self.device)
        npts = torch.empty(n, dtype=torch.int32, device=self.device)
        tip_idx = torch.empty((3, n), dtype=torch.int32, device=self.device)
        tip = torch.empty((3, n), dtype=torch.float64, device=self.device)
        _lib.check(self.lib.ctr_fk_shape(self.handle, _lib.ptr(q), _lib.ptr(s), None, n, int(max_pts), _lib.ptr(shape),
                                         _lib.ptr(npts), _lib.ptr(tip_idx), _lib.ptr(tip), self._stream()))
        return shape, npts, tip_idx.t(), tip.t()

    @staticmethod
    def split_backbone(r, npts, tip_idx):
        """One configuration's (r [max_pts][3], npts, tip_idx [3]) -> r1, r2, r3 exactly as model.py:66-68 slices them."""
        r = r[:int(npts)]
        t0, t1, t2 = (int(v) for v in tip_idx)
        return r[t1:t0 + 1], r[t2:t1 + 1], r[:t2 + 1]

    def jacobian(self, joints, system=None, eps=1e-4):
        """Finite-difference position Jacobian (CTR_Python/CTR_Model.py:251-262 over Model.forward_kinematics): joints (n, 6)
        float64 -> J (n, 3, 6), tip (n, 3)."""
        q = torch.as_tensor(joints, dtype=torch.float64, device=self.device).reshape(-1, 6).t().contiguous()
        n = q.shape[1]
        s = None if system is None else torch.as_tensor(system, dtype=torch.int32, device=self.device).expand(n).contiguous()
        J = torch.empty((18, n), dtype=torch.float64, device=self.device)
        tip = torch.empty((3, n), dtype=torch.float64, device=self.device)
        _lib.check(self.lib.ctr_jacobian(self.handle, _lib.ptr(q), _lib.ptr(s), n, float(eps), _lib.ptr(J), _lib.ptr(tip),
                                         self._stream()))
        return J.t().reshape(n, 3, 6), tip.t()

    JACOBIAN_NOALLOC_MAX = 16384          # CTR_JACOBIAN_NOALLOC_MAX of include/ctr_b200.h

    def jacobian_noalloc(self, joints_soa, J_out, tip_out=None, system=None, eps=1e-4):
        """ctr_jacobian_noalloc: the Jacobian for callers that may not allocate (no counterpart in the reference).  Everything
        is the caller's: joints_soa float64 [6][n], J_out float64 [18][n], tip_out float64 [3][n] or None, system int32 [n]
        or None, all contiguous on this env's device.  The scratch is a static device array of the library shared by every
        call on the device (n <= JACOBIAN_NOALLOC_MAX; overlapping calls race).  No torch allocation happens here either."""
        n = joints_soa.shape[1]
        for t, rows, dt in ((joints_soa, 6, torch.float64), (J_out, 18, torch.float64), (tip_out, 3, torch.float64)):
            if t is not None and not (t.is_cuda and t.dtype == dt and t.is_contiguous() and tuple(t.shape) == (rows, n)):
                raise ValueError("jacobian_noalloc takes contiguous device tensors of shape (%d, n), %s" % (rows, dt))
        if system is not None and not (system.is_cuda and system.dtype == torch.int32 and system.is_contiguous() and system.numel() == n):
            raise ValueError("system must be a contiguous int32 device tensor of n elements")
        _lib.check(self.lib.ctr_jacobian_noalloc(self.handle, _lib.ptr(joints_soa), _lib.ptr(system), n, float(eps), _lib.ptr(J_out),
                                                 _lib.ptr(tip_out), self._stream()))

    def ik_dls(self, target, q0, system=None, lam=0.25, eps=1e-4, tol=1e-3, max_iters=500):
        """Damped-least-squares position IK of the reference's baseline controller (src/jacobian_controller.py:19-73), batched:
        target (n, 3), q0 (n, 6) -> q (n, 6), err (n,), iters (n,).  Joints are not clipped (the reference does not)."""
        q = torch.as_tensor(q0, dtype=torch.float64, device=self.device).reshape(-1, 6).t().contiguous().clone()
        n = q.shape[1]
        tg = torch.as_tensor(target, dtype=torch.float64, device=self.device).reshape(n, 3).t().contiguous()
        s = None if system is None else torch.as_tensor(system, dtype=torch.int32, device=self.device).expand(n).contiguous()
        err = torch.empty(n, dtype=torch.float64, device=self.device)
        iters = torch.empty(n, dtype=torch.int32, device=self.device)
        _lib.check(self.lib.ctr_ik_dls(self.handle, _lib.ptr(tg), _lib.ptr(q), _lib.ptr(s), n, float(lam), float(eps), float(tol),
                                       int(max_iters), _lib.ptr(err), _lib.ptr(iters), self._stream()))
        return q.t(), err, iters

    # ---- BVP back-end: CTR_Python/CTR_Model.py (the reference's full model with tip load) -------------------------------
    def _soa(self, x, rows, n=None, dtype=torch.float64):
        if x is None:
            return None
        t = torch.as_tensor(x, dtype=dtype, device=self.device)
        t = t.reshape(-1, rows) if n is None else t.expand(n, rows) if t.dim() == 2 else t.reshape(1, rows).expand(n, rows)
        return t.t().contiguous()

    def bvp_pass(self, joints, force=None, u_init=None, system=None, params=None, max_step=0.01):
        """CTR_Model(..., f, q, 0, Tol = max_step, 1).ode_solver(u_init) batched: joints (n, 6) float64 -> dict(tip (n, 3),
        R (n, 9), cost (n, 5), steps (n,)).  force (n, 3) or None, u_init (n, 5) or None (zeros)."""
        q = self._soa(joints, 6)
        n = q.shape[1]
        f, u = self._soa(force, 3, n), self._soa(u_init, 5, n)
        s = None if system is None else torch.as_tensor(system, dtype=torch.int32, device=self.device).expand(n).contiguous()
        p = self._soa(params, 18, n)
        tip = torch.empty((3, n), dtype=torch.float64, device=self.device)
        R = torch.empty((9, n), dtype=torch.float64, device=self.device)
        cost = torch.empty((5, n), dtype=torch.float64, device=self.device)
        steps = torch.empty(n, dtype=torch.int32, device=self.device)
        _lib.check(self.lib.ctr_bvp_pass(self.handle, _lib.ptr(q), _lib.ptr(f), _lib.ptr(u), _lib.ptr(s), _lib.ptr(p), n,
                                         float(max_step), _lib.ptr(tip), _lib.ptr(R), _lib.ptr(cost), _lib.ptr(steps), self._stream()))
        return dict(tip=tip.t(), R=R.t(), cost=cost.t(), steps=steps)

    def bvp_jacobian(self, joints, force=None, u_init=None, system=None, params=None, max_step=0.01, compliance=True):
        """CTR_Model.jac (and comp): J (n, 3, 6), C (n, 3, 3) or None, tip (n, 3)."""
        q = self._soa(joints, 6)
        n = q.shape[1]
        f, u = self._soa(force, 3, n), self._soa(u_init, 5, n)
        s = None if system is None else torch.as_tensor(system, dtype=torch.int32, device=self.device).expand(n).contiguous()
        p = self._soa(params, 18, n)
        J = torch.empty((18, n), dtype=torch.float64, device=self.device)
        C = torch.empty((9, n), dtype=torch.float64, device=self.device) if compliance else None
        tip = torch.empty((3, n), dtype=torch.float64, device=self.device)
        _lib.check(self.lib.ctr_bvp_jacobian(self.handle, _lib.ptr(q), _lib.ptr(f), _lib.ptr(u), _lib.ptr(s), _lib.ptr(p), n,
                                             float(max_step), _lib.ptr(J), _lib.ptr(C), _lib.ptr(tip), self._stream()))
        return J.t().reshape(n, 3, 6), (C.t().reshape(n, 3, 3) if compliance else None), tip.t()

    def bvp_solve(self, joints, force=None, system=None, params=None, max_step=0.01, f_tol=1e-3, max_iters=40):
        """The shooting problem of CTR_Model.minimize: dict(u (n, 5), tip (n, 3), cost (n, 5), residual (n,), iters (n,))."""
        q = self._soa(joints, 6)
        n = q.shape[1]
        f = self._soa(force, 3, n)
        s = None if system is None else torch.as_tensor(system, dtype=torch.int32, device=self.device).expand(n).contiguous()
        p = self._soa(params, 18, n)
        u = torch.zeros((5, n), dtype=torch.float64, device=self.device)
        tip = torch.empty((3, n), dtype=torch.float64, device=self.device)
        cost = torch.empty((5, n), dtype=torch.float64, device=self.device)
        res = torch.empty(n, dtype=torch.float64, device=self.device)
        iters = torch.empty(n, dtype=torch.int32, device=self.device)
        _lib.check(self.lib.ctr_bvp_solve(self.handle, _lib.ptr(q), _lib.ptr(f), _lib.ptr(s), _lib.ptr(p), n, float(max_step),
                                          float(f_tol), int(max_iters), _lib.ptr(u), _lib.ptr(tip), _lib.ptr(cost), _lib.ptr(res),
                                          _lib.ptr(iters), self._stream()))
        return dict(u=u.t(), tip=tip.t(), cost=cost.t(), residual=res, iters=iters)

    def bvp_ik_dls(self, target, q0, system=None, lam=0.25, tol=1e-3, max_iters=500, max_step=0.01):
        """src/jacobian_controller.py:19-53 (dls_ik_position_only) on the model it was written for: q (n, 6), err, iters."""
        q = self._soa(q0, 6).clone()
        n = q.shape[1]
        tg = self._soa(target, 3, n)
        s = None if system is None else torch.as_tensor(system, dtype=torch.int32, device=self.device).expand(n).contiguous()
        err = torch.empty(n, dtype=torch.float64, device=self.device)
        iters = torch.empty(n, dtype=torch.int32, device=self.device)
        _lib.check(self.lib.ctr_bvp_ik_dls(self.handle, _lib.ptr(tg), _lib.ptr(q), _lib.ptr(s), n, float(lam), float(tol),
                                           int(max_iters), float(max_step), _lib.ptr(err), _lib.ptr(iters), self._stream()))
        return q.t(), err, iters

    def set_tube_params(self, params):
        """Per-env derived constants (N, 18) = L, Lc, EI, GJ, Ux, Uy (x3 each) or None (model.py:20-28)."""
        if params is None:
            _lib.check(self.lib.ctr_set_tube_params(self.handle, None, self._stream()))
            return
        p = torch.as_tensor(params, dtype=torch.float64, device=self.device).reshape(self.num_envs, 18).t().contiguous()
        _lib.check(self.lib.ctr_set_tube_params(self.handle, _lib.ptr(p), self._stream()))
        torch.cuda.current_stream(self.device).synchronize()

    def set_state(self, joints=None, desired_goal=None, system=None, t=None):
        """Overwrite env state from host arrays (N, 6) f32 / (N, 3) f64 / (N,) / (N,) (tests, checkpoint restore)."""
        def soa(x, dt, cols):
            if x is None:
                return None
            x = np.asarray(x, dtype=dt)
            return np.ascontiguousarray(x.reshape(self.num_envs, cols).T) if cols else np.ascontiguousarray(x.reshape(self.num_envs))
        j, g = soa(joints, np.float32, 6), soa(desired_goal, np.float64, 3)
        s, tt = soa(system, np.int32, 0), soa(t, np.int32, 0)
        _lib.check(self.lib.ctr_set_state_host(self.handle, _lib.ptr(j), _lib.ptr(g), _lib.ptr(s), _lib.ptr(tt)))

    def update_goal_tolerance(self, timestep):
        self.goal_tolerance.update(timestep)
        _lib.check(self.lib.ctr_set_tolerance(self.handle, float(self.goal_tolerance.get_tol())))

    def get_goal_tolerance(self):
        return self.goal_tolerance.get_tol()

    def stats(self, reduce=True):
        """Episode statistics since the last stats_reset().  With torch.distributed initialised (one process per GPU)
        the 8 partial sums are all-reduced (the only collective on this path)."""
        out = (ctypes.c_double * 8)()
        _lib.check(self.lib.ctr_stats(self.handle, out, self._stream()))
        v = torch.tensor(list(out), dtype=torch.float64)
        from ..distributed import all_reduce_stats
        if reduce:
            v = all_reduce_stats(v, self.device)
        keys = ('episodes', 'successes', 'sum_error', 'sum_return', 'sum_length', 'env_steps', 'resets', 'goal_draws')
        d = dict(zip(keys, v.tolist()))
        ep = max(d['episodes'], 1.0)
        d['success_rate'] = d['successes'] / ep
        d['mean_error'] = d['sum_error'] / ep
        d['mean_episode_length'] = d['sum_length'] / ep
        return d

    def work_counters(self):
        """FK work since the last stats_reset(): dict(nfev, attempts, sections, fk_solves)."""
        out = (ctypes.c_double * 4)()
        _lib.check(self.lib.ctr_work_counters(self.handle, out, self._stream()))
        return dict(zip(('nfev', 'attempts', 'sections', 'fk_solves'), list(out)))

    def checkpoint(self):
        """The complete env state as one numpy uint8 array (ctr_checkpoint_save), including the goal-tolerance schedule position."""
        nbytes = int(self.lib.ctr_checkpoint_bytes(self.handle))
        buf = np.empty(nbytes + 16, dtype=np.uint8)
        _lib.check(self.lib.ctr_set_tolerance(self.handle, float(self.goal_tolerance.get_tol())))
        _lib.check(self.lib.ctr_checkpoint_save(self.handle, _lib.ptr(buf), nbytes))
        buf[nbytes:] = np.frombuffer(np.array([self.goal_tolerance.training_step, self.goal_tolerance.current_tol], dtype=np.float64).tobytes(), dtype=np.uint8)
        return buf

    def restore(self, buf):
        """Restore a checkpoint() taken from an env created with the same kwargs and num_envs."""
        buf = np.ascontiguousarray(buf, dtype=np.uint8)
        nbytes = int(self.lib.ctr_checkpoint_bytes(self.handle))
        _lib.check(self.lib.ctr_checkpoint_load(self.handle, _lib.ptr(buf), nbytes))
        step, tol = np.frombuffer(buf[nbytes:nbytes + 16].tobytes(), dtype=np.float64)
        self.goal_tolerance.training_step = int(step) if float(step).is_integer() else float(step)
        self.goal_tolerance.current_tol = float(tol)

    def fault_counters(self):
        """What the reference raises on, counted since the last stats_reset(): dict(fk_failed, goal_stuck, invalid_actions)."""
        out = (ctypes.c_double * 4)()
        _lib.check(self.lib.ctr_fault_counters(self.handle, out, self._stream()))
        return dict(zip(('fk_failed', 'goal_stuck', 'invalid_actions'), list(out)[:3]))

    def debug_buffer(self, which):
        """Copy of one of the step path's internal scheduling buffers (ctr_debug_buffer; invariant tests): 0 perm, 1 window
        counters, 2 queue counters, 3-5 refill queues, 6 slot tags (uint32 as int64), 7 -> the step index (int)."""
        p, n = ctypes.c_void_p(), ctypes.c_int64()
        _lib.check(self.lib.ctr_debug_buffer(self.handle, int(which), ctypes.byref(p), ctypes.byref(n)))
        if which == 7:
            return int(n.value)
        if not p.value:
            return None
        torch.cuda.synchronize(self.device)
        t = torch.as_tensor(_DeviceBuffer(p.value, (n.value,), '<i4'), device=self.device).clone()
        return (t.to(torch.int64) & 0xFFFFFFFF) if which == 6 else t

    def alloc_host_outputs(self):
        """Pinned host arrays for step_host(), laid out like the device's result block (achieved_goal | reward | error |
        observation | desired_goal as float64 rows, done | is_success as uint8 rows) so that ctr_step_host moves
        each chunk's results with two DMA transfers instead of seven."""
        N, D = self.num_envs, self.obs_dim
        blk = torch.empty((D + 8, N), dtype=torch.float64).pin_memory()
        flg = torch.empty((2, N), dtype=torch.uint8).pin_memory()
        b, f = blk.numpy(), flg.numpy()
        self._host_blocks = (blk, flg)           # keep the pinned storage alive
        self._packed_out = {'achieved_goal': b[0:3], 'reward': b[3], 'error': b[4], 'observation': b[5:5 + D],
                            'desired_goal': b[5 + D:8 + D], 'done': f[0], 'is_success': f[1]}
        return self._packed_out

    def alloc_host_outputs_f32(self):
        """The same block as alloc_host_outputs() in float32 (opt-in: half the bytes over PCIe; the reference DECLARES float32
        spaces, obs.py:89-132, but returns float64, so the float64 block stays the default contract)."""
        N, D = self.num_envs, self.obs_dim
        blk = torch.empty((D + 8, N), dtype=torch.float32).pin_memory()
        flg = torch.empty((2, N), dtype=torch.uint8).pin_memory()
        b, f = blk.numpy(), flg.numpy()
        self._host_blocks_f32 = (blk, flg)
        self._packed_out_f32 = {'achieved_goal': b[0:3], 'reward': b[3], 'error': b[4], 'observation': b[5:5 + D],
                                'desired_goal': b[5 + D:8 + D], 'done': f[0], 'is_success': f[1]}
        return self._packed_out_f32

    def step_host(self, action_soa, out, delta=False):
        """End-to-end step with HOST buffers (ctr_step_host): action_soa float32 [6][N] numpy (ideally pinned);
        out: dict of preallocated numpy arrays among observation [D][N], achieved_goal [3][N], desired_goal [3][N],
        reward [N], done [N] u8, is_success [N] u8, error [N].  Synchronous.
        delta=True (only with the block of alloc_host_outputs(), which the caller must leave untouched between steps): the
        same float64 content with fewer bytes moved -- goals are patched for the envs that were reset, the tolerance row is
        rewritten on the host when it changed (ctr_step_host_delta)."""
        g = lambda k: _lib.ptr(out.get(k))
        _lib.check(self.lib.ctr_set_tolerance(self.handle, float(self.goal_tolerance.get_tol())))
        if out is getattr(self, '_packed_out_f32', None):     # float32 block of alloc_host_outputs_f32()
            _lib.check(self.lib.ctr_step_host_packed_f32(self.handle, _lib.ptr(action_soa), self._host_blocks_f32[0].data_ptr(),
                                                         self._host_blocks_f32[1].data_ptr()))
            return out
        if delta:
            if out is not getattr(self, '_packed_out', None):
                raise ValueError("delta=True needs the block returned by alloc_host_outputs()")
            _lib.check(self.lib.ctr_step_host_delta(self.handle, _lib.ptr(action_soa), self._host_blocks[0].data_ptr(),
                                                    self._host_blocks[1].data_ptr()))
            return out
        if out is getattr(self, '_packed_out', None):         # the block layout of alloc_host_outputs()
            _lib.check(self.lib.ctr_step_host_packed(self.handle, _lib.ptr(action_soa), self._host_blocks[0].data_ptr(),
                                                     self._host_blocks[1].data_ptr()))
            return out
        _lib.check(self.lib.ctr_step_host(self.handle, _lib.ptr(action_soa), g('observation'), g('achieved_goal'),
                                          g('desired_goal'), g('reward'), g('done'), g('is_success'), g('error')))
        return out

    def stats_reset(self):
        _lib.check(self.lib.ctr_stats_reset(self.handle, self._stream()))

    def launch_count(self):
        return int(self.lib.ctr_launch_count(self.handle))

    # SB-style conveniences (stable_baselines VecEnv.get_attr / env_method / set_attr: one value per selected env)
    _PER_ENV = {'t': 't', 'system': 'system', 'desired_goal': 'desired_goal', 'desired_joints': 'desired_joints',
                'starting_joints': 'starting_joints', 'starting_position': 'starting_position', 'joints': 'joints',
                'achieved_goal': 'achieved_goal'}

    def _indices(self, indices):
        if indices is None:
            return list(range(self.num_envs))
        if isinstance(indices, (int, np.integer)):
            return [int(indices)]
        return [int(i) for i in indices]

    def get_attr(self, name, indices=None):
        """Per-env values for the attributes the reference keeps per env (t, system, desired_goal, desired_joints,
        starting_joints, starting_position; also joints, achieved_goal), read from the device; shared attributes
        (action_space, goal_tolerance, ...) are repeated."""
        idx = self._indices(indices)
        if name in self._PER_ENV:
            buf = self.buffer(self._PER_ENV[name])
            sel = torch.as_tensor(idx, device=self.device, dtype=torch.long)
            vals = (buf[:, sel].t() if buf.dim() == 2 else buf[sel]).cpu().numpy()
            return [v.copy() if isinstance(v, np.ndarray) else v.item() for v in vals]
        return [getattr(self, name) for _ in idx]

    def env_method(self, name, *args, indices=None, **kwargs):
        """The reference env's methods that make sense per env: compute_reward / get_goal_tolerance / update_goal_tolerance act
        on shared state and return one value per selected env; forward_kinematics-style calls go through the batched API."""
        idx = self._indices(indices)
        out = getattr(self, name)(*args, **kwargs)
        return [out for _ in idx]

    def render(self, mode='empty', **kwargs):
        return None

    def close(self):
        if not self._closed:
            self._closed = True
            torch.cuda.synchronize(self.device)
            self._buf.clear()
            self.lib.ctr_destroy(self.handle)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
