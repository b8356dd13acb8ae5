"""CtrReachEnv -- the reference's single-environment gym API (envs/ctr_reach_env.py:13-210) as an N = 1 view of the
CUDA batched environment.  Same constructor kwargs, same return types (numpy arrays, python/numpy scalars), same
assertion behaviour; every numerical step runs in libctr_b200.so."""
import numpy as np

from .vec_env import CtrReachVecEnv

NUM_TUBES = 3


class _TrigObjView(object):
    """Stands in for `env.trig_obj` (reference Obs): exposes `.joints` and `.obs` read from the device."""

    def __init__(self, env):
        self._env = env
        self.obs = None

    @property
    def joints(self):
        return self._env._vec.buffer('joints')[:, 0].cpu().numpy()

    @joints.setter
    def joints(self, value):
        self._env._vec.set_state(joints=np.asarray(value, dtype=np.float32).reshape(1, 6))


class _ModelView(object):
    """Stands in for `env.model`: forward_kinematics(joint, system) -> tip xyz (model.py:30-70), and the backbone of the LAST
    solve as the reference keeps it for its renderer and analysis scripts -- `r` (30 samples per section) and its per-tube
    slices `r1`, `r2`, `r3` (model.py:66-68, read at ctr_reach_env.py:178-182 and src/utils.py:52-54).  The samples come from
    ctr_fk_shape (the integrator's dense output) and are computed on first access after a solve, not on every step."""

    def __init__(self, env):
        self._env = env
        self._last = None          # (joints float64[6], system) of the last forward-kinematics solve
        self._shape = None

    def _note(self, joint, system):
        self._last = (np.asarray(joint, dtype=np.float64).reshape(6).copy(), int(system))
        self._shape = None

    def forward_kinematics(self, joint, system=0, **kwargs):
        self._note(joint, system)
        return self._env._vec.forward_kinematics(np.asarray(joint, dtype=np.float64).reshape(1, 6),
                                                 system=[int(system)])[0].cpu().numpy()

    def _backbone(self):
        if self._last is None:
            raise AttributeError("'Model' object has no attribute 'r' (no forward_kinematics call yet)")
        if self._shape is None:
            q, system = self._last
            r, npts, tip_idx, _ = self._env._vec.backbone_shape(q.reshape(1, 6), system=[system])
            r, n, idx = r[0].cpu().numpy(), int(npts[0]), tip_idx[0].cpu().numpy()
            self._shape = (r[:n].copy(),) + tuple(a.copy() for a in self._env._vec.split_backbone(r, n, idx))
        return self._shape

    r = property(lambda self: self._backbone()[0])
    r1 = property(lambda self: self._backbone()[1])
    r2 = property(lambda self: self._backbone()[2])
    r3 = property(lambda self: self._backbone()[3])


class CtrReachEnv(object):
    metadata = {'render.modes': []}
    reward_range = (-float('inf'), float('inf'))
    spec = None

    def __init__(self, ctr_systems_parameters, goal_tolerance_parameters, noise_parameters, joint_representation,
                 initial_joints, constrain_alpha, extension_action_limit, rotation_action_limit, max_steps_per_episode,
                 n_substeps, evaluation, select_systems, resample_joints=True, length_based_sample=False,
                 domain_rand=0.0, **b200_options):
        b200_options.setdefault('auto_reset', False)   # gym semantics: the caller resets
        self._vec = CtrReachVecEnv(1, ctr_systems_parameters, goal_tolerance_parameters, noise_parameters,
                                   joint_representation, initial_joints, constrain_alpha, extension_action_limit,
                                   rotation_action_limit, max_steps_per_episode, n_substeps, evaluation, select_systems,
                                   resample_joints=resample_joints, length_based_sample=False, domain_rand=domain_rand,
                                   **b200_options)
        v = self._vec
        self.ctr_system_parameters = v.systems
        self.select_systems = select_systems
        self.noise_parameters = noise_parameters
        self.joint_representation = joint_representation
        self.max_steps_per_episode = max_steps_per_episode
        self.n_substeps = n_substeps
        self.starting_joints = initial_joints
        self.desired_joints = initial_joints
        self.constrain_alpha = constrain_alpha
        self.evaluation = evaluation
        self.resample_joints = resample_joints
        self.length_based_sample = length_based_sample
        self.domain_rand = domain_rand
        self.model = _ModelView(self)
        self._host_out = None
        self.visualization = None
        self.t = 0
        self.trig_obj = _TrigObjView(self)
        self.observation_space = v.observation_space
        self.action_space = v.action_space
        self.extension_action_limit = extension_action_limit
        self.rotation_action_limit = rotation_action_limit
        self.system = 0
        self.goal_tolerance = v.goal_tolerance
        # ctr_reach_env.py:65-66: FK at the initial joints
        v.set_state(joints=np.asarray(initial_joints, dtype=np.float32).reshape(1, 6))
        self.starting_position = self.model.forward_kinematics(np.asarray(initial_joints, dtype=np.float64), 0)
        self.desired_goal = self.starting_position

    # ------------------------------------------------------------------------------------------------------
    def _pull(self, names):
        return {k: self._vec.buffer(k).cpu().numpy() for k in names}

    def _obs(self):
        b = self._pull(('observation', 'achieved_goal', 'desired_goal'))
        obs = {'observation': b['observation'][:, 0].copy(), 'achieved_goal': b['achieved_goal'][:, 0].copy(),
               'desired_goal': b['desired_goal'][:, 0].copy()}
        self.trig_obj.obs = obs
        return obs

    def reset(self, goal=None, system=None):
        if system is None and self.length_based_sample:
            # reference: ctr_reach_env.py:88 reads self.system_parameters, which does not exist
            raise AttributeError("'CtrReachEnv' object has no attribute 'system_parameters'")
        g = None if goal is None else np.asarray(goal, dtype=np.float64).reshape(1, 3)
        s = None if system is None else [int(system)]
        self._vec.reset(goal=g, system=s)
        self.t = 0
        b = self._pull(('system', 'desired_goal', 'desired_joints', 'starting_joints', 'starting_position'))
        self.system = int(b['system'][0])
        self.desired_goal = b['desired_goal'][:, 0].copy()
        if goal is None:
            self.desired_joints = b['desired_joints'][:, 0].copy()
        self.starting_joints = b['starting_joints'][:, 0].copy()
        self.starting_position = b['starting_position'][:, 0].copy()
        self.model._note(self.trig_obj.joints, self.system)      # ctr_reach_env.py:108: the last solve is the start state
        return self._obs()

    def seed(self, seed=None):
        if seed is not None:
            np.random.seed(seed)   # the reference seeds the global numpy RNG (ctr_reach_env.py:121-122)
            self._vec.seed(seed)

    def step(self, action):
        assert not np.all(np.isnan(action))
        assert self.action_space.contains(action)
        a = np.asarray(action, dtype=np.float32).reshape(6, 1)
        if self._host_out is None:
            self._host_out = self._vec.alloc_host_outputs()
        b = self._vec.step_host(a, self._host_out)               # one synchronous call: results arrive in pinned host arrays
        self.t += 1
        self.model._note(self.trig_obj.joints, self.system)      # ctr_reach_env.py:137
        obs = {'observation': b['observation'][:, 0].copy(), 'achieved_goal': b['achieved_goal'][:, 0].copy(),
               'desired_goal': b['desired_goal'][:, 0].copy()}
        self.trig_obj.obs = obs
        reward = np.float64(b['reward'][0])
        done = bool(b['done'][0])
        if self.evaluation:
            tol = self.goal_tolerance.get_tol()
            info = {'is_success': np.bool_(b['is_success'][0]), 'errors_pos': np.float64(b['error'][0]), 'errors_orient': 0,
                    'system_idx': self.select_systems[self.system], 'position_tolerance': tol, 'orientation_tolerance': 0,
                    'achieved_goal': obs['achieved_goal'], 'desired_goal': self.desired_goal,
                    'starting_position': self.starting_position, 'q_desired': self.desired_joints,
                    'q_achieved': self.trig_obj.joints, 'q_starting': self.starting_joints}
        else:
            info = {'is_success': np.bool_(b['is_success'][0]), 'error': np.float64(b['error'][0])}
        return obs, reward, done, info

    def compute_reward(self, achieved_goal, desired_goal, info):
        achieved_goal, desired_goal = np.asarray(achieved_goal), np.asarray(desired_goal)
        assert achieved_goal.shape == desired_goal.shape
        r = self._vec.compute_reward(achieved_goal, desired_goal).cpu().numpy()
        return r if r.ndim else np.float64(r)

    def render(self, mode='empty', **kwargs):
        """ctr_reach_env.py:172-182.  The reference's 'live' mode hands (t, achieved, desired, r1, r2, r3) to a matplotlib figure
        (ctr_3d_graph.py; no graphics stack on this box): here the same six values are returned as a dict, headless."""
        if mode == 'live':
            obs = self.trig_obj.obs if self.trig_obj.obs is not None else self._obs()
            return {'t': self.t, 'achieved_goal': obs['achieved_goal'], 'desired_goal': obs['desired_goal'],
                    'r1': self.model.r1, 'r2': self.model.r2, 'r3': self.model.r3}
        return None

    def close(self):
        self._vec.close()
        raise SystemExit(0)   # reference behaviour (ctr_reach_env.py:191)

    def print_parameters(self):
        print("----Observation and q_space----")
        print("----Goal tolerance parameters----")

    def update_goal_tolerance(self, timestep):
        self._vec.update_goal_tolerance(timestep)

    def get_goal_tolerance(self):
        return self.goal_tolerance.get_tol()
