"""ctypes binding of libctr_b200.so (include/ctr_b200.h).  There is no CPU fallback: if the CUDA library is
missing or no device is present, construction fails loudly."""
import ctypes
import os

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.environ.get('CTR_B200_LIB') or os.path.join(_ROOT, 'libctr_b200.so')   # override: kernel-variant experiments

ABI_VERSION = 2
MAX_SYSTEMS = 8

REP = {'egocentric': 0, 'proprioceptive': 1}
INTEGRATOR = {'dp45': 0, 'rk4': 1}
PRECISION = {'f64': 0, 'f32': 1}
REWARD = {'sparse': 0, 'dense': 1}

# enum ctr_field
(F_JOINTS, F_DESIRED_GOAL, F_DESIRED_JOINTS, F_STARTING_JOINTS, F_STARTING_POSITION, F_ACHIEVED_GOAL, F_ORIENTATION,
 F_T, F_SYSTEM, F_EPISODE, F_OBS, F_REWARD, F_DONE, F_SUCCESS, F_ERROR, F_TERMINAL_OBS, F_TERMINAL_ACHIEVED,
 F_TUBE_PARAMS, F_NFEV, F_STATS, F_COUNT) = range(21)

EXPORTS = ['ctr_last_error', 'ctr_abi_version', 'ctr_create', 'ctr_destroy', 'ctr_num_envs', 'ctr_obs_dim', 'ctr_buffer',
           'ctr_system_constants', 'ctr_set_tolerance', 'ctr_get_tolerance', 'ctr_seed', 'ctr_set_tube_params', 'ctr_fk',
           'ctr_fk_host', 'ctr_set_action', 'ctr_set_state_host', 'ctr_reset', 'ctr_step', 'ctr_step_strided',
           'ctr_step_host', 'ctr_step_host_packed', 'ctr_step_host_packed_f32', 'ctr_step_host_delta', 'ctr_reset_host', 'ctr_compute_reward', 'ctr_relabel', 'ctr_stats', 'ctr_stats_reset',
           'ctr_launch_count', 'ctr_work_counters', 'ctr_fault_counters', 'ctr_debug_buffer', 'ctr_segment_tables', 'ctr_checkpoint_bytes', 'ctr_checkpoint_save',
           'ctr_checkpoint_load', 'ctr_measure_fma_peak', 'ctr_her_begin', 'ctr_her_store',
           'ctr_her_sample', 'ctr_fk_shape', 'ctr_jacobian', 'ctr_jacobian_noalloc', 'ctr_ik_dls', 'ctr_bvp_pass', 'ctr_bvp_jacobian', 'ctr_bvp_solve',
           'ctr_bvp_ik_dls']


class CtrConfig(ctypes.Structure):
    _fields_ = [
        ('abi_version', ctypes.c_int32),
        ('n_systems', ctypes.c_int32),
        ('tube', ctypes.c_double * (MAX_SYSTEMS * 3 * 8)),
        ('n_substeps', ctypes.c_int32),
        ('max_steps_per_episode', ctypes.c_int32),
        ('joint_representation', ctypes.c_int32),
        ('constrain_alpha', ctypes.c_int32),
        ('extension_action_limit', ctypes.c_double),
        ('rotation_action_limit_deg', ctypes.c_double),
        ('resample_joints', ctypes.c_int32),
        ('auto_reset', ctypes.c_int32),
        ('integrator', ctypes.c_int32),
        ('precision', ctypes.c_int32),
        ('rk4_steps', ctypes.c_int32),
        ('reward_mode', ctypes.c_int32),
        ('rtol', ctypes.c_double),
        ('atol', ctypes.c_double),
        ('initial_tol', ctypes.c_double),
        ('domain_rand', ctypes.c_double),
        ('seed', ctypes.c_uint64),
        ('env_offset', ctypes.c_int64),
        ('domain_rand_length', ctypes.c_double),
        ('action_scale', ctypes.c_int32),
        ('action_clip', ctypes.c_int32),
    ]


class CtrHerRing(ctypes.Structure):
    """ctr_her_ring (include/ctr_b200.h): caller-owned, time-major, record-major replay ring."""
    _fields_ = [('n_envs', ctypes.c_int64), ('capacity', ctypes.c_int64), ('obs_dim', ctypes.c_int32), ('pad_', ctypes.c_int32)] + [
        (k, ctypes.c_void_p) for k in ('rec', 'done', 't_in_ep', 'ep_len', 'cur_obs', 'cur_achieved', 'cur_desired', 'cur_t')]


HER_RECORD = 44   # CTR_HER_RECORD: doubles per replay record


class CtrHerBatch(ctypes.Structure):
    """ctr_her_batch (include/ctr_b200.h): one sampled batch, SoA [rows][m]."""
    _fields_ = [('m', ctypes.c_int64)] + [
        (k, ctypes.c_void_p) for k in ('obs', 'next_obs', 'achieved', 'next_achieved', 'desired', 'action', 'reward', 'done',
                                       'src', 'goal_src')]


_lib = None


def load():
    """Load the CUDA library; raises if it has not been built (python gym-ctr-reach_b200/build.py)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("libctr_b200.so is missing (%s); build it with `python gym-ctr-reach_b200/build.py` -- "
                           "this package has no CPU fallback" % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    vp, i32, i64, f64 = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64, ctypes.c_double
    lib.ctr_last_error.restype = ctypes.c_char_p
    lib.ctr_last_error.argtypes = []
    lib.ctr_abi_version.restype = ctypes.c_int
    lib.ctr_create.argtypes = [ctypes.POINTER(CtrConfig), i64, i32, ctypes.POINTER(vp)]
    lib.ctr_destroy.argtypes = [vp]
    lib.ctr_num_envs.restype = i64
    lib.ctr_num_envs.argtypes = [vp]
    lib.ctr_obs_dim.restype = i32
    lib.ctr_obs_dim.argtypes = [vp]
    lib.ctr_buffer.argtypes = [vp, i32, ctypes.POINTER(vp), ctypes.POINTER(i64)]
    lib.ctr_system_constants.argtypes = [vp, i32, vp]
    lib.ctr_set_tolerance.argtypes = [vp, f64]
    lib.ctr_get_tolerance.restype = f64
    lib.ctr_get_tolerance.argtypes = [vp]
    lib.ctr_seed.argtypes = [vp, ctypes.c_uint64]
    lib.ctr_set_tube_params.argtypes = [vp, vp, vp]
    lib.ctr_fk.argtypes = [vp, vp, vp, vp, i64, i32, i32, vp, vp, vp, vp]
    lib.ctr_fk_host.argtypes = [vp, vp, vp, i64, i32, i32, vp, vp, vp]
    lib.ctr_set_action.argtypes = [vp, vp, vp, vp, i64, i32, vp]
    lib.ctr_set_state_host.argtypes = [vp, vp, vp, vp, vp]
    lib.ctr_reset.argtypes = [vp, vp, vp, vp, vp]
    lib.ctr_step.argtypes = [vp, vp, vp]
    lib.ctr_step_strided.argtypes = [vp, vp, i64, i64, vp]
    lib.ctr_step_host.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.ctr_step_host_packed.argtypes = [vp, vp, vp, vp]
    lib.ctr_step_host_packed_f32.argtypes = [vp, vp, vp, vp]
    lib.ctr_step_host_delta.argtypes = [vp, vp, vp, vp]
    lib.ctr_reset_host.argtypes = [vp, vp, vp, vp]
    lib.ctr_compute_reward.argtypes = [vp, vp, vp, i64, f64, i32, vp, vp]
    lib.ctr_relabel.argtypes = [vp, vp, vp, i64, f64, i32, vp, vp, vp]
    lib.ctr_stats.argtypes = [vp, vp, vp]
    lib.ctr_stats_reset.argtypes = [vp, vp]
    lib.ctr_work_counters.argtypes = [vp, vp, vp]
    lib.ctr_fault_counters.argtypes = [vp, vp, vp]
    lib.ctr_segment_tables.argtypes = [vp, vp, vp, vp, i64, vp, vp, vp, vp, vp]
    lib.ctr_checkpoint_bytes.restype = i64
    lib.ctr_checkpoint_bytes.argtypes = [vp]
    lib.ctr_checkpoint_save.argtypes = [vp, vp, i64]
    lib.ctr_checkpoint_load.argtypes = [vp, vp, i64]
    lib.ctr_debug_buffer.argtypes = [vp, i32, ctypes.POINTER(vp), ctypes.POINTER(i64)]
    lib.ctr_measure_fma_peak.argtypes = [i32, i32, ctypes.POINTER(f64)]
    lib.ctr_launch_count.restype = i64
    lib.ctr_launch_count.argtypes = [vp]
    lib.ctr_fk_shape.argtypes = [vp, vp, vp, vp, i64, i32, vp, vp, vp, vp, vp]
    lib.ctr_jacobian.argtypes = [vp, vp, vp, i64, f64, vp, vp, vp]
    lib.ctr_jacobian_noalloc.argtypes = [vp, vp, vp, i64, f64, vp, vp, vp]
    lib.ctr_ik_dls.argtypes = [vp, vp, vp, vp, i64, f64, f64, f64, i32, vp, vp, vp]
    lib.ctr_bvp_pass.argtypes = [vp, vp, vp, vp, vp, vp, i64, f64, vp, vp, vp, vp, vp]
    lib.ctr_bvp_jacobian.argtypes = [vp, vp, vp, vp, vp, vp, i64, f64, vp, vp, vp, vp]
    lib.ctr_bvp_solve.argtypes = [vp, vp, vp, vp, vp, i64, f64, f64, i32, vp, vp, vp, vp, vp, vp]
    lib.ctr_bvp_ik_dls.argtypes = [vp, vp, vp, vp, i64, f64, f64, i32, f64, vp, vp, vp]
    lib.ctr_her_begin.argtypes = [vp, ctypes.POINTER(CtrHerRing), vp]
    lib.ctr_her_store.argtypes = [vp, ctypes.POINTER(CtrHerRing), vp, i64, i64, i64, vp]
    lib.ctr_her_sample.argtypes = [vp, ctypes.POINTER(CtrHerRing), i64, i64, i32, i32, f64, ctypes.c_uint64, ctypes.c_uint64,
                                   ctypes.POINTER(CtrHerBatch), vp]
    for name in EXPORTS:
        if name not in ('ctr_last_error', 'ctr_num_envs', 'ctr_obs_dim', 'ctr_get_tolerance', 'ctr_launch_count', 'ctr_checkpoint_bytes'):
            getattr(lib, name).restype = ctypes.c_int
    if lib.ctr_abi_version() != ABI_VERSION:
        raise RuntimeError("libctr_b200.so ABI version mismatch")
    _lib = lib
    return lib


class CtrError(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        raise CtrError(load().ctr_last_error().decode())


def make_config(systems, n_substeps=10, max_steps_per_episode=150, joint_representation='egocentric',
                constrain_alpha=False, extension_action_limit=0.001, rotation_action_limit=5, resample_joints=True,
                auto_reset=True, integrator='dp45', precision='f64', rk4_steps=8, reward_mode='sparse', rtol=1e-3,
                atol=1e-6, initial_tol=0.02, domain_rand=0.0, seed=0, env_offset=0, domain_rand_length=0.0, action_scale=False,
                action_clip=False):
    """systems: float64 [n_systems][3][8] in Tube argument order (Tube.py:7-8)."""
    systems = np.ascontiguousarray(systems, dtype=np.float64).reshape(-1, 3, 8)
    if not 1 <= systems.shape[0] <= MAX_SYSTEMS:
        raise ValueError("between 1 and %d tube systems are supported" % MAX_SYSTEMS)
    cfg = CtrConfig()
    cfg.abi_version = ABI_VERSION
    cfg.n_systems = systems.shape[0]
    flat = np.zeros(MAX_SYSTEMS * 24)
    flat[:systems.size] = systems.ravel()
    cfg.tube = (ctypes.c_double * (MAX_SYSTEMS * 24))(*flat)
    cfg.n_substeps = int(n_substeps)
    cfg.max_steps_per_episode = int(max_steps_per_episode)
    assert joint_representation in REP, "Incorrect joint representation selected."
    cfg.joint_representation = REP[joint_representation]
    cfg.constrain_alpha = int(bool(constrain_alpha))
    cfg.extension_action_limit = float(extension_action_limit)
    cfg.rotation_action_limit_deg = float(rotation_action_limit)
    cfg.resample_joints = int(bool(resample_joints))
    cfg.auto_reset = int(bool(auto_reset))
    cfg.integrator = INTEGRATOR[integrator]
    cfg.precision = PRECISION[precision]
    cfg.rk4_steps = int(rk4_steps)
    cfg.reward_mode = REWARD[reward_mode]
    cfg.rtol, cfg.atol = float(rtol), float(atol)
    cfg.initial_tol = float(initial_tol)
    cfg.domain_rand = float(domain_rand)
    cfg.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
    cfg.env_offset = int(env_offset)
    cfg.domain_rand_length = float(domain_rand_length)
    cfg.action_scale = int(bool(action_scale))
    cfg.action_clip = int(bool(action_clip))
    return cfg


def ptr(x):
    """Raw pointer of a torch tensor / numpy array / None."""
    if x is None:
        return None
    if isinstance(x, np.ndarray):
        return x.ctypes.data
    return x.data_ptr()
