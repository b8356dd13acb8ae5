"""ctr_reach_envs -- B200-native drop-in for the env-step path of keshaviyengar/gym-ctr-reach.

Same import path, gym id and default kwargs as the reference (ctr_reach_envs/__init__.py:4-96); the step / reset /
compute_reward path runs in hand-written sm_100a kernels behind libctr_b200.so (include/ctr_b200.h)."""
import numpy as np

from .registration import register, make, spec, registry  # noqa: F401

TUBE_FIELDS = ('length', 'length_curved', 'diameter_inner', 'diameter_outer', 'stiffness', 'torsional_stiffness',
               'x_curvature', 'y_curvature')

# The four 3-tube systems the reference registers (__init__.py:9-70), one row per tube in TUBE_FIELDS order
# [m, m, m, m, Pa, Pa, 1/m, 1/m]: ctr_0 Khadem (autonomous steering), ctr_1 Grassmann (6-DOF FK/IK learning),
# ctr_2 RViM lab, ctr_3 origin unknown.
_TUBE_TABLE = {
    'ctr_0': [(431e-3, 103e-3, 0.7e-3, 1.10e-3, 10.25e+10, 18.79e+10, 21.3, 0),
              (332e-3, 113e-3, 1.4e-3, 1.8e-3, 68.6e+10, 11.53e+10, 13.1, 0),
              (174e-3, 134e-3, 2e-3, 2.4e-3, 16.96e+10, 14.25e+10, 3.5, 0)],
    'ctr_1': [(370e-3, 45e-3, 0.3e-3, 0.4e-3, 50e+10, 2.3e+10, 15.8, 0),
              (305e-3, 100e-3, 0.7e-3, 0.9e-3, 50e+10, 2.3e+10, 9.27, 0),
              (170e-3, 100e-3, 1.2e-3, 1.5e-3, 50e+10, 2.3e+10, 4.37, 0)],
    'ctr_2': [(309e-3, 145e-3, 0.7e-3, 1.1e-3, 75e+9, 25e+9, 13.52, 0),
              (275e-3, 114e-3, 1.4e-3, 1.8e-3, 75e+9, 25e+9, 11.68, 0),
              (173e-3, 173e-3, 1.83e-3, 2.39e-3, 75e+9, 25e+9, 10.8, 0)],
    'ctr_3': [(150e-3, 100e-3, 1.0e-3, 2.4e-3, 5e+10, 2.3e+10, 15.82, 0),
              (100e-3, 21.6e-3, 3.0e-3, 3.8e-3, 5e+10, 2.3e+10, 11.8, 0),
              (70e-3, 8.8e-3, 4.4e-3, 5.4e-3, 5.0e+10, 2.3e+10, 20.04, 0)],
}


def default_kwargs():
    """A fresh copy of the kwargs `CTR-Reach-v0` is registered with (__init__.py:6-94)."""
    systems = {name: {'tube_%d' % i: dict(zip(TUBE_FIELDS, row)) for i, row in enumerate(rows)}
               for name, rows in _TUBE_TABLE.items()}
    return dict(
        ctr_systems_parameters=systems,
        extension_action_limit=0.001, rotation_action_limit=5,           # m / degrees per substep
        max_steps_per_episode=150, n_substeps=10,
        goal_tolerance_parameters=dict(inc_tol_obs=False, final_tol=0.001, initial_tol=0.020, N_ts=200000,
                                       function='constant', set_tol=0),
        noise_parameters=dict(rotation_std=np.deg2rad(0), extension_std=0.001 * np.deg2rad(0), tracking_std=0.0),
        select_systems=[0], constrain_alpha=False,
        initial_joints=np.array([0, 0, 0, 0, 0, 0]),                     # [beta_0..2, alpha_0..2]
        joint_representation='egocentric', resample_joints=True, evaluation=False, length_based_sample=False,
        domain_rand=0.0)


register(id='CTR-Reach-v0', entry_point='ctr_reach_envs.envs:CtrReachEnv', kwargs=default_kwargs(), max_episode_steps=150)
