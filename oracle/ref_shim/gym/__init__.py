"""Minimal stand-in for the `gym` package (NOT installed in this image).

TEST INFRASTRUCTURE ONLY.  It exists so that the *unmodified* reference sources under
/root/reference can be imported in the build container to generate golden vectors
(oracle/make_golden.py) and to validate the oracle restatement.  It models only what the
reference touches: `gym.GoalEnv`, `gym.spaces.Box/Dict`,
`gym.envs.registration.register`, `gym.make`, `gym.spec`.
"""
from . import spaces  # noqa: F401
from .envs import registration as _registration
from .envs.registration import make, spec, register  # noqa: F401


class Env(object):
    metadata = {'render.modes': []}
    reward_range = (-float('inf'), float('inf'))
    spec = None
    action_space = None
    observation_space = None


class GoalEnv(Env):
    pass
