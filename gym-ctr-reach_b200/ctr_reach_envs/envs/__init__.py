from .goal_tolerance import GoalTolerance  # noqa: F401
from .vec_env import CtrReachVecEnv  # noqa: F401
from .ctr_reach_env import CtrReachEnv  # noqa: F401
