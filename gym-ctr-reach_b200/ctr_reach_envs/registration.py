"""gym-free registry with the reference's id (`CTR-Reach-v0`, __init__.py:4-96).  When `gym` is importable the id is
registered there as well, so `gym.make('CTR-Reach-v0')` keeps working."""
import importlib

registry = {}


class EnvSpec(object):
    def __init__(self, id, entry_point, kwargs=None, max_episode_steps=None):
        self.id, self.entry_point, self.kwargs = id, entry_point, dict(kwargs or {})
        self.max_episode_steps = max_episode_steps

    def make(self, **kwargs):
        mod, name = self.entry_point.split(':')
        cls = getattr(importlib.import_module(mod), name)
        kw = dict(self.kwargs)
        kw.update(kwargs)
        env = cls(**kw)
        env.spec = self
        return env


def register(id, entry_point, kwargs=None, max_episode_steps=None):
    registry[id] = EnvSpec(id, entry_point, kwargs, max_episode_steps)
    try:  # optional: mirror into a real gym installation
        from gym.envs.registration import register as gym_register
        gym_register(id=id, entry_point=entry_point, kwargs=kwargs, max_episode_steps=max_episode_steps)
    except Exception:
        pass


def spec(id):
    return registry[id]


def make(id, **kwargs):
    return registry[id].make(**kwargs)
