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


def register(id, entry_point, kwargs=None, max_episode_steps=None, **_):
    registry[id] = EnvSpec(id, entry_point, kwargs, max_episode_steps)


def spec(id):
    return registry[id]


def make(id, **kwargs):
    return registry[id].make(**kwargs)
