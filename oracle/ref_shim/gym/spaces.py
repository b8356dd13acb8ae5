"""gym.spaces stand-in: Box (bounds cast to dtype, default float32; uniform sampling in float64 then cast,
as gym 0.1x does for bounded boxes) and Dict."""
import numpy as np

_rng = np.random.RandomState(0)


def seed(s):
    global _rng
    _rng = np.random.RandomState(s)


class Box(object):
    def __init__(self, low, high, shape=None, dtype=np.float32):
        self.dtype = np.dtype(dtype)
        self.low = np.asarray(low).astype(self.dtype)
        self.high = np.asarray(high).astype(self.dtype)
        self.shape = self.low.shape

    def sample(self):
        return _rng.uniform(low=self.low, high=self.high, size=self.shape).astype(self.dtype)

    def contains(self, x):
        x = np.asarray(x)
        return bool(x.shape == self.shape and np.all(x >= self.low) and np.all(x <= self.high))


class Dict(object):
    def __init__(self, spaces=None, **kw):
        self.spaces = dict(spaces or {}, **kw)

    def __getitem__(self, k):
        return self.spaces[k]

    def contains(self, x):
        return all(self.spaces[k].contains(v) for k, v in x.items())
