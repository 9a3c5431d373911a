"""gym.spaces.Box / Discrete: shape, dtype, n, sample(), contains().

The reference saves its policy with the spaces inside the checkpoint and reads it back with a plain `th.load(path)`
(actor_critic_ts.py:227,263); torch >= 2.6 unpickles with an allow-list by default, so the two classes pickle as plain
Python values and are registered as safe globals here."""
import numpy as np


class Space:
    def __init__(self, shape=None, dtype=None):
        self.shape = None if shape is None else tuple(shape)
        self.dtype = None if dtype is None else np.dtype(dtype)

    def __contains__(self, x):
        return self.contains(x)


class Box(Space):
    def __init__(self, low, high, shape=None, dtype=np.float32):
        if shape is None:
            shape = np.asarray(low).shape
        super().__init__(shape, dtype)
        self.low = np.full(self.shape, low, dtype=self.dtype) if np.isscalar(low) else np.asarray(low, self.dtype)
        self.high = np.full(self.shape, high, dtype=self.dtype) if np.isscalar(high) else np.asarray(high, self.dtype)

    def sample(self):
        return np.random.uniform(self.low, self.high, self.shape).astype(self.dtype)

    def contains(self, x):
        x = np.asarray(x)
        return x.shape == self.shape and bool(np.all(x >= self.low)) and bool(np.all(x <= self.high))

    def __repr__(self):
        return f"Box({self.low.min()}, {self.high.max()}, {self.shape}, {self.dtype})"

    def __getstate__(self):
        uniform = bool((self.low == self.low.flat[0]).all() and (self.high == self.high.flat[0]).all())
        return {"low": float(self.low.flat[0]) if uniform else self.low.tolist(),
                "high": float(self.high.flat[0]) if uniform else self.high.tolist(), "shape": list(self.shape), "dtype": self.dtype.name}

    def __setstate__(self, st):
        self.__init__(st["low"], st["high"], tuple(st["shape"]), np.dtype(st["dtype"]))

    def __eq__(self, other):
        return isinstance(other, Box) and self.shape == other.shape and np.allclose(self.low, other.low) and np.allclose(self.high, other.high)


class Discrete(Space):
    def __init__(self, n):
        super().__init__((), np.int64)
        self.n = int(n)

    def sample(self):
        return int(np.random.randint(self.n))

    def contains(self, x):
        return 0 <= int(x) < self.n

    def __repr__(self):
        return f"Discrete({self.n})"

    def __getstate__(self):
        return {"n": self.n}

    def __setstate__(self, st):
        self.__init__(st["n"])

    def __eq__(self, other):
        return isinstance(other, Discrete) and self.n == other.n


try:                                             # torch >= 2.6: allow-listed unpickling (see the module docstring)
    import torch.serialization as _ts
    _ts.add_safe_globals([Box, Discrete, Space])
except (ImportError, AttributeError):
    pass
