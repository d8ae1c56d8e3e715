"""gymnasium shims.  gymnasium 0.29.1 is the reference's pinned dependency (requirements.txt:1)
but is not installed on this image; when it is importable the real classes are used, otherwise
these minimal stand-ins provide the metadata surface (spaces + VectorEnv base) callers touch.
No arithmetic lives here.
"""
from __future__ import annotations

import numpy as np

try:  # pragma: no cover - depends on the image
    import gymnasium as _gym
    from gymnasium import spaces as _spaces
    from gymnasium.vector import VectorEnv as _VectorEnv

    HAVE_GYMNASIUM = True
    Box = _spaces.Box
    Env = _gym.Env
    VectorEnv = _VectorEnv

    def batch_box(box, n):
        from gymnasium.vector.utils import batch_space
        return batch_space(box, n)

except Exception:  # gymnasium absent
    HAVE_GYMNASIUM = False

    class Box:  # the slice of gymnasium.spaces.Box the reference and SB3-style callers read
        def __init__(self, low, high, shape=None, dtype=np.float32):
            low = np.asarray(low, dtype=dtype)
            high = np.asarray(high, dtype=dtype)
            if shape is None:
                shape = low.shape if low.shape else high.shape
            self.low = np.broadcast_to(low, shape).astype(dtype).copy()
            self.high = np.broadcast_to(high, shape).astype(dtype).copy()
            self.shape = tuple(shape)
            self.dtype = np.dtype(dtype)
            self._rng = np.random.default_rng()

        def seed(self, seed=None):
            self._rng = np.random.default_rng(seed)

        def sample(self):
            return self._rng.uniform(self.low, self.high).astype(self.dtype)

        def contains(self, x):
            x = np.asarray(x)
            return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))

        def __repr__(self):
            return f"Box({self.low.min()}, {self.high.max()}, {self.shape}, {self.dtype})"

    class Env:
        metadata: dict = {}
        render_mode = None

        def reset(self, *, seed=None, options=None):
            return None

        def close(self):
            pass

    class VectorEnv:
        """Gymnasium-0.29-shaped base: num_envs, (single_)observation/action_space, step_async/wait."""
        metadata: dict = {}
        closed = False

        def reset(self, *, seed=None, options=None):
            raise NotImplementedError

        def step_async(self, actions):
            self._pending_actions = actions

        def step_wait(self):
            return self.step(self._pending_actions)

        def close(self, **kwargs):
            if not self.closed:
                self.close_extras(**kwargs)
                self.closed = True

        def close_extras(self, **kwargs):
            pass

    def batch_box(box, n):
        return Box(np.broadcast_to(box.low, (n,) + box.shape), np.broadcast_to(box.high, (n,) + box.shape),
                   shape=(n,) + box.shape, dtype=box.dtype)
