"""Gymnasium when it is installed, otherwise the few names of it this package needs (spaces.Box/Dict,
Env, register/make with the TimeLimit the reference's registrations ask for).  gymnasium==1.1.1 is a
dependency of the reference (pyproject.toml:11) but is absent from this image and cannot be installed
offline, so the registry falls back to this module; the public API is the same either way."""
import importlib
import numpy as np

try:  # pragma: no cover - depends on the environment
    import gymnasium as _gym
    from gymnasium import spaces
    from gymnasium.envs.registration import register as _register, registry as _registry
    HAVE_GYMNASIUM = True
    Env = _gym.Env

    def register(id, entry_point, max_episode_steps=None, **kw):
        if id not in _registry:
            _register(id=id, entry_point=entry_point, max_episode_steps=max_episode_steps, **kw)

    make = _gym.make
except ImportError:
    HAVE_GYMNASIUM = False

    class _Box:
        def __init__(self, low, high, shape=None, dtype=np.float32):
            self.shape = tuple(shape) if shape is not None else (1,)
            self.dtype = np.dtype(dtype)
            self.low = np.full(self.shape, low, dtype=self.dtype)
            self.high = np.full(self.shape, high, dtype=self.dtype)
            self._rng = np.random.default_rng()

        def seed(self, seed=None):
            self._rng = np.random.default_rng(seed)

        def sample(self):
            lo = np.where(np.isfinite(self.low), self.low, -1.0)
            hi = np.where(np.isfinite(self.high), self.high, 1.0)
            return self._rng.uniform(lo, hi).astype(self.dtype)

        def contains(self, x):
            x = np.asarray(x)
            return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))

        def __repr__(self):
            return f"Box({self.low.min()}, {self.high.max()}, {self.shape}, {self.dtype})"

    class _Dict:
        def __init__(self, spaces):
            self.spaces = dict(spaces)

        def __getitem__(self, k):
            return self.spaces[k]

        def keys(self):
            return self.spaces.keys()

        def items(self):
            return self.spaces.items()

        def sample(self):
            return {k: s.sample() for k, s in self.spaces.items()}

        def contains(self, x):
            return set(x.keys()) == set(self.spaces.keys()) and all(s.contains(x[k]) for k, s in self.spaces.items())

    class spaces:  # noqa: N801 - mirrors the gymnasium module name
        Box = _Box
        Dict = _Dict

    class Env:
        metadata = {}
        render_mode = None
        np_random = None

        def reset(self, *, seed=None, options=None):
            if seed is not None or self.np_random is None:
                self.np_random = np.random.default_rng(seed)

        def close(self):
            pass

        @property
        def unwrapped(self):
            return self

    _REGISTRY = {}

    def register(id, entry_point, max_episode_steps=None, **kw):
        _REGISTRY[id] = dict(entry_point=entry_point, max_episode_steps=max_episode_steps, kwargs=kw)

    class TimeLimit:
        """gymnasium.wrappers.TimeLimit: truncated=True once max_episode_steps is reached."""

        def __init__(self, env, max_episode_steps):
            self.env, self._max, self._n = env, max_episode_steps, 0

        def __getattr__(self, name):
            return getattr(self.env, name)

        @property
        def unwrapped(self):
            return self.env

        def reset(self, **kw):
            self._n = 0
            return self.env.reset(**kw)

        def step(self, action):
            obs, rew, term, trunc, info = self.env.step(action)
            self._n += 1
            if self._max is not None and self._n >= self._max:
                trunc = True
            return obs, rew, term, trunc, info

    def make(id, **kwargs):
        spec = _REGISTRY[id]
        mod, cls = spec['entry_point'].split(':')
        env = getattr(importlib.import_module(mod), cls)(**{**spec['kwargs'], **kwargs})
        return TimeLimit(env, spec['max_episode_steps']) if spec['max_episode_steps'] else env
