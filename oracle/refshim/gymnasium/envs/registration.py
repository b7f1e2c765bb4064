"""ORACLE shim: gymnasium.envs.registration (id -> entry point + TimeLimit)."""
import importlib

registry = {}


def register(id, entry_point, max_episode_steps=None, **kwargs):
    registry[id] = dict(entry_point=entry_point, max_episode_steps=max_episode_steps, kwargs=kwargs)


class _TimeLimit:
    def __init__(self, env, max_episode_steps):
        self.env = env
        self.unwrapped = env
        self._max = max_episode_steps
        self._n = 0

    def __getattr__(self, name):
        return getattr(self.env, name)

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
    spec = registry[id]
    mod, cls = spec['entry_point'].split(':')
    env = getattr(importlib.import_module(mod), cls)(**{**spec['kwargs'], **kwargs})
    return _TimeLimit(env, spec['max_episode_steps'])
