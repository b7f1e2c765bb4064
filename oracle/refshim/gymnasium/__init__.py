"""ORACLE shim: the sliver of gymnasium==1.1.1 the reference env modules import
(``gym.Env``, ``spaces.Box/Dict``, ``envs.registration.register``).  Used only to
let the unmodified reference classes import in a container without gymnasium."""
import numpy as np
from . import spaces
from .envs import registration
from .envs.registration import register, make, registry


class Env:
    metadata = {}
    render_mode = None
    np_random = None

    def reset(self, *, seed=None, options=None):
        if seed is not None or self.np_random is None:
            self.np_random = np.random.default_rng(seed)

    def close(self):
        pass


class Wrapper(Env):
    def __init__(self, env):
        self.env = env

    def __getattr__(self, name):
        return getattr(self.env, name)
