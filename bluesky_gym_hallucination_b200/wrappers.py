"""The reference's wrappers, with their constructor signatures, in front of the CUDA core:

  WindFieldWrapper(env, lat, lon, vnorth, veast, alt=None, augment_obs=False)   bluesky_gym/wrappers/wind.py:8-64
  NoisyObservationWrapper(env, noise_level=0.1)                                bluesky_gym/wrappers/uncertainty.py:4-31

`env` is one of this package's single-env facades (envs.py, also when wrapped by the registry's TimeLimit) or a
BlueSkyVecEnv.  Nothing is computed on the host: the wrappers switch the feature on in the core (bsg_set_wind /
bsg_set_obs_noise), which runs it inside the fused step kernel -- wind in the kinematics and the autopilot, wind_u /
wind_v appended to the observation, N(0, sigma) added to every observation slot from a Philox stream.
An altitude profile (`alt=`) makes the point's vectors a function of altitude (bsg_set_wind_profile).
"""
import numpy as np
from .core import obs_slices
from .vec_env import BlueSkyVecEnv, single_spaces


def _venv(env):
    e = env
    for _ in range(8):                       # unwrap TimeLimit / other wrappers down to the facade
        if isinstance(e, BlueSkyVecEnv):
            return e
        if hasattr(e, '_v'):
            return e._v
        e = getattr(e, 'env', None)
        if e is None:
            break
    raise TypeError("the wrapped env must be a bluesky_gym_hallucination_b200 env (single facade or BlueSkyVecEnv)")


class _Wrapper:
    def __init__(self, env):
        self.env = env
        self._v = _venv(env)

    def __getattr__(self, name):
        return getattr(self.env, name)

    def reset(self, **kwargs):
        return self.env.reset(**kwargs)

    def step(self, action):
        return self.env.step(action)

    def close(self):
        return self.env.close()


class WindFieldWrapper(_Wrapper):
    def __init__(self, env, lat, lon, vnorth, veast, alt=None, augment_obs=False):
        super().__init__(env)
        v = self._v
        v.core.set_wind(lat, lon, vnorth, veast, alt=alt, augment_obs=augment_obs)     # alt: one point with a profile
        v._refresh_spaces()
        self.observation_space = v.single_observation_space
        if hasattr(env, 'observation_space') and env is not v:
            try:
                env.observation_space = v.single_observation_space
            except AttributeError:
                pass


class NoisyObservationWrapper(_Wrapper):
    def __init__(self, env, noise_level=0.1, seed=None):
        super().__init__(env)
        self.noise_level = noise_level
        s = int(np.random.randint(0, 2 ** 31 - 1)) if seed is None else int(seed)
        self._v.core.set_obs_noise(noise_level, seed=s)
