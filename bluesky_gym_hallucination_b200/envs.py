"""Single-env facades registered under the reference's ids: the unchanged Gymnasium API
(gym.make('HorizontalCREnv-v0' | 'VerticalCREnv-v0' | 'SectorCREnv-v0' | 'MergeEnv-v0'), reset/step, obs dicts of
float64 arrays, info keys) in front of the batched CUDA core with num_envs = 1.
Reference classes: bluesky_gym/envs/horizontal_cr_env.py:32-125, vertical_cr_env.py:47-305,
sector_cr_env.py:39-132, merge_env.py:47-146.  render_mode="human" needs an SDL window: unavailable on this box."""
import numpy as np
import torch
from . import _lib as L
from .gym_compat import Env
from .vec_env import BlueSkyVecEnv, INFO_KEYS


class _SingleEnv(Env):
    metadata = {"render_modes": ["rgb_array", "human"], "render_fps": 120}
    ENV_ID = None

    def __init__(self, render_mode=None, precision='fp64', device=0, asas=False, mvp=False, **kw):
        assert render_mode is None or render_mode in self.metadata["render_modes"]
        if render_mode == "human":
            raise NotImplementedError("render_mode='human' opens a pygame/SDL window: unavailable on this box")
        self.render_mode = render_mode
        self._v = BlueSkyVecEnv(self.ENV_ID, 1, device=device, precision=precision, asas=asas, mvp=mvp,
                                max_episode_steps=0, **kw)        # TimeLimit is applied by the registry wrapper
        self.observation_space, self.action_space = self._v.single_observation_space, self._v.single_action_space
        self._episode = 0

    def reset(self, seed=None, options=None):
        super().reset(seed=seed)
        # like the reference, scenario draws do not come from self.np_random; an explicit seed pins them
        s = int(seed) if seed is not None else int(np.random.randint(0, 2 ** 31 - 1))
        self._v.core.reset(seed=s)
        self._v.core.set_autoreset(False)
        obs = self._v._obs_dict(self._v.core.observe().cpu().numpy())
        keys = INFO_KEYS[self._v.kind]
        third = 0 if self._v.kind == L.KIND_VERTICAL else float('nan')
        info = {keys[0]: 0, keys[1]: 0, keys[2]: third}
        if len(keys) > 3:
            info[keys[3]] = 0
        return {k: v[0] for k, v in obs.items()}, info

    def step(self, action):
        obs, rew, term, trunc, info = self._v.step(np.asarray(action, dtype=np.float64).reshape(1, -1))
        o = {k: v[0] for k, v in obs.items()}
        i = {k: (int(v[0]) if v.dtype.kind == 'i' else float(v[0])) for k, v in info.items()}
        return o, float(rew[0]), int(term[0]), bool(trunc[0]), i

    def render(self):
        return None

    def close(self):
        self._v.close()


class HorizontalCREnv(_SingleEnv):
    ENV_ID = 'HorizontalCREnv-v0'


class VerticalCREnv(_SingleEnv):
    ENV_ID = 'VerticalCREnv-v0'


class SectorCREnv(_SingleEnv):
    ENV_ID = 'SectorCREnv-v0'

    def __init__(self, render_mode=None, ac_density_mode="normal", **kw):
        self.density_mode = ac_density_mode                              # sector_cr_env.py:47
        super().__init__(render_mode=render_mode, ac_density_mode=ac_density_mode, **kw)


class MergeEnv(_SingleEnv):
    """bluesky_gym/envs/merge_env.py:47-146 (never truncates itself: TimeLimit 50 comes from the registry)."""
    ENV_ID = 'MergeEnv-v0'
