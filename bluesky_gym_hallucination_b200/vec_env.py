"""BlueSkyVecEnv -- thousands of independent BlueSky-Gym envs behind the Gymnasium VectorEnv API, one
fused CUDA launch per step (libbsgcore.so).  Host-side mirror of the reference's env interface:
same env ids, observation keys, action shapes, reward / terminated / truncated / info semantics as
  bluesky_gym/envs/horizontal_cr_env.py:49-62,82-125,215-223
  bluesky_gym/envs/vertical_cr_env.py:67-85,203-211,258-305
  bluesky_gym/envs/sector_cr_env.py:52-67,87-132,219-225
  bluesky_gym/envs/merge_env.py:57-76,100-146,238-245
and the registrations of bluesky_gym/__init__.py:18-46 (max_episode_steps 300 / 300 / 200 / 50).  The reference's
wrappers (bluesky_gym/wrappers/wind.py, uncertainty.py) are constructor options: `wind=`, `obs_noise=`.
PyTorch appears only as the tensor hand-off to RL libraries (`step_tensors`)."""
from collections.abc import Mapping

import numpy as np
import torch
from . import _lib as L
from .core import BsgCore, obs_slices
from .gym_compat import spaces

ENV_SPECS = {
    # id: (kind, reference aircraft count, max_episode_steps, action dim)
    'HorizontalCREnv-v0': (L.KIND_HORIZONTAL, 6, 300, 1),
    'VerticalCREnv-v0': (L.KIND_VERTICAL, 6, 300, 1),
    'SectorCREnv-v0': (L.KIND_SECTOR, 32, 200, 2),
    'MergeEnv-v0': (L.KIND_MERGE, 20, 50, 2),
}
INFO_KEYS = {L.KIND_HORIZONTAL: ('total_reward', 'total_intrusions', 'average_drift'),
             L.KIND_VERTICAL: ('total_reward', 'total_intrusions', 'final_altitude'),
             L.KIND_SECTOR: ('total_reward', 'total_intrusions', 'average_drift'),
             L.KIND_MERGE: ('total_reward', 'total_intrusions', 'average_drift', 'faf_reach')}   # merge_env.py:238-245


def single_spaces(kind, n_ac, wind_obs=False):
    """The reference's Dict observation space (float64 Boxes) and action Box for one env."""
    d = {}
    for key, a, b in obs_slices(kind, n_ac, wind_obs):
        bounded = (kind == L.KIND_SECTOR and key in ('cos(drift)', 'sin(drift)', 'airspeed')) or \
                  (kind == L.KIND_MERGE and key in ('cos(drift)', 'sin(drift)'))   # sector_cr_env.py:54-56, merge_env.py:60-61
        lo, hi = (-1, 1) if bounded else ((0, 1) if key == 'faf_reached' else (-np.inf, np.inf))
        d[key] = spaces.Box(lo, hi, shape=(b - a,), dtype=np.float64)
    act = spaces.Box(-1, 1, shape=(2 if kind in (L.KIND_SECTOR, L.KIND_MERGE) else 1,), dtype=np.float64)
    return spaces.Dict(d), act


# default ND-map grid of the env layer: 320 x 320 km around (52 N, 4 E) in 5 km cells, 12 km of altitude in 1.5 km layers,
# 30 look-ahead steps of 10 s
NDMAP_DEFAULT = dict(nx=64, ny=64, nz=8, nt=30, x0=-160e3, y0=-160e3, z0=0.0, cx=5e3, cy=5e3, cz=1.5e3, dt=10.0,
                     lat0=52.0, lon0=4.0)


def _batch_space(space, n):
    """gymnasium.vector.utils.batch_space for the two space types the envs use: Box -> Box [n, ...], Dict of them"""
    if hasattr(space, 'spaces'):
        return spaces.Dict({k: _batch_space(b, n) for k, b in space.spaces.items()})
    shape = (n,) + tuple(space.shape)
    return spaces.Box(np.broadcast_to(space.low, shape).copy(), np.broadcast_to(space.high, shape).copy(), shape=shape,
                      dtype=space.dtype.type)


class BlueSkyVecEnv:
    """Batched env.  reset()/step() speak numpy (Gymnasium VectorEnv semantics, next-step auto-reset);
    step_tensors() keeps everything on the GPU for RL libraries that take torch tensors."""
    metadata = {"render_modes": [], "autoreset_mode": "next_step"}

    def __init__(self, env_id, num_envs, num_aircraft=None, device=0, precision='fp32', asas=False, mvp=False,
                 seed=0, cd_outputs=False, max_episode_steps=None, wind=None, wind_obs=False, obs_noise=0.0,
                 ac_density_mode='normal', fixed_nac=False, ndmap_obs=None, obs_dtype=np.float64):
        """wind = (lat, lon, vnorth, veast) like WindFieldWrapper(env, lat, lon, vnorth, veast, augment_obs=wind_obs);
        obs_noise = NoisyObservationWrapper(env, noise_level); ac_density_mode = SectorCREnv's "normal" | "uniform"
        (sector_cr_env.py:45,98-103); fixed_nac = every aircraft slot live (the synthetic scale-up configs).
        mvp = True | 'horizontal' (BlueSky's default resolution method, heading + speed) | 'vertical' (RMETHV V/S) |
        'both' (horizontal + vertical): ASAS ON + RESO MVP for the intruders, RESOOFF for the agent.
        obs_dtype = np.float64 (the reference's declared dtype: every step converts the batch) | np.float32 with the fp32
        core: the observation dict holds VIEWS of the pinned host block the step filled -- no conversion, no copy; valid
        until the second next step.
        ndmap_obs = True | dict of BsgCore.ndmap keywords: every reset / step also builds the ND-map (4D conflict grid over
        the look-ahead trajectories) of the live traffic and appends the observation key 'ndmap_conflict_cells' [nt]: the
        share of grid cells in conflict at each look-ahead step (the map itself stays on the device: ndmap_last)."""
        reso_mode = {True: 0, 'horizontal': 0, 'vertical': 1, 'both': 2}[mvp] if mvp else 0
        if ac_density_mode not in ('normal', 'uniform'):
            raise ValueError(f"ac_density_mode must be 'normal' or 'uniform', not {ac_density_mode!r}")
        if env_id not in ENV_SPECS:
            raise ValueError(f"unknown env id {env_id!r}; the batched core covers {sorted(ENV_SPECS)}")
        kind, n_ref, max_steps, act_dim = ENV_SPECS[env_id]
        self.env_id, self.kind, self.num_envs = env_id, kind, int(num_envs)
        self.num_aircraft = int(num_aircraft or n_ref)
        self.core = BsgCore(kind, self.num_envs, self.num_aircraft, precision=precision, device=device,
                            asas_on=bool(asas or mvp), reso_on=bool(mvp), reso_off_own=bool(mvp), reso_mode=reso_mode,
                            max_episode_steps=max_steps if max_episode_steps is None else max_episode_steps,
                            cd_outputs=cd_outputs, density_mode=ac_density_mode, fixed_nac=fixed_nac)
        if wind is not None:
            self.core.set_wind(*wind, augment_obs=wind_obs)
        if obs_noise:
            self.core.set_obs_noise(obs_noise, seed=int(seed) + 0x5EED)
        self._seed = int(seed)
        self._resets = 0
        self._nd = None
        self._obs_dtype = np.dtype(obs_dtype)
        if self._obs_dtype not in (np.dtype(np.float64), np.dtype(np.float32)):
            raise ValueError("obs_dtype must be float64 or float32")
        self._refresh_spaces()
        if ndmap_obs:
            nd = dict(NDMAP_DEFAULT)                   # the default grid, overridden key by key
            if isinstance(ndmap_obs, dict):
                nd.update(ndmap_obs)
            self._nd, self._nd_out, self.ndmap_last = nd, None, None
            self._nd_cells = float(nd['nx'] * nd['ny'] * nd['nz'])
            self._refresh_spaces()
        self._act = torch.zeros((self.num_envs, act_dim), dtype=self.core.rdtype, device=self.core.device)
        self.closed = False

    def _refresh_spaces(self):
        """observation / action spaces and the slices of the flat observation (again after a wrapper added wind slots)"""
        wind_obs = self.core.wind_obs
        self.single_observation_space, self.single_action_space = single_spaces(self.kind, self.num_aircraft, wind_obs)
        d = dict(self.single_observation_space.spaces)
        if self._obs_dtype != np.dtype(np.float64):     # same keys and shapes, the dtype the caller asked for
            d = {k: spaces.Box(b.low, b.high, shape=b.shape, dtype=self._obs_dtype.type) for k, b in d.items()}
        if self._nd is not None:
            d['ndmap_conflict_cells'] = spaces.Box(0.0, 1.0, shape=(self._nd['nt'],), dtype=self._obs_dtype.type)
        self.single_observation_space = spaces.Dict(d)
        # Gymnasium VectorEnv: observation_space / action_space are the BATCHED spaces (leading axis num_envs)
        self.observation_space = _batch_space(self.single_observation_space, self.num_envs)
        self.action_space = _batch_space(self.single_action_space, self.num_envs)
        self._slices = obs_slices(self.kind, self.num_aircraft, wind_obs)

    # ------------------------------------------------------------------ Gymnasium VectorEnv API
    def _slice_obs(self, flat):
        """the keys of the flat observation rows (any subset of envs); no ND-map key"""
        flat = np.asarray(flat, dtype=self._obs_dtype)          # float32 from the fp32 core: a view, not a copy
        return {k: flat[:, a:b] for k, a, b in self._slices}

    def _obs_dict(self, flat):
        """observation dict of the WHOLE batch (flat: [E, obs_dim]); with ndmap_obs also builds the ND-map of the live state"""
        d = self._slice_obs(flat)
        if self._nd is not None:      # the ND-map of the state this observation describes, consumed as an observation key
            self._nd_out = self.core.ndmap(out=self._nd_out, **self._nd)
            self.ndmap_last = self._nd_out
            d['ndmap_conflict_cells'] = self._nd_out[2].cpu().numpy().astype(self._obs_dtype) / self._obs_dtype.type(self._nd_cells)
        return d

    def _infos(self, info):
        info = np.asarray(info, dtype=np.float64)
        keys = INFO_KEYS[self.kind]
        d = {keys[0]: info[:, 0], keys[1]: info[:, 1].astype(np.int64), keys[2]: info[:, 2]}
        if len(keys) > 3:
            d[keys[3]] = info[:, 3].astype(np.int64)
        return d

    def reset(self, seed=None, options=None):
        # Gymnasium: reset(seed=s) restarts the generator (the same batch every time), reset() continues it -- every
        # unseeded reset after the first one draws a fresh batch (the seed walks on by a 64-bit odd constant)
        if seed is not None:
            self._seed = int(seed)
        elif self._resets:
            self._seed = (self._seed + 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
        self._resets += 1
        self.core.reset(seed=self._seed)
        self.core.set_autoreset(True, seed=(self._seed + 1) & 0xFFFFFFFFFFFFFFFF)
        obs = self.core.observe().cpu().numpy()
        E = self.num_envs
        nan = np.full(E, np.nan)
        third = np.zeros(E) if self.kind == L.KIND_VERTICAL else nan       # average_drift of an empty list is nan
        keys = INFO_KEYS[self.kind]
        d = {keys[0]: np.zeros(E), keys[1]: np.zeros(E, np.int64), keys[2]: third}
        if len(keys) > 3:
            d[keys[3]] = np.zeros(E, np.int64)
        return self._obs_dict(obs), d

    def step(self, actions):
        # host path of the C-ABI: pinned H2D of the actions, one launch, one D2H of every output into pinned host
        # memory; the views are converted (copied) to the reference's float64 / bool types before they are returned
        a = np.ascontiguousarray(np.asarray(actions, dtype=self.core.npdtype).reshape(self.num_envs, -1))
        obs, rew, term, trunc, info = self.core.step_host_views(a)
        return (self._obs_dict(obs), rew.astype(np.float64), term.astype(bool), trunc.astype(bool), self._infos(info))

    def step_tensors(self, actions):
        """actions: CUDA tensor [E, act_dim] in the core's dtype.  Returns the core's output tensors
        (obs [E, obs_dim], reward, terminated, truncated, info) without leaving the device or syncing."""
        return self.core.step(actions)

    def ndmap(self, **kw):
        """ND-map (4D occupancy / conflict grid over the look-ahead trajectories) of the batch's LIVE traffic:
        (occ u8 [E, nt, nz, ny, nx], conf bitmask, ncell [E, nt]) device tensors -- BsgCore.ndmap / bsg_ndmap_state.
        Keywords override the default grid NDMAP_DEFAULT."""
        return self.core.ndmap(**{**NDMAP_DEFAULT, **kw})

    def close(self):
        if not self.closed:
            self.core.close()
            self.closed = True


class _InfoView(Mapping):
    """infos[i] of the SB3 protocol for an env whose episode goes on: reads env i's slots of the adapter's current info
    arrays on demand, so a step builds no dict for it (SB3 only looks inside the infos of finished episodes, and those
    are real dicts).  Wrappers that write into an info (VecMonitor, VecNormalize, VecFrameStack) can: assignments go to a
    per-step overlay, and copy() gives a plain dict.  The values are those of the CURRENT step: copy() to keep them."""
    __slots__ = ('_o', '_i')

    def __init__(self, owner, i):
        self._o, self._i = owner, i

    def __getitem__(self, k):
        o, i = self._o, self._i
        over = o._over.get(i)
        if over is not None and k in over:
            return over[k]
        if k in o._info:
            v = o._info[k][i]
            return float(v) if o._info[k].dtype.kind == 'f' else int(v)
        if k == 'TimeLimit.truncated':
            return bool(o._tl[i])
        raise KeyError(k)

    def __setitem__(self, k, v):
        self._o._over.setdefault(self._i, {})[k] = v

    def __iter__(self):
        yield from self._o._info
        yield 'TimeLimit.truncated'
        for k in self._o._over.get(self._i, ()):
            if k not in self._o._info and k != 'TimeLimit.truncated':
                yield k

    def __len__(self):
        return sum(1 for _ in self)

    def copy(self):
        return dict(self)


try:  # the real base class when stable-baselines3 is installed (absent from this image): isinstance(env, VecEnv) holds
    import stable_baselines3 as _sb3
    if not hasattr(_sb3, '__version__'):
        raise ImportError('a stand-in, not stable-baselines3')
    from stable_baselines3.common.vec_env import VecEnv as _SB3Base
except ImportError:
    _SB3Base = object


class SB3VecEnvAdapter(_SB3Base):
    """stable-baselines3 VecEnv protocol (dict obs batched on axis 0, dones = terminated | truncated,
    same-step auto-reset with infos[i]['terminal_observation'] and infos[i]['TimeLimit.truncated']), the
    contract scripts/multi_processing_example.py:51-54 and bluesky_gym/utils/logger.py:15-35 rely on.
    Vectorised: infos[i] of an env whose episode goes on is a mapping view over the step's info arrays (no dict is built
    for it); a FINISHED env gets a real dict (SB3's wrappers copy and extend those), and only the finished envs' rows
    travel for the terminal / fresh observations.  With ndmap_obs the key 'ndmap_conflict_cells' of the returned batch is
    the map of the state after the device's resets (the fresh episode of a finished env); the terminal observation
    carries that same row, since the terminal traffic is gone when the step returns."""

    def __init__(self, venv: BlueSkyVecEnv):
        if _SB3Base is not object:
            _SB3Base.__init__(self, venv.num_envs, venv.single_observation_space, venv.single_action_space)
        self.venv, self.num_envs = venv, venv.num_envs
        self.observation_space, self.action_space = venv.single_observation_space, venv.single_action_space
        self.render_mode = None                      # no graphics stack: nothing to render
        self._actions = None
        self._info, self._tl, self._over = {}, np.zeros(self.num_envs, bool), {}
        self._views = [_InfoView(self, i) for i in range(self.num_envs)]

    def reset(self):
        obs, _ = self.venv.reset()
        self.venv.core.set_autoreset(2, seed=(self.venv._seed + 1) & 0xFFFFFFFFFFFFFFFF)          # same-step: reset now, no skipped step
        return obs

    def step_async(self, actions):
        self._actions = actions

    def step_wait(self):
        v = self.venv
        obs, rew, term, trunc, info = v.step(self._actions)
        dones = term | trunc
        self._info, self._tl, self._over = info, trunc & ~term, {}
        infos = self._views
        idx = np.flatnonzero(dones)
        if idx.size:
            # the device already reset those envs: their fresh observations, rows of the finished envs only
            rows = v.core.observe()[torch.as_tensor(idx, device=v.core.device)].cpu().numpy()
            fresh = v._slice_obs(rows)
            infos = list(self._views)
            for i in idx:
                d = dict(self._views[i])
                d['terminal_observation'] = {k: obs[k][i].copy() for k in obs}
                infos[int(i)] = d
            for k in fresh:
                obs[k][idx] = fresh[k]
        return obs, rew, dones, infos

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    def close(self):
        self.venv.close()

    # the rest of SB3's VecEnv surface.  There are no per-env Python objects behind the batch: an attribute or method
    # of "env i" is the batch's, answered once per requested index.
    def _indices(self, indices):
        if indices is None:
            return list(range(self.num_envs))
        return [int(indices)] if np.isscalar(indices) else [int(i) for i in indices]

    def get_attr(self, attr_name, indices=None):
        v = getattr(self, attr_name) if attr_name in ('render_mode',) else getattr(self.venv, attr_name)
        return [v for _ in self._indices(indices)]

    def set_attr(self, attr_name, value, indices=None):
        if len(self._indices(indices)) != self.num_envs:
            raise ValueError("the envs of a device batch share their attributes: set_attr needs indices=None")
        setattr(self.venv, attr_name, value)

    def env_method(self, method_name, *method_args, indices=None, **method_kwargs):
        out = getattr(self.venv, method_name)(*method_args, **method_kwargs)
        return [out for _ in self._indices(indices)]

    def env_is_wrapped(self, wrapper_class, indices=None):
        return [False for _ in self._indices(indices)]

    def seed(self, seed=None):
        """seed of the NEXT reset (env i draws stream (seed, i) on the device); returns the per-env seeds SB3 expects"""
        if seed is not None:
            self.venv._seed, self.venv._resets = int(seed), 0
        return [self.venv._seed + i for i in range(self.num_envs)]

    def get_images(self):
        return [None for _ in range(self.num_envs)]

    def render(self, mode=None):
        return None


class RLlibVectorEnvAdapter:
    """ray.rllib.env.vector_env.VectorEnv protocol over one device batch: vector_reset / reset_at /
    vector_step with per-env LISTS of observations, the shape RLlib's sampler consumes.  The device resets finished
    envs itself (same-step mode), so reset_at(i) returns the observation env i already holds.  ray is not installed in
    this image: the class follows the published protocol; tests/test_gpu_vecenv.py::test_rllib_adapter drives it."""

    def __init__(self, venv: BlueSkyVecEnv):
        self.venv, self.num_envs = venv, venv.num_envs
        self.observation_space, self.action_space = venv.single_observation_space, venv.single_action_space
        self._obs = None

    def _split(self, obs):
        return [{k: obs[k][i] for k in obs} for i in range(self.num_envs)]

    def vector_reset(self, *, seeds=None, options=None):
        obs, info = self.venv.reset(seed=None if not seeds else seeds[0])
        self.venv.core.set_autoreset(2, seed=(self.venv._seed + 1) & 0xFFFFFFFFFFFFFFFF)
        self._obs = obs
        return self._split(obs), [{} for _ in range(self.num_envs)]

    def reset_at(self, index=None, *, seed=None, options=None):
        i = 0 if index is None else int(index)
        fresh = self.venv._slice_obs(self.venv.core.observe()[i:i + 1].cpu().numpy())
        out = {k: fresh[k][0] for k in fresh}
        if self._obs is not None and 'ndmap_conflict_cells' in self._obs:      # the map built for this step, env i's row
            out['ndmap_conflict_cells'] = self._obs['ndmap_conflict_cells'][i]
        return out, {}

    def vector_step(self, actions):
        a = np.stack([np.asarray(x, dtype=np.float64).reshape(-1) for x in actions])
        obs, rew, term, trunc, info = self.venv.step(a)
        self._obs = obs
        keys = list(info)
        cols = [info[k].tolist() for k in keys]                   # one conversion per key, not per env and key
        infos = [dict(zip(keys, row)) for row in zip(*cols)]
        return self._split(obs), rew.tolist(), term.tolist(), trunc.tolist(), infos

    def get_sub_environments(self):
        return []

    def close(self):
        self.venv.close()
