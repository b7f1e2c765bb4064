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
        self.observation_space, self.action_space = self.single_observation_space, self.single_action_space
        self._slices = obs_slices(self.kind, self.num_aircraft, wind_obs)

    # ------------------------------------------------------------------ Gymnasium VectorEnv API
    def _obs_dict(self, flat):
        flat = np.asarray(flat, dtype=self._obs_dtype)          # float32 from the fp32 core: a view, not a copy
        d = {k: flat[:, a:b] for k, a, b in self._slices}
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
        if seed is not None:
            self._seed = int(seed)
        self.core.reset(seed=self._seed)
        self.core.set_autoreset(True, seed=self._seed + 1)
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
    """infos[i] of the SB3 protocol as a VIEW: reads env i's slots of the adapter's current info arrays on demand, so a
    step allocates nothing per env (SB3 only looks inside the infos of finished episodes)."""
    __slots__ = ('_o', '_i')

    def __init__(self, owner, i):
        self._o, self._i = owner, i

    def __getitem__(self, k):
        o, i = self._o, self._i
        if k in o._info:
            v = o._info[k][i]
            return float(v) if o._info[k].dtype.kind == 'f' else int(v)
        if k == 'TimeLimit.truncated':
            return bool(o._tl[i])
        if k == 'terminal_observation' and i in o._term_obs:
            return o._term_obs[i]
        raise KeyError(k)

    def __iter__(self):
        yield from self._o._info
        yield 'TimeLimit.truncated'
        if self._i in self._o._term_obs:
            yield 'terminal_observation'

    def __len__(self):
        return len(self._o._info) + 1 + (self._i in self._o._term_obs)


class SB3VecEnvAdapter:
    """stable-baselines3 VecEnv protocol (dict obs batched on axis 0, dones = terminated | truncated,
    same-step auto-reset with infos[i]['terminal_observation'] and infos[i]['TimeLimit.truncated']), the
    contract scripts/multi_processing_example.py:51-54 and bluesky_gym/utils/logger.py:15-35 rely on.
    Vectorised: `infos` is a list of E persistent mapping views over the step's info arrays (no per-env dict is built),
    and only the finished envs' rows travel for the terminal / fresh observations."""

    def __init__(self, venv: BlueSkyVecEnv):
        self.venv, self.num_envs = venv, venv.num_envs
        self.observation_space, self.action_space = venv.single_observation_space, venv.single_action_space
        self._actions = None
        self._info, self._tl, self._term_obs = {}, np.zeros(self.num_envs, bool), {}
        self._views = [_InfoView(self, i) for i in range(self.num_envs)]

    def reset(self):
        obs, _ = self.venv.reset()
        self.venv.core.set_autoreset(2, seed=self.venv._seed + 1)          # same-step: reset now, no skipped step
        return obs

    def step_async(self, actions):
        self._actions = actions

    def step_wait(self):
        v = self.venv
        obs, rew, term, trunc, info = v.step(self._actions)
        dones = term | trunc
        self._info, self._tl, self._term_obs = info, trunc & ~term, {}
        idx = np.flatnonzero(dones)
        if idx.size:
            # the device already reset those envs: their fresh observations, rows of the finished envs only
            rows = v.core.observe()[torch.as_tensor(idx, device=v.core.device)].cpu().numpy()
            fresh = v._obs_dict(rows)
            for n, i in enumerate(idx):
                self._term_obs[int(i)] = {k: obs[k][i].copy() for k in obs}
            for k in obs:
                obs[k][idx] = fresh[k]
        return obs, rew, dones, self._views

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    def close(self):
        self.venv.close()


class RLlibVectorEnvAdapter:
    """ray.rllib.env.vector_env.VectorEnv protocol over one device batch: vector_reset / reset_at /
    vector_step with per-env LISTS of observations, the shape RLlib's sampler consumes.  The device resets finished
    envs itself (same-step mode), so reset_at(i) returns the observation env i already holds.  ray is not installed in
    this image: the class follows the published protocol and is exercised by tests/test_gpu_vecenv.py only."""

    def __init__(self, venv: BlueSkyVecEnv):
        self.venv, self.num_envs = venv, venv.num_envs
        self.observation_space, self.action_space = venv.single_observation_space, venv.single_action_space
        self._obs = None

    def _split(self, obs):
        return [{k: obs[k][i] for k in obs} for i in range(self.num_envs)]

    def vector_reset(self, *, seeds=None, options=None):
        obs, info = self.venv.reset(seed=None if not seeds else seeds[0])
        self.venv.core.set_autoreset(2, seed=self.venv._seed + 1)
        self._obs = obs
        return self._split(obs), [{} for _ in range(self.num_envs)]

    def reset_at(self, index=None, *, seed=None, options=None):
        i = 0 if index is None else int(index)
        fresh = self.venv._obs_dict(self.venv.core.observe()[i:i + 1].cpu().numpy())
        return {k: fresh[k][0] for k in fresh}, {}

    def vector_step(self, actions):
        a = np.stack([np.asarray(x, dtype=np.float64).reshape(-1) for x in actions])
        obs, rew, term, trunc, info = self.venv.step(a)
        self._obs = obs
        infos = [{k: (float(info[k][i]) if info[k].dtype.kind == 'f' else int(info[k][i])) for k in info}
                 for i in range(self.num_envs)]
        return self._split(obs), rew.tolist(), term.tolist(), trunc.tolist(), infos

    def get_sub_environments(self):
        return []

    def close(self):
        self.venv.close()
