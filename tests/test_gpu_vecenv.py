"""GPU tests of the host-side mirror of the reference interface: the registered single-env ids, the batched
VectorEnv (next-step auto-reset) and the SB3-style adapter (same-step auto-reset, terminal_observation)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_registered_single_envs_speak_the_reference_api():
    import bluesky_gym_hallucination_b200 as bluesky_gym
    bluesky_gym.register_envs()
    for env_id, act_dim, third in (('HorizontalCREnv-v0', 1, 'average_drift'), ('VerticalCREnv-v0', 1, 'final_altitude'),
                                   ('SectorCREnv-v0', 2, 'average_drift')):
        env = bluesky_gym.make(env_id, render_mode=None)
        obs, info = env.reset(seed=3)
        assert set(obs) == set(env.observation_space.spaces.keys())
        for k, v in obs.items():
            assert v.dtype == np.float64 and v.shape == env.observation_space[k].shape
        assert set(info) == {'total_reward', 'total_intrusions', third}
        if third == 'average_drift':
            assert np.isnan(info[third])                       # mean of an empty list right after reset
        tot = 0.0
        for _ in range(5):
            obs, rew, term, trunc, info = env.step(np.zeros(act_dim))
            assert isinstance(rew, float) and term in (0, 1) and isinstance(trunc, bool)
            tot += rew
        assert abs(info['total_reward'] - tot) < 1e-9
        env.close()
    with pytest.raises(NotImplementedError):
        bluesky_gym.make('HorizontalCREnv-v0', render_mode='human')


def test_horizontal_straight_flight_reaches_the_waypoint_like_the_reference_logs():
    """Anchor from the shipped logs: flying straight, Horizontal episodes end after ~24-37 steps."""
    import bluesky_gym_hallucination_b200 as bluesky_gym
    bluesky_gym.register_envs()
    env = bluesky_gym.make('HorizontalCREnv-v0')
    lens = []
    for ep in range(4):
        env.reset(seed=100 + ep)
        for n in range(1, 301):
            _, _, term, trunc, _ = env.step(np.zeros(1))
            if term or trunc:
                break
        lens.append(n)
    assert all(22 <= n <= 40 for n in lens), lens
    env.close()


def test_time_limit_truncates_single_env():
    from bluesky_gym_hallucination_b200.vec_env import BlueSkyVecEnv
    v = BlueSkyVecEnv('SectorCREnv-v0', 8, max_episode_steps=3)
    v.reset(seed=1)
    for s in range(3):
        obs, rew, term, trunc, info = v.step(np.zeros((8, 2)))
    assert trunc.all()
    v.close()


def test_vector_env_next_step_autoreset():
    from bluesky_gym_hallucination_b200.vec_env import BlueSkyVecEnv
    E = 64
    v = BlueSkyVecEnv('HorizontalCREnv-v0', E, num_aircraft=6, precision='fp32', seed=5)
    obs, info = v.reset(seed=5)
    assert obs['intruder_distance'].shape == (E, 5) and obs['cos_drift'].shape == (E, 1)
    done_prev = np.zeros(E, bool)
    ended = 0
    for s in range(60):
        obs, rew, term, trunc, info = v.step(np.zeros((E, 1)))
        # the step after a done returns the first observation of the new episode: reward 0, flags 0, counters 0
        assert np.all(rew[done_prev] == 0.0) and not term[done_prev].any() and not trunc[done_prev].any()
        assert np.all(info['total_reward'][done_prev] == 0.0)
        assert np.all(obs['waypoint_distance'][done_prev, 0] > 95.0 / 150.0)      # fresh waypoint 100-149 km away
        done_prev = term | trunc
        ended += int(done_prev.sum())
    assert ended >= E            # straight flight ends every episode within ~37 steps
    t = v.step_tensors(torch.zeros((E, 1), dtype=torch.float32, device='cuda'))
    assert t[0].is_cuda and t[0].shape == (E, 28)
    v.close()


def test_sb3_adapter_same_step_autoreset():
    from bluesky_gym_hallucination_b200.vec_env import BlueSkyVecEnv, SB3VecEnvAdapter
    E = 32
    a = SB3VecEnvAdapter(BlueSkyVecEnv('HorizontalCREnv-v0', E, precision='fp32', seed=9))
    obs = a.reset()
    seen = 0
    for s in range(45):
        obs, rew, dones, infos = a.step(np.zeros((E, 1)))
        assert set(infos[0]) >= {'total_reward', 'total_intrusions', 'average_drift', 'TimeLimit.truncated'}
        for i in np.where(dones)[0]:
            seen += 1
            assert 'terminal_observation' in infos[i]
            assert infos[i]['terminal_observation']['waypoint_distance'][0] < 5.0 / 150.0    # reached the waypoint
            assert obs['waypoint_distance'][i, 0] > 95.0 / 150.0                             # already the new episode
    assert seen >= E // 2
    a.close()


def _sb3_base():
    """the restated abstract surface of stable_baselines3's VecEnv (oracle/refshim: test infrastructure)"""
    import importlib.util
    import os
    from conftest import ROOT
    path = os.path.join(ROOT, 'oracle', 'refshim', 'stable_baselines3', 'common', 'vec_env', '__init__.py')
    spec = importlib.util.spec_from_file_location('_sb3_vec_env_surface', path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.VecEnv


def test_sb3_adapter_has_the_vecenv_surface_and_survives_its_wrappers():
    """scripts/multi_processing_example.py:51-54 hands the env to SB3: every abstract method of VecEnv is there with
    SB3's signature, and the infos behave like the dicts VecMonitor / VecNormalize copy and extend."""
    import inspect
    from bluesky_gym_hallucination_b200.vec_env import BlueSkyVecEnv, SB3VecEnvAdapter
    base = _sb3_base()
    for name in base.__abstractmethods__ | {'step', 'seed', 'get_images', 'render'}:
        want = list(inspect.signature(getattr(base, name)).parameters)
        got = list(inspect.signature(getattr(SB3VecEnvAdapter, name)).parameters)
        assert got == want, (name, got, want)
    E = 8
    a = SB3VecEnvAdapter(BlueSkyVecEnv('HorizontalCREnv-v0', E, precision='fp32', seed=4, max_episode_steps=2))
    assert a.render_mode is None and a.get_attr('render_mode') == [None] * E
    assert a.get_attr('num_aircraft', indices=[0, 3]) == [6, 6] and a.env_is_wrapped(object) == [False] * E
    assert a.seed(21) == list(range(21, 21 + E))
    first = a.reset()
    assert len(a.env_method('_infos', np.zeros((E, 4)), indices=0)) == 1
    obs, rew, dones, infos = a.step(np.zeros((E, 1)))
    assert not dones.any()
    kept = infos[0].copy()                                   # VecMonitor: info.copy(), then info['episode'] = ...
    assert isinstance(kept, dict) and 'total_reward' in kept
    infos[1]['episode'] = {'r': 1.0}                         # a wrapper writing into a running env's info
    assert infos[1]['episode'] == {'r': 1.0} and 'episode' in dict(infos[1]) and 'episode' not in dict(infos[2])
    obs, rew, dones, infos = a.step(np.zeros((E, 1)))
    assert dones.all()                                       # TimeLimit at 2 steps
    for i in range(E):
        assert type(infos[i]) is dict and infos[i]['TimeLimit.truncated'] is True
        infos[i]['terminal_observation'] = dict(infos[i]['terminal_observation'])      # VecNormalize rewrites it
        ep = infos[i].copy(); ep['episode'] = {'r': float(rew[i]), 'l': 2}
    assert 'episode' not in infos[1]                         # last step's overlay is gone
    a.reset()                                                # seed(21) again? no: unseeded reset continues the stream
    a.seed(21)
    again = a.reset()
    for k in first:
        assert np.array_equal(first[k], again[k]), k
    a.close()


def test_sb3_adapter_with_ndmap_obs():
    """with ndmap_obs the adapters slice only the finished envs' flat rows; the ND-map key keeps the batch shape
    (it used to raise a broadcast error whenever 1 < finished < E)."""
    from bluesky_gym_hallucination_b200.vec_env import BlueSkyVecEnv, SB3VecEnvAdapter, RLlibVectorEnvAdapter
    E = 12
    v = BlueSkyVecEnv('HorizontalCREnv-v0', E, precision='fp32', seed=2, ndmap_obs={'nt': 6})
    a = SB3VecEnvAdapter(v)
    obs = a.reset()
    n_partial = 0
    for s in range(45):                                     # flying straight, episodes end after 24-37 steps, env by env
        obs, rew, dones, infos = a.step(np.zeros((E, 1)))
        assert obs['ndmap_conflict_cells'].shape == (E, 6)
        if 0 < dones.sum() < E:
            n_partial += 1
            i = int(np.flatnonzero(dones)[0])
            assert infos[i]['terminal_observation']['ndmap_conflict_cells'].shape == (6,)
            # the fresh rows are the device's current observation of those envs
            cur = v.core.observe().cpu().numpy()
            for k, lo, hi in v._slices:
                assert np.array_equal(obs[k][i], cur[i, lo:hi].astype(obs[k].dtype)), k
    assert n_partial > 0
    a.close()
    r = RLlibVectorEnvAdapter(BlueSkyVecEnv('SectorCREnv-v0', 4, precision='fp32', seed=2, ndmap_obs={'nt': 6}))
    r.vector_reset(seeds=[5])
    o, *_ = r.vector_step([np.zeros(2)] * 4)
    o2, _ = r.reset_at(2)
    assert np.array_equal(o2['ndmap_conflict_cells'], o[2]['ndmap_conflict_cells'])       # env 2's row, not env 0's
    r.close()


def test_rllib_adapter():
    """ray.rllib VectorEnv protocol (f-2): lists of per-env observations / rewards / flags / info dicts; finished envs are
    reset on the device and reset_at(i) hands out env i's first observation of the new episode."""
    from bluesky_gym_hallucination_b200.vec_env import BlueSkyVecEnv, RLlibVectorEnvAdapter
    E = 6
    r = RLlibVectorEnvAdapter(BlueSkyVecEnv('HorizontalCREnv-v0', E, precision='fp32', seed=8, max_episode_steps=3))
    obs, infos = r.vector_reset(seeds=[8])
    assert len(obs) == E and len(infos) == E and r.get_sub_environments() == []
    assert set(obs[0]) == set(r.observation_space.spaces) and obs[0]['intruder_distance'].shape == (5,)
    tot = np.zeros(E)
    for s in range(3):
        obs, rew, term, trunc, infos = r.vector_step([np.zeros(1) for _ in range(E)])
        assert all(isinstance(x, list) and len(x) == E for x in (obs, rew, term, trunc, infos))
        tot += np.array(rew)
        assert all(type(i) is dict and set(i) == {'total_reward', 'total_intrusions', 'average_drift'} for i in infos)
        assert all(type(i['total_intrusions']) is int and type(i['total_reward']) is float for i in infos)
    assert all(trunc) and not any(term)
    assert np.allclose([i['total_reward'] for i in infos], tot, atol=1e-5)
    fresh, _ = r.reset_at(4)
    assert fresh['waypoint_distance'][0] > 95.0 / 150.0 and not np.array_equal(fresh['waypoint_distance'], obs[4]['waypoint_distance'])
    cur = r.venv.core.observe().cpu().numpy()
    for k, lo, hi in r.venv._slices:
        assert np.array_equal(fresh[k], cur[4, lo:hi].astype(fresh[k].dtype)), k
    r.close()


def test_vector_env_spaces_are_batched():
    """Gymnasium VectorEnv: observation_space / action_space are the batched spaces, single_* the per-env ones."""
    from bluesky_gym_hallucination_b200.vec_env import BlueSkyVecEnv
    v = BlueSkyVecEnv('SectorCREnv-v0', 5, precision='fp32', seed=1)
    obs, _ = v.reset()
    assert v.action_space.shape == (5, 2) and v.single_action_space.shape == (2,)
    for k, b in v.observation_space.spaces.items():
        assert b.shape == (5,) + v.single_observation_space[k].shape == obs[k].shape
    v.close()


def test_same_step_autoreset_draws_a_new_scenario_every_episode():
    """the same-step auto-reset (SB3 / RLlib adapters) must advance the env's episode counter -- two
    consecutive auto-reset episodes of one env start from different scenarios, and so do different envs."""
    from bluesky_gym_hallucination_b200.vec_env import BlueSkyVecEnv, SB3VecEnvAdapter
    E = 16
    a = SB3VecEnvAdapter(BlueSkyVecEnv('HorizontalCREnv-v0', E, precision='fp32', seed=5, max_episode_steps=3))
    a.reset()
    starts = [[] for _ in range(E)]          # first observation of every auto-reset episode, per env
    for s in range(12):
        obs, rew, dones, infos = a.step(np.zeros((E, 1)))
        for i in np.where(dones)[0]:
            starts[i].append(np.concatenate([obs[k][i] for k in sorted(obs)]))
    a.close()
    assert all(len(x) >= 3 for x in starts)
    for x in starts:
        for u, v in zip(x[:-1], x[1:]):
            assert not np.allclose(u, v), "an env replayed the same scenario in consecutive episodes"
    assert not np.allclose(starts[0][0], starts[1][0])


def test_reset_with_a_seed_repeats_and_without_a_seed_continues():
    """Gymnasium contract: reset(seed=s) twice gives the same batch; reset() after it draws a new one."""
    from bluesky_gym_hallucination_b200.vec_env import BlueSkyVecEnv
    v = BlueSkyVecEnv('SectorCREnv-v0', 8, precision='fp32', seed=3)
    flat = lambda o: np.concatenate([o[k].ravel() for k in sorted(o)])
    a = flat(v.reset(seed=11)[0]); b = flat(v.reset(seed=11)[0])
    c = flat(v.reset()[0]); d = flat(v.reset()[0]); e = flat(v.reset(seed=11)[0])
    v.close()
    assert np.array_equal(a, b) and np.array_equal(a, e)
    assert not np.allclose(a, c) and not np.allclose(c, d)


def _episodes(kind, N, E, steps, seed, max_steps, act_dim, **kw):
    """Random-policy episodes of a device batch: returns (lengths, total_intrusions, third info slot) of every finished episode."""
    from bluesky_gym_hallucination_b200.core import BsgCore
    c = BsgCore(kind, E, N, precision='fp32', max_episode_steps=max_steps, **kw)
    c.reset(seed=seed); c.set_autoreset(True, seed=seed + 1)
    gen = torch.Generator(device='cuda'); gen.manual_seed(seed)
    age = np.zeros(E, np.int64)
    lens, intr, third = [], [], []
    for s in range(steps):
        a = torch.rand((E, act_dim), device='cuda', generator=gen) * 2 - 1
        c.step(a)
        done = (c.terminated | c.truncated).cpu().numpy().astype(bool)
        info = c.info.cpu().numpy()
        # next-step auto-reset: an env that ended in the previous step spends this step drawing its episode
        fresh = age < 0
        age[~fresh] += 1
        age[fresh] = 0
        for e in np.where(done & ~fresh)[0]:
            lens.append(age[e]); intr.append(info[e, 1]); third.append(info[e, 2])
        age[done & ~fresh] = -1
    c.close()
    return np.array(lens), np.array(intr), np.array(third)


def test_vertical_episode_lengths_match_the_reference_logs():
    """Envelope anchor (scripts/common/results/logs_backup/VerticalCREnv-v0/*.csv, 200 000 episodes of four
    agents): a Vertical episode is a 200 km run at CAS 150 m/s in 30 s steps -- median length 40, NEVER more than 43-44
    (TAS >= 155 m/s near the ground), and never fewer than 6 steps (a crash needs 2000 m of descent at 12.5 m/s)."""
    from bluesky_gym_hallucination_b200 import _lib as L
    lens, intr, final_alt = _episodes(L.KIND_VERTICAL, 6, 512, 100, 41, 300, 1)
    assert len(lens) >= 512
    assert lens.max() <= 44 and lens.min() >= 6, (lens.min(), lens.max())
    assert 37 <= np.median(lens) <= 42, np.median(lens)
    # logged first-300-episode means (near-random policies): 0.98 - 1.66 intrusion steps per episode
    assert 0.3 <= intr.mean() <= 3.0, intr.mean()


def test_sector_random_policy_statistics_match_the_reference_logs():
    """Anchor (logs_backup/SectorCREnv-v0/*.csv, first 300 episodes of each agent = near-random policies): 14 - 40
    intrusion steps per episode on average, episodes from a few steps (the agent leaves the sector) to the 200-step
    TimeLimit; the agent holds ~150 m/s CAS, so the speed slot of the observation (tas - 150) / 6 stays O(1)."""
    from bluesky_gym_hallucination_b200 import _lib as L
    from bluesky_gym_hallucination_b200.core import BsgCore
    lens, intr, drift = _episodes(L.KIND_SECTOR, 32, 256, 420, 43, 200, 2)
    assert len(lens) >= 256
    assert lens.max() <= 200 and 30 <= np.median(lens) <= 200, (np.median(lens), lens.max())
    assert 8.0 <= intr.mean() <= 60.0, intr.mean()
    assert 0.2 <= np.nanmean(drift) <= 2.0, np.nanmean(drift)
    c = BsgCore(L.KIND_SECTOR, 64, 32, precision='fp32')
    c.reset(seed=5)
    obs = c.observe().cpu().numpy()
    ac, _, nac, _ = c.get_state()
    # sector_cr_env.py:303: "airspeed": (tas - 150) / 6 -- CAS 150 m/s at the sector's flight level is TAS ~ 150 - 190
    tas0 = ac[:, L.F['tas'], 0]
    assert np.all((tas0 > 140) & (tas0 < 200)), (tas0.min(), tas0.max())
    c.close()


def _reset_stats(kind, ac, nac, ev, F, EV):
    """Per-kind derived quantities of a batch of freshly reset envs (same code for the reference samples and the device
    generator): dict name -> 1-D sample."""
    lat, lon, alt, hdg, tas = (ac[:, F[k]] for k in ('lat', 'lon', 'alt', 'hdg', 'tas'))
    E = ac.shape[0]
    live = np.arange(ac.shape[2])[None, :] < nac[:, None]
    intr = live.copy(); intr[:, 0] = False
    c0 = np.cos(np.radians(lat[:, :1]))
    dx = (lon - lon[:, :1]) * 60.0 * c0; dy = (lat - lat[:, :1]) * 60.0               # NM from the agent
    out = {'intruder_dist_nm': np.hypot(dx, dy)[intr], 'intruder_bearing': (np.degrees(np.arctan2(dx, dy)) % 360)[intr],
           'intruder_hdg_rel': ((hdg - hdg[:, :1]) % 360)[intr], 'intruder_tas': tas[intr]}
    if kind == 'H':
        wd = np.hypot((ev[:, EV['wpt_lon']] - lon[:, 0]) * 60 * c0[:, 0], (ev[:, EV['wpt_lat']] - lat[:, 0]) * 60)
        out['waypoint_dist_nm'] = wd
    if kind == 'V':
        out['own_alt'] = alt[:, 0]; out['target_minus_alt'] = ev[:, EV['target_alt']] - alt[:, 0]
        out['intruder_dalt'] = (alt - alt[:, :1])[intr]
    if kind == 'S':
        out['nac'] = nac.astype(np.float64); out['own_tas'] = tas[:, 0]; out['own_hdg'] = hdg[:, 0]
        out['all_alt'] = alt[live]
    return out


@pytest.mark.parametrize('kind', ['H', 'V', 'S'])
def test_reset_distributions_match_the_reference(kind):
    """The device scenario generators (csrc/bsg_reset.cuh, Philox) draw from the same laws as the UNMODIFIED reference
    reset() (horizontal_cr_env.py:121-160, vertical_cr_env.py:182-199,262-281, sector_cr_env.py:141-230): two-sample
    Kolmogorov-Smirnov of every derived quantity against samples recorded from the reference classes
    (tests/golden/reset_samples_*.npz, made by tests/golden/make_reset_samples.py)."""
    import os
    from scipy import stats
    from bluesky_gym_hallucination_b200.core import BsgCore
    from bluesky_gym_hallucination_b200 import _lib as L
    g = np.load(os.path.join(os.path.dirname(__file__), 'golden', f'reset_samples_{kind}.npz'))
    fields = [str(x) for x in g['fields']]
    Fg = {k: i for i, k in enumerate(fields)}
    n = g['ac'].shape[2]
    ref = _reset_stats(kind, g['ac'], g['nac'], g['ev'], Fg, L.EV)
    E = 4096
    c = BsgCore({'H': L.KIND_HORIZONTAL, 'V': L.KIND_VERTICAL, 'S': L.KIND_SECTOR}[kind], E, n, precision='fp32')
    c.reset(seed=77)
    torch.cuda.synchronize()
    ac, _, nac, ev = c.get_state()
    c.close()
    dev = _reset_stats(kind, ac, nac, ev, L.F, L.EV)
    report = {}
    for k in ref:
        a, b = np.asarray(ref[k], np.float64), np.asarray(dev[k], np.float64)
        a, b = a[np.isfinite(a)], b[np.isfinite(b)]
        if max(a.std(), b.std()) <= 1e-6 * max(abs(a.mean()), 1.0):       # a constant of the generator: compare it as such
            assert abs(a.mean() - b.mean()) <= 1e-6 * max(abs(a.mean()), 1.0), (k, a.mean(), b.mean())
            continue
        ks = stats.ks_2samp(a, b)
        # critical value of the two-sample statistic at alpha = 1e-3 (draws inside one env are not independent, so the
        # effective sample is taken as the number of ENVS, not aircraft)
        na, nb = min(len(a), len(g['nac'])), min(len(b), E)
        crit = 1.95 * np.sqrt((na + nb) / (na * nb))
        report[k] = (round(float(ks.statistic), 4), round(float(crit), 4), float(a.mean()), float(b.mean()))
    print(f"[reset {kind}] KS statistic / critical (alpha 1e-3) / reference mean / device mean: {report}")
    bad = {k: v for k, v in report.items() if v[0] > v[1]}
    assert not bad, bad


def test_ndmap_is_consumed_as_an_observation_key():
    """ndmap_obs: every reset / step builds the ND-map of the live traffic and the observation carries the share of grid
    cells in conflict per look-ahead step; the key equals a stand-alone build of the same state, follows the traffic from
    step to step, and the registered single envs take the option through make()."""
    import bluesky_gym_hallucination_b200 as bluesky_gym
    from bluesky_gym_hallucination_b200 import BlueSkyVecEnv
    E = 16
    v = BlueSkyVecEnv('SectorCREnv-v0', E, num_aircraft=24, ndmap_obs=True, seed=5, fixed_nac=True)
    assert v.single_observation_space['ndmap_conflict_cells'].shape == (30,)
    obs, _ = v.reset(seed=5)
    key0 = obs['ndmap_conflict_cells']
    assert key0.shape == (E, 30) and key0.dtype == np.float64 and (key0 >= 0).all() and (key0 <= 1).all()
    occ, conf, ncell = v.ndmap(**v._nd)                                   # stand-alone build of the same live state
    assert np.array_equal(key0, ncell.cpu().numpy() / float(64 * 64 * 8))
    assert (key0 > 0).any()                                               # 24 aircraft in one sector: conflicts ahead
    assert int(occ.sum().item()) > 0 and v.ndmap_last is not None
    rng = np.random.default_rng(0)
    for _ in range(3):
        obs, rew, term, trunc, info = v.step(rng.uniform(-1, 1, (E, 2)))
    assert not np.array_equal(obs['ndmap_conflict_cells'], key0)          # the map follows the traffic
    assert set(obs) == set(v.single_observation_space.spaces)
    v.close()
    bluesky_gym.register_envs()
    env = bluesky_gym.make('HorizontalCREnv-v0', ndmap_obs=dict(nt=12))
    o, _ = env.reset(seed=1)
    assert o['ndmap_conflict_cells'].shape == (12,) and 'ndmap_conflict_cells' in env.observation_space.spaces
    o, r, te, tr, i = env.step(np.zeros(1))
    assert o['ndmap_conflict_cells'].shape == (12,)
    env.close()


def test_float32_observations_are_views_of_the_pinned_block():
    """obs_dtype=float32: same keys, shapes and values as the float64 dict, declared float32, and no copy -- the arrays
    are views of the pinned host block the step filled."""
    from bluesky_gym_hallucination_b200 import BlueSkyVecEnv
    E = 64
    v64 = BlueSkyVecEnv('HorizontalCREnv-v0', E, num_aircraft=20, asas=True, seed=3)
    v32 = BlueSkyVecEnv('HorizontalCREnv-v0', E, num_aircraft=20, asas=True, seed=3, obs_dtype=np.float32)
    assert all(b.dtype == np.float32 for b in v32.single_observation_space.spaces.values())
    o64, _ = v64.reset(seed=3); o32, _ = v32.reset(seed=3)
    rng = np.random.default_rng(0)
    for _ in range(3):
        a = rng.uniform(-1, 1, (E, 1))
        o64, r64, t64, tr64, i64 = v64.step(a)
        o32, r32, t32, tr32, i32 = v32.step(a)
    assert set(o32) == set(o64)
    views = [b for b in v32.core._blocks if b is not None]
    for k in o64:
        assert o32[k].dtype == np.float32 and o32[k].shape == o64[k].shape
        assert np.array_equal(o32[k].astype(np.float64), o64[k], equal_nan=True)
        assert any(np.shares_memory(o32[k], blk[0]) for blk in views), k      # a view, not a copy
    assert np.array_equal(r32, r64) and np.array_equal(t32, t64)
    v64.close(); v32.close()

