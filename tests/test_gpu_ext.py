"""GPU parity tests of the widened rows (SURVEY 8 f-3 MergeEnv, f-4 wind field + noisy observations): the EXT
instantiations of the fused step kernel, through the C-ABI, against the goldens recorded from the UNMODIFIED
reference classes (bluesky_gym/envs/merge_env.py, bluesky_gym/wrappers/wind.py) and against the CPU oracle.
Tolerances as in test_gpu_parity.py: fp64 1e-12 (x accumulated libm differences), fp32 1e-5."""
import numpy as np
import pytest
import torch
from conftest import load_golden
from test_gpu_parity import relerr, obs_tol32, run_step, RT64, RT32

pytestmark = pytest.mark.gpu
KIND = {'H': 0, 'V': 1, 'S': 2, 'M': 4}
EXT_GOLDEN = ['M', 'M_wind', 'S_wind', 'V_wind', 'H_wind', 'S_asas_wind', 'V_windalt', 'S_asas_windalt']
STATE_FIELDS = ['lat', 'lon', 'alt', 'hdg', 'tas', 'vs', 'selspd', 'selalt', 'selvs', 'aptrk', 'ax', 'asas_trk',
                'asas_tas', 'asas_vs', 'asas_alt', 'asas_active', 'phase', 'gsn', 'gse', 'lnav']


def golden_core(name, g, precision):
    from bluesky_gym_hallucination_b200.core import BsgCore
    kind, asas = name[0], 'asas' in name
    T, _, n = g['ac0'].shape
    core = BsgCore(KIND[kind], T, n, precision=precision, asas_on=asas, reso_on=asas, reso_off_own=asas, cd_outputs=True)
    if 'wind' in g:
        w = g['wind']
        core.set_wind(w[:, 0], w[:, 1], w[:, 2], w[:, 3], augment_obs=True)
    if 'windprof' in g:      # WindFieldWrapper(env, lat, lon, vnorth[], veast[], alt=[...]): one point, altitude profile
        p = g['windprof']
        core.set_wind(float(g['windpt'][0]), float(g['windpt'][1]), p[:, 1], p[:, 2], alt=p[:, 0], augment_obs=True)
    return core


def ref_obs_of(g, obs):
    """The reference's obs with the wind slots of terminated rows taken from ours: on termination the reference
    deletes aircraft before the wrapper looks up 'kl001' (id2idx -> -1 -> the last remaining aircraft)."""
    ref = g['obs1'].copy()
    if 'wind' in g or 'windprof' in g:
        t = g['term'] == 1
        ref[t, -2:] = obs[t, -2:]
    return ref


@pytest.mark.parametrize('name', EXT_GOLDEN)
def test_fp64_ext_replays_reference_goldens(orc, name):
    g = load_golden(name)
    kind = name[0]
    na = 2 if kind in 'SM' else 1
    core = golden_core(name, g, 'fp64')
    assert core.obs_dim == g['obs1'].shape[1]
    nac = g['nac'].astype(np.int32)
    obs, rew, term, trunc, info, ac1, reso1, ev1, cd = run_step(core, g['ac0'], g['reso0'], nac, g['ev0'], g['act'][:, :na])
    tol = RT64 * 200
    assert relerr(obs, ref_obs_of(g, obs)) < tol, relerr(obs, ref_obs_of(g, obs))
    assert relerr(rew, g['rew']) < tol
    assert np.array_equal(term, g['term'])
    if kind != 'M':        # MergeEnv's recorded truncations are the TimeLimit wrapper's
        assert np.array_equal(trunc, g['trunc'])
    live = g['term'] == 0
    idx = [orc.F[k] for k in STATE_FIELDS]
    assert relerr(ac1[live][:, idx], g['ac1'][live][:, idx]) < tol
    assert np.array_equal(ac1[live][:, orc.F['lnav']], g['ac1'][live][:, orc.F['lnav']])
    assert np.array_equal(reso1[live], g['reso1'][live])
    assert relerr(info[:, 0], g['info'][:, 0]) < tol and np.array_equal(info[:, 1], g['info'][:, 1])
    if kind == 'M':
        assert np.array_equal(info[:, 3], g['info'][:, 3])                # faf_reach
        assert ((g['ac0'][:, orc.F['lnav']] == 0) & (ac1[:, orc.F['lnav']] == 1)).any()   # a waypoint switch on the GPU
    core.close()


@pytest.mark.parametrize('name', EXT_GOLDEN)
def test_fp32_ext_replays_reference_goldens(orc, name):
    g = load_golden(name)
    kind = name[0]
    na = 2 if kind in 'SM' else 1
    T = g['ac0'].shape[0]
    core = golden_core(name, g, 'fp32')
    nac = g['nac'].astype(np.int32)
    obs, rew, term, trunc, info, ac1, reso1, ev1, cd = run_step(core, g['ac0'], g['reso0'], nac, g['ev0'], g['act'][:, :na])
    ref = ref_obs_of(g, obs)
    tol = RT32 * np.maximum(np.abs(ref), 1.0)
    if kind == 'S':
        tol = obs_tol32(kind, ref)
    if kind == 'M':        # cos/sin(track): direction of a velocity difference (merge_env.py:213-215), as in Sector
        vrel = np.hypot(ref[:, 15:20] * 150, ref[:, 20:25] * 150)
        tol[:, 25:30] *= np.maximum(1.0, 600.0 / vrel); tol[:, 30:35] *= np.maximum(1.0, 600.0 / vrel)
    # thresholds crossed within fp32 rounding are ENUMERATED, not counted: a reward / termination / truncation flip must
    # sit within EPS32 of an env-layer threshold in the ORACLE's post-step state (intrusion, waypoint, runway, polygon),
    # or -- with MVP on -- belong to a step whose resolution decisions the oracle's trace places at a threshold /
    # ill-conditioned pair; an ASAS-activity flip must be such a step; LNAV switches (a turn-radius test on the route) are
    # listed and bounded by count
    from test_oracle_golden import golden_cfg
    from parity_helpers import env_threshold_margin
    from test_gpu_parity import EPS32, KAPPA0
    cfg = golden_cfg(orc, name, g)
    o_ac, o_reso, o_ev = g['ac0'].copy(), g['reso0'].copy(), g['ev0'].copy()
    tr = orc.step_trace(cfg, o_ac, o_reso, nac, o_ev, np.ascontiguousarray(g['act'][:, :na]), EPS32, nthreads=4)[5]
    near_env = env_threshold_margin(KIND[kind], o_ac, o_ev, nac, orc.F, orc.EV) < EPS32
    kap = np.nan_to_num(tr['kappa'], nan=0.0, posinf=1e30).max(1)
    near_res = ((tr['res_margin'] < EPS32).any(1) | (kap >= KAPPA0)) if 'asas' in name else np.zeros(T, bool)
    lnav_flip = (ac1[:, orc.F['lnav']] != g['ac1'][:, orc.F['lnav']]).any(1) & (g['term'] == 0)
    act_flip = (ac1[:, orc.F['asas_active']] != g['ac1'][:, orc.F['asas_active']]).any(1) & (g['term'] == 0)
    hard = (term != g['term']) | (np.abs(rew - g['rew']) > 0.5)
    if kind != 'M':
        hard |= trunc != g['trunc']
    print(f"[{name}] fp32 flips enumerated: reward / termination {np.where(hard)[0].tolist()}, ASAS activity {np.where(act_flip)[0].tolist()}, "
          f"LNAV switch {np.where(lnav_flip)[0].tolist()}; steps at an env threshold {int(near_env.sum())}, at a resolution threshold {int(near_res.sum())} of {T}")
    assert not (hard & ~near_env & ~near_res).any(), np.where(hard & ~near_env & ~near_res)[0]
    assert not (act_flip & ~near_res).any(), np.where(act_flip & ~near_res)[0]
    assert lnav_flip.sum() <= max(2, T // 25)
    flip = hard | lnav_flip | act_flip
    err = np.abs(obs - ref)
    if 'asas' in name:     # MVP conditioning, as in test_gpu_parity.test_batched_step_vs_oracle: the bound per step is
        # 1e-5 max(|ref|, 1) + 8 eps_f32 kappa / vnorm; 99 % of the entries sit inside the plain bound
        keep = ~flip & ~near_res
        wid = tol + ((8 * 2.0 ** -24 / 32.0) * kap)[:, None]
        assert not (err > wid)[keep].any(), np.argwhere((err > wid) & keep[:, None])[:5]
        assert np.quantile((err / tol)[keep], 0.99) <= 1.0
    else:
        bad = err > tol
        assert bad[~flip].sum() == 0, (np.argwhere(bad)[:5], relerr(obs[~flip], ref[~flip]))
    assert relerr(rew[~flip], g['rew'][~flip]) < 5e-5
    live = (g['term'] == 0) & ~flip
    for k in ['lat', 'lon', 'alt', 'tas', 'vs', 'gsn', 'gse']:
        floor = 150.0 if k in ('gsn', 'gse') else 1.0      # components of a ~150 m/s vector
        assert relerr(ac1[live][:, orc.F[k]], g['ac1'][live][:, orc.F[k]], floor=floor) < RT32 * (2 if 'asas' in name else 1), k
    core.close()


def test_merge_batched_vs_oracle(orc):
    """512 seeded Merge envs x 20 aircraft: chained fp64 steps == oracle (LNAV switches included), one fp32 step within
    tolerance; with and without a wind field."""
    from bluesky_gym_hallucination_b200 import scenario as sc
    from bluesky_gym_hallucination_b200.core import BsgCore
    E, N = 512, 20
    for wind in (None, ([53.0, 55.5], [4.0, 5.0], [15.0, -10.0], [-12.0, 8.0])):
        ac, reso, nac, ev = sc.merge(E, N, seed=21)
        # in every third env, six intruders start close to the FAF so that waypoint switches happen within a few steps
        near = np.arange(E) % 3 == 0
        lat, lon = sc.point_at_distance(sc.FAF[0], sc.FAF[1], np.random.default_rng(1).uniform(1.0, 12.0, (E, 6)),
                                        ac[:, orc.F['hdg'], 1:7] + 180.0)
        ac[near, orc.F['lat'], 1:7] = lat[near]; ac[near, orc.F['lon'], 1:7] = lon[near]
        cfg = orc.make_cfg(orc.KIND_M, N, wind=wind, wind_obs=int(wind is not None))
        c64 = BsgCore(4, E, N, precision='fp64'); c32 = BsgCore(4, E, N, precision='fp32', cd_outputs=True)
        if wind:
            c64.set_wind(*wind, augment_obs=True); c32.set_wind(*wind, augment_obs=True)
        rng = np.random.default_rng(3)
        o_ac, o_reso, o_ev = ac.copy(), reso.copy(), ev.copy()
        c64.set_state(ac, reso, nac, ev)
        switched = 0
        for s in range(4):
            act = rng.uniform(-1, 1, (E, 2))
            pre = (o_ac.copy(), o_reso.copy(), o_ev.copy())
            o_obs, o_rew, o_term, o_trunc, _ = orc.step(cfg, o_ac, o_reso, nac, o_ev, act, nthreads=8)
            switched += int((pre[0][:, orc.F['lnav']] != o_ac[:, orc.F['lnav']]).sum())
            c64.step(torch.as_tensor(act, dtype=torch.float64, device=c64.device))
            torch.cuda.synchronize()
            assert relerr(c64.obs.cpu().numpy(), o_obs) < RT64 * 1000
            assert relerr(c64.reward.cpu().numpy(), o_rew) < RT64 * 1000
            assert np.array_equal(c64.terminated.cpu().numpy(), o_term)
            g_ac = c64.get_state()[0]
            assert np.array_equal(g_ac[:, orc.F['lnav']], o_ac[:, orc.F['lnav']])
            obs, rew, term, trunc, info, ac1, _, _, _ = run_step(c32, pre[0], pre[1], nac, pre[2], act)
            flip = (term != o_term) | (np.abs(rew - o_rew) > 0.5) | (ac1[:, orc.F['lnav']] != o_ac[:, orc.F['lnav']]).any(1)
            assert flip.sum() <= E // 50, flip.sum()
            tol = RT32 * np.maximum(np.abs(o_obs), 1.0)
            vrel = np.hypot(o_obs[:, 15:20] * 150, o_obs[:, 20:25] * 150)
            tol[:, 25:30] *= np.maximum(1.0, 600.0 / vrel); tol[:, 30:35] *= np.maximum(1.0, 600.0 / vrel)
            # LNAV steers by the BEARING to the active waypoint: an aircraft passing at distance d carries the fp32
            # position difference of the step (~1e-5 of 800 m) as a heading difference ~ 1e-2 m / d, so the velocity
            # slots of an env scale with 20 km / (closest waypoint passage of the step, before or after it)
            dmin = np.full(E, 1e9)
            for st in (pre[0], o_ac):
                ln = st[:, orc.F['lnav']]
                wl = np.where(ln == 0, sc.FAF[0], sc.RWY[0]); wo = np.where(ln == 0, sc.FAF[1], sc.RWY[1])
                dkm = 111.2 * np.hypot(st[:, orc.F['lat']] - wl, (st[:, orc.F['lon']] - wo) * np.cos(np.radians(wl)))
                dmin = np.minimum(dmin, np.where(ln >= 0, dkm, 1e9).min(1))
            dmin[(pre[0][:, orc.F['lnav']] != o_ac[:, orc.F['lnav']]).any(1)] = 0.2      # passed over it within the step
            tol[:, 15:35] *= np.maximum(1.0, 20.0 / np.maximum(dmin, 0.2))[:, None]
            bad = np.abs(obs - o_obs) > 2 * tol
            assert bad[~flip].sum() == 0, (np.argwhere(bad)[:5], relerr(obs[~flip], o_obs[~flip]))
        assert switched > 50, switched
        c64.close(); c32.close()


def test_merge_device_reset_value_laws(orc):
    """bsg_reset for the Merge env: the reference's spawn laws (merge_env.py:110-115,148-156)."""
    from bluesky_gym_hallucination_b200.core import BsgCore
    from bluesky_gym_hallucination_b200 import scenario as sc
    E, N = 2048, 20
    c = BsgCore(4, E, N, precision='fp64')
    c.reset(seed=5)
    ac, reso, nac, ev = c.get_state()
    F = orc.F
    assert (nac == N).all()
    assert np.allclose(ev[:, orc.EV['wpt_lat']], sc.FAF[0], atol=1e-12) and np.allclose(ev[:, orc.EV['wpt_x']], sc.RWY[0])
    brg = ac[:, F['hdg']] + 180.0
    assert brg.min() >= -15.0 and brg.max() <= 15.0 and abs(brg.mean()) < 0.3
    # distance from the FAF along the radial: invert get_point_at_distance with the haversine on the 6371 km sphere
    la, lo = np.radians(ac[:, F['lat']]), np.radians(ac[:, F['lon']])
    fa, fo = np.radians(sc.FAF[0]), np.radians(sc.FAF[1])
    d = 2 * 6371.0 * np.arcsin(np.sqrt(np.sin((la - fa) / 2) ** 2 + np.cos(la) * np.cos(fa) * np.sin((lo - fo) / 2) ** 2))
    assert d[:, 0].min() >= 50 - 1e-6 and d[:, 0].max() <= 200 + 1e-6 and abs(d[:, 0].mean() - 125) < 4
    assert d[:, 1:].min() >= 20 - 1e-6 and d[:, 1:].max() <= 500 + 1e-6 and abs(d[:, 1:].mean() - 260) < 3
    lat2, lon2 = sc.point_at_distance(sc.FAF[0], sc.FAF[1], d, brg)
    assert np.abs(lat2 - ac[:, F['lat']]).max() < 1e-9 and np.abs(lon2 - ac[:, F['lon']]).max() < 1e-9
    assert (ac[:, F['lnav'], 0] == -1).all() and (ac[:, F['lnav'], 1:] == 0).all()
    assert np.allclose(ac[:, F['alt']], 10000.0) and np.allclose(ac[:, F['cas']], 100.0)
    assert np.allclose(ac[:, F['tas']], sc.vcas2tas(100.0, 10000.0), rtol=1e-12)
    assert relerr(np.hypot(ac[:, F['gsn']], ac[:, F['gse']]), ac[:, F['tas']]) < 1e-12
    # a reset + observe equals the oracle's observation of the exported state
    cfg = orc.make_cfg(orc.KIND_M, N)
    assert relerr(c.observe().cpu().numpy(), orc.observe(cfg, ac, nac, ev)) < RT64 * 10
    c.close()


def test_obs_noise_statistics_and_determinism():
    """NoisyObservationWrapper on device: obs + N(0, sigma) on every slot; sigma = 0 is the identity; the stream is
    a pure function of (seed, env, launch, slot)."""
    from bluesky_gym_hallucination_b200 import scenario as sc
    from bluesky_gym_hallucination_b200.core import BsgCore
    E, N, sigma = 4096, 20, 0.1
    ac, reso, nac, ev = sc.horizontal(E, N, seed=3)
    act = np.random.default_rng(0).uniform(-1, 1, (E, 1))
    outs = []
    for sg, seed in ((0.0, 1), (sigma, 1), (sigma, 1), (sigma, 2)):
        c = BsgCore(0, E, N, precision='fp32', asas_on=True)
        c.set_obs_noise(sg, seed=seed)
        c.set_state(ac, reso, nac, ev)
        c.step(torch.as_tensor(act, dtype=c.rdtype, device=c.device))
        torch.cuda.synchronize()
        outs.append((c.obs.cpu().numpy().astype(np.float64), c.reward.cpu().numpy()))
        c.close()
    clean, n1, n1b, n2 = outs
    assert np.array_equal(n1[0], n1b[0]) and not np.array_equal(n1[0], n2[0])
    assert np.array_equal(clean[1], n1[1])                       # the reward is computed from the true state
    d = n1[0] - clean[0]
    assert abs(d.mean()) < 5e-4 and abs(d.std() - sigma) < 5e-4
    assert np.abs(d.mean(0)).max() < 0.01 and np.abs(d.std(0) - sigma).max() < 0.01          # every slot, across envs
    assert abs(np.corrcoef(d[:, 0], d[:, 1])[0, 1]) < 0.06 and abs(np.corrcoef(d[0], d[1])[0, 1]) < 0.3
    k = ((d / sigma) ** 4).mean()
    assert abs(k - 3.0) < 0.1, k                                  # Gaussian kurtosis


def test_merge_and_wrappers_public_api():
    """gym.make('MergeEnv-v0') and the reference's wrapper signatures in front of the CUDA core."""
    import bluesky_gym_hallucination_b200 as bluesky_gym
    from bluesky_gym_hallucination_b200.wrappers import WindFieldWrapper, NoisyObservationWrapper
    bluesky_gym.register_envs()
    env = bluesky_gym.make('MergeEnv-v0', render_mode=None)
    obs, info = env.reset(seed=4)
    assert list(obs) == ['cos(drift)', 'sin(drift)', 'airspeed', 'waypoint_dist', 'faf_reached', 'x_r', 'y_r', 'vx_r',
                         'vy_r', 'cos(track)', 'sin(track)', 'distances']
    assert obs['x_r'].shape == (5,) and obs['faf_reached'][0] == 0 and set(info) == {'total_reward', 'faf_reach', 'average_drift', 'total_intrusions'}
    steps = 0
    while True:
        obs, r, term, trunc, info = env.step(np.zeros(2))
        steps += 1
        assert all(v.dtype == np.float64 for v in obs.values())
        if term or trunc:
            break
    assert steps <= 50 and (term or steps == 50)
    env.close()
    env = WindFieldWrapper(bluesky_gym.make('SectorCREnv-v0', render_mode=None), 52.0, 4.0, 20.0, 0.0, augment_obs=True)
    assert 'wind_u' in env.observation_space.spaces
    obs, info = env.reset(seed=1)
    obs, r, term, trunc, info = env.step(np.zeros(2))
    assert obs['wind_u'].shape == (1,) and abs(float(np.hypot(obs['wind_u'][0], obs['wind_v'][0])) - 20.0 / 50) < 1e-9
    env.close()
    env = NoisyObservationWrapper(bluesky_gym.make('HorizontalCREnv-v0', render_mode=None), noise_level=0.5, seed=3)
    o1, _ = env.reset(seed=2)
    base = bluesky_gym.make('HorizontalCREnv-v0', render_mode=None)
    o0, _ = base.reset(seed=2)
    dd = np.concatenate([(o1[k] - o0[k]).ravel() for k in o0])
    assert 0.2 < dd.std() < 1.0
    env.close(); base.close()
    v = bluesky_gym.BlueSkyVecEnv('MergeEnv-v0', 64, wind=(52.0, 4.0, 5.0, 5.0), wind_obs=True, obs_noise=0.01)
    obs, info = v.reset(seed=0)
    obs, rew, term, trunc, info = v.step(np.zeros((64, 2)))
    assert obs['wind_u'].shape == (64, 1) and info['faf_reach'].shape == (64,) and obs['distances'].shape == (64, 5)
    v.close()
