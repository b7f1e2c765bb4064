"""CPU: the C oracle (oracle/c) against the golden fixtures recorded from the UNMODIFIED
reference env classes run on oracle/refshim (tests/golden/make_golden.py).  One reference
step is replayed from every recorded pre-state; obs / reward / terminated / truncated / info
and the full post-state must agree to 1e-11 relative."""
import sys
import os
import numpy as np
import pytest
from conftest import load_golden, ROOT

KIND = {'H': 0, 'V': 1, 'S': 2, 'M': 4}
TOL = 1e-11


def golden_cfg(orc, name, g):
    """Oracle config of a golden file: env kind from the first letter, ASAS+MVP / wind from the tags."""
    kind = name[0]
    asas = int('asas' in name or 'mvp' in name)
    mode = 2 if 'mvphv' in name else (1 if 'mvpv' in name else 0)     # MVP resolution method (RMETHH / RMETHV)
    wind = tuple(g['wind'].T) if 'wind' in g else None
    if 'windprof' in g:      # one point with an altitude profile: (lat, lon, vnorth[], veast[], alt[])
        wind = (g['windpt'][0], g['windpt'][1], g['windprof'][:, 1], g['windprof'][:, 2], g['windprof'][:, 0])
    return orc.make_cfg(KIND[kind], g['ac0'].shape[2], asas_on=asas, reso_on=asas, reso_off_own=asas,
                        wind=wind, wind_obs=int(wind is not None), reso_mode=mode)


def rel(a, b):
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1.0))) if a.size else 0.0


def vertical_degenerate_steps(orc, cfg, g, na, name, eps=1e-9):
    """MVP with the vertical term (RMETHV / combined): `abs(vrel_z) > 0.0` is an EXACT-zero switch.  Two aircraft that
    resolve against each other symmetrically hold vertical speeds that agree to the last bit -- or differ by one ulp --
    depending on how tLOS(i,j) and tLOS(j,i) round (numpy's SIMD sin/cos vs libm vs CUDA), and the resolution flips
    sign on that bit.  Steps in which the oracle's trace puts a resolving aircraft within eps of such a switch are
    ENUMERATED and left out of the exact comparison; returns the mask of the steps that are compared."""
    nac = g['nac'].astype(np.int32)
    tr = orc.step_trace(cfg, g['ac0'].copy(), g['reso0'].copy(), nac, g['ev0'].copy(),
                        np.ascontiguousarray(g['act'][:, :na]), eps, nthreads=2)[5]
    deg = (tr['res_margin'] < eps).any(1)
    print(f"[{name}] steps at an exact-zero switch of the vertical term: {np.where(deg)[0].tolist()} ({int(deg.sum())} of {len(deg)})")
    assert deg.mean() < 0.6, deg.mean()      # (Sector: level, co-altitude traffic -- every conflict starts symmetric)
    return ~deg


GOLDEN = ['H', 'V', 'S', 'H_asas', 'S_asas', 'M', 'M_wind', 'S_wind', 'V_wind', 'H_wind', 'S_asas_wind',
          'V_mvpv', 'V_mvphv', 'H_mvphv', 'S_mvpv', 'V_windalt', 'S_asas_windalt']


@pytest.mark.parametrize('name', GOLDEN)
def test_c_oracle_replays_reference_steps(orc, name):
    g = load_golden(name)
    kind = name[0]
    asas = int('asas' in name or 'mvp' in name)
    n = g['ac0'].shape[2]
    cfg = golden_cfg(orc, name, g)
    nac = g['nac'].astype(np.int32)
    na = 2 if kind in 'SM' else 1
    obs0 = orc.observe(cfg, g['ac0'].copy(), nac, g['ev0'].copy())
    keep = np.ones(len(nac), bool)
    if kind == 'M':   # the reference observes BEFORE _check_waypoint flips wpt_reach (merge_env.py:139-140): the obs
        keep = g['obs0'][:, 4] == g['ev0'][:, orc.EV['wpt_reach']]     # a step returned lags the flag by one step
        assert (~keep).sum() >= 1
    assert rel(obs0[keep], g['obs0'][keep]) < TOL
    ac, reso, ev = g['ac0'].copy(), g['reso0'].copy(), g['ev0'].copy()
    obs, rew, term, trunc, cd = orc.step(cfg, ac, reso, nac, ev, np.ascontiguousarray(g['act'][:, :na]), nthreads=2)
    ref_obs = g['obs1'].copy()
    sel = np.ones(len(nac), bool)
    if 'mvp' in name:
        sel = vertical_degenerate_steps(orc, cfg, g, na, name)
        obs, ref_obs, rew, term, trunc, ac, reso, ev = (x[sel] for x in (obs, ref_obs, rew, term, trunc, ac, reso, ev))
        g = {k: (v[sel] if getattr(v, 'shape', ()) and v.shape[0] == len(sel) else v) for k, v in g.items()}
    if 'wind' in name:   # on termination the reference deletes aircraft before the wrapper looks up 'kl001'
        t = g['term'] == 1   # (id2idx -> -1 -> the LAST remaining aircraft): wind slots of that row are not KL001's
        ref_obs[t, -2:] = obs[t, -2:]
    assert rel(obs, ref_obs) < TOL
    assert rel(rew, g['rew']) < TOL
    assert np.array_equal(term, g['term'])
    if kind == 'M':   # MergeEnv never truncates itself; the recorded flags are the TimeLimit wrapper's (50 steps)
        assert not trunc.any() and np.all(g['ev0'][g['trunc'] == 1, orc.EV['steps']] == 0)
    else:
        assert np.array_equal(trunc, g['trunc'])
    live = g['term'] == 0
    assert rel(ac[live], g['ac1'][live]) < TOL
    assert np.array_equal(reso[live], g['reso1'][live])
    assert rel(ev[:, orc.EV['total_reward']], g['info'][:, 0]) < TOL
    assert np.array_equal(ev[:, orc.EV['total_intr']], g['info'][:, 1])
    if kind in 'HSM':   # average_drift = mean of the appended drifts (nan right after reset)
        avg = ev[:, orc.EV['drift_sum']] / ev[:, orc.EV['drift_cnt']]
        assert rel(avg, g['info'][:, 2]) < 1e-9
    else:
        assert rel(ev[:, orc.EV['final_alt']], g['info'][:, 2]) < TOL
    if kind == 'M':     # info['faf_reach']; both waypoint branches and the LNAV switch must occur
        assert np.array_equal(ev[:, orc.EV['final_alt']], g['info'][:, 3])
        assert g['term'].sum() >= 1 and (g['info'][:, 3] == 1).any()
        lnav0, lnav1 = g['ac0'][:, orc.F['lnav']], g['ac1'][:, orc.F['lnav']]
        assert ((lnav0 == 0) & (lnav1 == 1)).any()
    if 'wind' in name:  # the wind must actually bend the ground speed (Horizontal flies at alt 0: below 50 ft)
        gs = np.hypot(g['ac1'][:, orc.F['gsn']], g['ac1'][:, orc.F['gse']])
        moved = np.abs(gs - g['ac1'][:, orc.F['tas']])[live].max()
        assert (moved < 1e-9) if kind == 'H' else (moved > 1.0)
    # every golden file must exercise its branches
    if name == 'H':
        assert g['term'].sum() >= 1
    if name == 'S':
        assert g['trunc'].sum() >= 1
    if asas:
        assert (g['ac1'][:, orc.F['asas_active']].sum(1) > 0).sum() > 10
    if 'mvp' in name:   # the vertical term must act: resolving aircraft are commanded off their vertical speed / altitude
        act = (g['ac1'][:, orc.F['asas_active']] > 0) & live[:, None]
        dvs = np.abs(g['ac1'][:, orc.F['asas_vs']] - g['ac1'][:, orc.F['vs']])[act]
        dal = np.abs(g['ac1'][:, orc.F['asas_alt']] - g['ac1'][:, orc.F['selalt']])[act]
        assert (dvs > 0.5).sum() > 5 and (dal > 50.0).sum() > 5, ((dvs > 0.5).sum(), (dal > 50.0).sum())


def test_functions_kat():
    """Closed-form helper KAT observed in the survey: get_point_at_distance(52,4,100,0)."""
    g = load_golden('functions')
    assert abs(g['gpad'][0] - 52.89932160591875) < 1e-12 and g['gpad'][1] == 4.0
    assert list(g['bound']) == [181.0, 160.0, 10.0, -179.0]


def test_numpy_detect_matches_c_dense(orc):
    """oracle/refshim's numpy StateBased.detect (A.5 line by line) == the C pair loop."""
    sys.path.insert(0, os.path.join(ROOT, 'oracle', 'refshim'))
    from bluesky.traffic.asas import detect
    rng = np.random.default_rng(5)
    n = 40
    lat = 52 + rng.uniform(-0.6, 0.6, n); lon = 4 + rng.uniform(-1, 1, n)
    alt = rng.choice([3000.0, 3300.0, 3600.0], n) + rng.uniform(-50, 50, n)
    trk = rng.uniform(0, 360, n); gs = rng.uniform(120, 250, n); vs = rng.choice([-8.0, 0.0, 8.0], n)
    rpz, hpz = 5 * 1852.0, 1000 * 0.3048
    r = detect(lat, lon, alt, trk, gs, vs, np.full(n, rpz), np.full(n, hpz), np.full(n, 300.0))
    d = orc.detect_dense(lat, lon, alt, trk, gs, vs, rpz, hpz, 300.0)
    off = ~np.eye(n, dtype=bool)
    assert r['swconfl'].sum() > 5
    assert np.array_equal(r['swconfl'], d['swconfl'].astype(bool))
    assert np.array_equal(r['swlos'][off], d['swlos'].astype(bool)[off])
    assert rel(d['tcpa'][off], r['tcpa'][off]) < 1e-12
    assert rel(d['dcpa2'][off], r['dcpa2'][off]) < 1e-10
    c = r['swconfl']
    assert rel(d['tinconf'][c], r['tinconf'][c]) < 1e-12


def test_creconfs_known_answer(orc):
    """Analytic KAT: an intruder built by creconfs(dpsi, dcpa, tlosh) must be detected with
    dcpa == cpa*1852 m and tinconf == tlosh on the first CD call."""
    sys.path.insert(0, os.path.join(ROOT, 'oracle', 'refshim'))
    import bluesky as bs
    bs.init()
    bs.traf.cre('KL001', actype='A320', acalt=3000, acspd=150)
    cases = [(90, 2, 200), (200, 0, 150), (300, 4, 280), (60, 3, 120)]
    for k, (dpsi, cpa, tlosh) in enumerate(cases):
        bs.traf.creconfs(acid=f'I{k}', actype='A320', targetidx=0, dpsi=dpsi, dcpa=cpa, tlosh=tlosh)
    t = bs.traf
    d = orc.detect_dense(t.lat, t.lon, t.alt, t.trk, t.gs, t.vs)
    for k, (dpsi, cpa, tlosh) in enumerate(cases):
        j = k + 1
        assert d['swconfl'][0, j] == 1
        # kwikpos scales dlon by cos(lat_ref), kwikqdrdist by cos(mean lat): ~0.3 % of the offset
        assert abs(np.sqrt(d['dcpa2'][0, j]) - cpa * 1852.0) < 300.0
        assert abs(d['tinconf'][0, j] - tlosh) < 3.0


def test_np_port_matches_reference():
    """oracle/envs_np.HorizontalNp (the Python port bench.py's reference arm times on the GPU box, where
    /root/reference does not exist) replays the unmodified reference's recorded episodes exactly: same global
    np.random draw order, same obs / reward / terminated."""
    from oracle import envs_np
    g = load_golden('H')
    np.random.seed(11)                         # the seed tests/golden/make_golden.py used for golden_H
    env = envs_np.HorizontalNp(5, asas=False)
    t = 0
    for ep in range(4):
        obs = env.reset()
        assert rel(obs, g['obs0'][t]) < 1e-12
        for s in range(45):
            a = np.random.uniform(-1, 1, (1,))
            if ep % 2 == 0:
                a = a * 0.0
            obs, r, term, trunc, info = env.step(a)
            assert rel(obs, g['obs1'][t]) < 1e-12 and abs(r - g['rew'][t]) < 1e-12 and term == g['term'][t]
            assert abs(info['total_reward'] - g['info'][t, 0]) < 1e-9
            t += 1
            if term:
                break
    assert t == len(g['rew'])


def test_np_vertical_port_matches_reference():
    """oracle/envs_np.VerticalNp (BASELINE.json configs[0] on the GPU box: one VerticalCREnv, random actions, 1000 steps on
    the CPU path) replays the episodes recorded from the unmodified reference class exactly."""
    from oracle import envs_np
    g = load_golden('V')
    np.random.seed(12)                         # tests/golden/make_golden.py: ('V', 12, 4, 50, ...)
    env = envs_np.VerticalNp()
    t = 0
    for ep in range(4):
        obs = env.reset()
        assert rel(obs, g['obs0'][t]) < 1e-12
        for s in range(50):
            a = np.random.uniform(-1, 1, (1,))
            obs, r, term, trunc, info = env.step(a)
            assert rel(obs, g['obs1'][t]) < 1e-12 and abs(r - g['rew'][t]) < 1e-9 and term == g['term'][t]
            assert abs(info['total_reward'] - g['info'][t, 0]) < 1e-9 and info['total_intrusions'] == g['info'][t, 1]
            t += 1
            if term:
                break
    assert t == len(g['rew'])
    v, eps, dt = envs_np.vertical_1000_steps(seed=1, steps=60)
    assert v > 0 and eps >= 1
    r = envs_np.detect_rates(sizes=(20,), budget_s=0.2)
    assert r[20] > 0


def test_reference_arm_runs_the_unmodified_reference_class():
    """bench.py --impl reference: when tools/make_oracle_ref.sh has shipped the reference's own package (oracle/_ref,
    git-ignored), the CPU arm steps the UNMODIFIED HorizontalCREnv class at the workload's aircraft count.  Runs in a
    subprocess: the simulator stand-in is process-global."""
    import hashlib
    import subprocess
    import sys
    from conftest import ROOT
    from oracle import envs_np
    if not envs_np.ref_available():
        pytest.skip("oracle/_ref is absent (run tools/make_oracle_ref.sh where /root/reference exists)")
    ref = os.path.join(ROOT, 'oracle', '_ref')
    for line in open(os.path.join(ref, 'SHA256SUMS')):          # the copy is what the recipe hashed at the source
        digest, rel = line.split()
        assert hashlib.sha256(open(os.path.join(ref, rel), 'rb').read()).hexdigest() == digest, rel
    code = (
        "import sys, numpy as np; sys.path.insert(0, %r)\n"
        "from oracle import envs_np\n"
        "np.random.seed(1)\n"
        "env = envs_np.ref_horizontal(19, True)\n"
        "import bluesky_gym.envs.horizontal_cr_env as H\n"
        "assert type(env) is H.HorizontalCREnv and H.__file__.startswith(%r), H.__file__\n"
        "obs, info = env.reset()\n"
        "assert obs['intruder_distance'].shape == (19,)\n"
        "for _ in range(3):\n"
        "    obs, r, term, trunc, info = env.step(np.zeros(1))\n"
        "assert obs['x_difference_speed'].shape == (19,) and np.isfinite(r)\n"
        "import bluesky as bs\n"
        "assert bs.traf.ntraf == 20 and bs.traf.cd.inconf.shape == (20,)\n"
        "print('ok')\n") % (ROOT, ref)
    out = subprocess.run([sys.executable, '-c', code], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and out.stdout.strip().endswith('ok'), out.stderr[-2000:]
