"""Generate the golden fixtures in this directory.  Run HERE (needs /root/reference):

    python tests/golden/make_golden.py

It imports the UNMODIFIED reference env classes from /root/reference/bluesky_gym on top
of oracle/refshim (the CPU stand-in for bluesky-simulator==1.0.3, which is un-vendored and
not installable offline) and records seeded episodes: before every step the full traffic
state, then the action, and after the step the reference's own obs / reward / terminated /
truncated / info.  The fixtures pin Tier A (env layer) to the reference's source; Tier B
(BlueSky arithmetic) remains a restatement: parity unpinned vs upstream.
"""
import os
import random
import sys
import warnings
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, 'oracle', 'refshim'))
sys.path.insert(0, '/root/reference')
sys.path.insert(0, ROOT)
warnings.filterwarnings('ignore')

import bluesky as bs                     # noqa: E402  (oracle shim)
import gymnasium as gym                  # noqa: E402  (oracle shim)
import bluesky_gym                       # noqa: E402  (the reference, unmodified)
from oracle import oracle as orc         # noqa: E402

OBS_ORDER = {
    'H': ['intruder_distance', 'cos_difference_pos', 'sin_difference_pos', 'x_difference_speed',
          'y_difference_speed', 'waypoint_distance', 'cos_drift', 'sin_drift'],
    'V': ['altitude', 'vz', 'target_altitude', 'runway_distance', 'intruder_distance', 'cos_difference_pos',
          'sin_difference_pos', 'altitude_difference', 'x_difference_speed', 'y_difference_speed',
          'z_difference_speed'],
    'S': ['cos(drift)', 'sin(drift)', 'airspeed', 'x_r', 'y_r', 'vx_r', 'vy_r', 'cos(track)', 'sin(track)',
          'distances'],
    'M': ['cos(drift)', 'sin(drift)', 'airspeed', 'waypoint_dist', 'faf_reached', 'x_r', 'y_r', 'vx_r', 'vy_r',
          'cos(track)', 'sin(track)', 'distances'],
}
ENV_ID = {'H': 'HorizontalCREnv-v0', 'V': 'VerticalCREnv-v0', 'S': 'SectorCREnv-v0', 'M': 'MergeEnv-v0'}


def flat_obs(kind, obs):
    keys = OBS_ORDER[kind] + (['wind_u', 'wind_v'] if 'wind_u' in obs else [])
    return np.concatenate([np.asarray(obs[k], dtype=np.float64).ravel() for k in keys])


def snapshot(kind, env, n):
    """Traffic + env-level state in the oracle/CUDA data model (oracle/c/bsgo.h)."""
    t = bs.traf
    u = env.unwrapped
    nl = t.ntraf
    ac = np.zeros((orc.NF, n))
    src = dict(lat=t.lat, lon=t.lon, alt=t.alt, hdg=t.hdg, tas=t.tas, vs=t.vs, cas=t.cas, selspd=t.selspd,
               selalt=t.selalt, selvs=t.selvs, aptrk=t.ap.trk, ax=t.ax, asas_trk=t.cr.trk, asas_tas=t.cr.tas,
               asas_vs=t.cr.vs, asas_alt=t.cr.alt, asas_active=t.cr.active.astype(float),
               phase=t.perf.phase.astype(float), gsn=t.gsnorth, gse=t.gseast,
               lnav=np.array([float(r.iactwp) if on else -1.0 for r, on in zip(t.ap.route, t.swlnav)]))
    for k, v in src.items():
        ac[orc.F[k], :nl] = v
    reso = np.zeros(n, np.uint64)
    for (i, j) in t.cr.resopairs:
        reso[i] |= np.uint64(1) << np.uint64(j)
    ev = np.zeros(orc.NE)
    ev[orc.EV['total_reward']] = u.total_reward
    ev[orc.EV['total_intr']] = u.total_intrusions
    if kind == 'H':
        ev[orc.EV['wpt_lat']] = u.wpt_lat[0]; ev[orc.EV['wpt_lon']] = u.wpt_lon[0]
        ev[orc.EV['wpt_reach']] = u.wpt_reach[0]; ev[orc.EV['ac_hdg']] = u.ac_hdg
    if kind == 'V':
        ev[orc.EV['target_alt']] = u.target_alt; ev[orc.EV['final_alt']] = u.final_altitude
    if kind == 'S':
        ev[orc.EV['wpt_x']] = u.wpts[0][0]; ev[orc.EV['wpt_y']] = u.wpts[0][1]
        poly = bs.tools.areafilter._areas[u.poly_name]
        nv = len(poly.vlat)
        assert nv <= orc.MAXV
        ev[orc.EV['nvert']] = nv
        ev[orc.EV['poly']:orc.EV['poly'] + 2 * nv:2] = poly.vlat
        ev[orc.EV['poly'] + 1:orc.EV['poly'] + 2 * nv:2] = poly.vlon
    if kind == 'M':   # slot reuse documented in oracle/c/bsgo.h
        ev[orc.EV['wpt_lat']] = u.wpt_lat; ev[orc.EV['wpt_lon']] = u.wpt_lon
        ev[orc.EV['wpt_x']] = u.rwy_lat; ev[orc.EV['wpt_y']] = u.rwy_lon
        ev[orc.EV['wpt_reach']] = u.wpt_reach; ev[orc.EV['final_alt']] = u.faf_reached
    if kind in 'HSM':
        ev[orc.EV['drift_sum']] = float(np.sum(u.average_drift)); ev[orc.EV['drift_cnt']] = len(u.average_drift)
    return ac, reso, nl, ev


def record(kind, seed, episodes, max_steps, asas=False, n_pad=None, wind=None):
    """wind = (lat, lon, vnorth, veast): the env is wrapped in the reference's WindFieldWrapper (augment_obs)."""
    np.random.seed(seed)
    random.seed(seed)           # MergeEnv draws from Python's `random`
    env = gym.make(ENV_ID[kind], render_mode=None)
    if wind is not None:
        from bluesky_gym.wrappers.wind import WindFieldWrapper
        env = WindFieldWrapper(env, *wind, augment_obs=True)
    R = dict(ac0=[], reso0=[], nac=[], ev0=[], act=[], ac1=[], reso1=[], ev1=[], obs0=[], obs1=[], rew=[],
             term=[], trunc=[], info=[])
    if wind is not None and len(wind) == 5:      # one point with an altitude profile: WindFieldWrapper(..., alt=[...])
        R['windpt'] = np.array([wind[0], wind[1]], dtype=np.float64)
        R['windprof'] = np.stack([np.asarray(wind[4], dtype=np.float64), np.asarray(wind[2], dtype=np.float64),
                                  np.asarray(wind[3], dtype=np.float64)], axis=1)          # [knots][alt, vnorth, veast]
    elif wind is not None:
        wl = np.atleast_1d(np.asarray(wind[0], dtype=np.float64))
        R['wind'] = np.stack([wl] + [np.broadcast_to(np.atleast_1d(np.asarray(w, dtype=np.float64)), wl.shape)
                                     for w in wind[1:]], axis=1)
    for ep in range(episodes):
        obs, info = env.reset()
        if asas:   # the north-star feature on top of the reference: ASAS ON + MVP for the intruders
            bs.stack.stack('ASAS ON'); bs.stack.stack('RESO MVP'); bs.stack.stack('RESOOFF KL001')
            # resolution method (BlueSky RMETHH / RMETHV): True = the default (horizontal, heading + speed),
            # 'v' = vertical speed only, 'hv' = horizontal + vertical
            if asas == 'v':
                bs.stack.stack('RMETHV V/S')
            elif asas == 'hv':
                bs.stack.stack('RMETHH OFF'); bs.stack.stack('RMETHV OFF')
            else:
                bs.stack.stack('RMETHV OFF'); bs.stack.stack('RMETHH BOTH')
            bs.stack.process()
        if kind == 'M':   # flush the route commands reset() queued (ADDWPT / DEST / RESO OFF) so that the
            bs.stack.process()   # snapshot below already holds the LNAV state the first sim.step() would set
        n = n_pad or {'S': 32, 'M': 20}.get(kind, 6)
        for s in range(max_steps):
            ac0, reso0, nl, ev0 = snapshot(kind, env, n)
            a = np.random.uniform(-1, 1, env.action_space.shape)
            if kind == 'H' and ep % 2 == 0:
                a = a * 0.0           # straight flight: reaches the waypoint (termination branch)
            if kind == 'M':
                a = a * (0.15 if ep % 2 == 0 else 1.0)      # gentle steering: reaches the FAF, then the runway
            obs1, r, term, trunc, info = env.step(a)
            R['ac0'].append(ac0); R['reso0'].append(reso0); R['nac'].append(nl); R['ev0'].append(ev0)
            R['act'].append(np.resize(a, 2) if kind in 'SM' else np.array([a[0], 0.0]))
            R['obs0'].append(flat_obs(kind, obs)); R['obs1'].append(flat_obs(kind, obs1))
            R['rew'].append(r); R['term'].append(int(term)); R['trunc'].append(int(bool(trunc)))
            R['info'].append([info['total_reward'], info['total_intrusions'],
                              info.get('average_drift', info.get('final_altitude', 0.0)), info.get('faf_reach', 0.0)])
            if term:       # the reference deletes all aircraft on termination: no post-state
                R['ac1'].append(np.zeros_like(ac0)); R['reso1'].append(reso0 * 0); R['ev1'].append(ev0 * 0)
                break
            ac1, reso1, _, ev1 = snapshot(kind, env, n)
            R['ac1'].append(ac1); R['reso1'].append(reso1); R['ev1'].append(ev1)
            obs = obs1
            if trunc:
                break
    return {k: np.array(v) for k, v in R.items()}


def main():
    bluesky_gym.register_envs()
    W1 = (52.0, 4.0, 12.0, -7.0)                                        # one point: uniform wind
    WP = (52.0, 4.0, np.array([4.0, 12.0, -6.0, 20.0]), np.array([-3.0, 8.0, 15.0, -10.0]),
          np.array([0.0, 1000.0, 2500.0, 6000.0]))                        # one point, four knots [m]
    W3 = ([51.5, 52.6, 52.0], [3.5, 4.2, 5.5], [10.0, -4.0, 6.0], [-8.0, 12.0, 3.0])   # three points: weighted field
    specs = [('H', 11, 4, 45, False, None, ''), ('V', 12, 4, 50, False, None, ''), ('S', 13, 4, 120, False, None, ''),
             ('H', 21, 3, 30, True, None, '_asas'), ('S', 23, 3, 60, True, None, '_asas'),
             ('M', 31, 4, 50, False, None, ''), ('M', 32, 2, 50, False, W1, '_wind'),
             ('S', 33, 2, 60, False, W1, '_wind'), ('V', 34, 2, 50, False, W3, '_wind'),
             ('H', 35, 2, 30, False, W1, '_wind'), ('S', 36, 2, 40, True, W3, '_asas_wind'),
             # MVP resolution methods beyond the default: vertical only / horizontal + vertical
             ('V', 41, 4, 50, 'v', None, '_mvpv'), ('V', 42, 4, 50, 'hv', None, '_mvphv'),
             ('H', 43, 3, 30, 'hv', None, '_mvphv'), ('S', 44, 2, 50, 'v', None, '_mvpv'),
             # wind with an altitude profile (the wrapper's alt= argument): the Vertical env climbs / descends through it
             ('V', 45, 3, 50, False, WP, '_windalt'), ('S', 46, 2, 40, True, WP, '_asas_windalt')]
    only = sys.argv[1:]          # e.g. `make_golden.py V_mvpv H_mvphv` regenerates those fixtures alone
    for kind, seed, eps, ms, asas, wind, tag in specs:
        if only and f"{kind}{tag}" not in only:
            continue
        R = record(kind, seed, eps, ms, asas, wind=wind)
        name = f"golden_{kind}{tag}.npz"
        np.savez_compressed(os.path.join(HERE, name), **R)
        print(name, 'steps', len(R['rew']), 'terms', int(R['term'].sum()), 'truncs', int(R['trunc'].sum()),
              'asas-active steps', int((R['ac1'][:, orc.F['asas_active']].sum(1) > 0).sum()))
    if only:
        return
    # closed-form helper KATs from the reference's functions.py
    import bluesky_gym.envs.common.functions as fn
    kat = dict(gpad=np.array(fn.get_point_at_distance(52, 4, 100, 0)),
               hdg=np.array([fn.get_hdg(np.array([52., 4.]), np.array([52.5, 5.0]))]),
               ll2nm=fn.latlong_to_nm(np.array([51.990426702297746, 4.376124857109851]), np.array([52.3, 4.9])),
               bound=np.array([fn.bound_angle_positive_negative_180(x) for x in (541.0, -200.0, 10.0, 181.0)]))
    np.savez(os.path.join(HERE, 'golden_functions.npz'), **kat)
    print('functions KAT', kat['gpad'])


if __name__ == '__main__':
    main()
