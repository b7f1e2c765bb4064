"""ORACLE (test infrastructure, not product code).

Python/numpy port of the reference env layer on top of oracle/refshim's BlueSky stand-in, generalised to
K intruders -- the reference's own execution style (one process-global simulator per process, per-step
Python loops over intruders, numpy on tiny arrays).  /root/reference does not exist on the GPU box, so
this port is what bench.py's `--impl reference` arm and `cpu_baseline` time there.  It restates:
  HorizontalCREnv  /root/reference/bluesky_gym/envs/horizontal_cr_env.py:82-275
  VerticalCREnv    /root/reference/bluesky_gym/envs/vertical_cr_env.py:113-305   (BASELINE.json configs[0])
and is checked against the unmodified reference (tests/test_oracle_golden.py::test_np_port_matches_reference).
"""
import os
import sys
import numpy as np

_SHIM = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'refshim')


def _bs():
    if _SHIM not in sys.path:
        sys.path.insert(0, _SHIM)
    import bluesky
    return bluesky


def _wrap180(a):
    if a > 180:
        return -(360 - a)
    if a < -180:
        return 360 + a
    return a


class HorizontalNp:
    """HorizontalCREnv with `n_intruders` creconfs intruders; asas=True adds BlueSky's 'ASAS ON'."""

    def __init__(self, n_intruders=5, asas=False):
        self.K = n_intruders
        self.bs = _bs()
        self.bs.init(mode='sim', detached=True)
        self.bs.stack.stack('DT 5;FF')
        if asas:
            self.bs.stack.stack('ASAS ON')
        self.total_reward = 0.0
        self.total_intrusions = 0
        self.drifts = []

    def reset(self):
        bs = self.bs
        bs.traf.reset()
        self.total_reward = 0.0; self.total_intrusions = 0; self.drifts = []
        bs.traf.cre('KL001', actype="A320", acspd=150)
        for i in range(self.K):
            bs.traf.creconfs(acid=f'{i}', actype="A320", targetidx=0, dpsi=np.random.randint(45, 315),
                             dcpa=np.random.randint(0, 5), tlosh=np.random.randint(100, 1000))
        d = np.random.randint(100, 150) / 6371
        la, lo = np.radians(bs.traf.lat[0]), np.radians(bs.traf.lon[0])
        la2 = np.arcsin(np.sin(la) * np.cos(d) + np.cos(la) * np.sin(d) * np.cos(0.0))
        lo2 = lo + np.arctan2(np.sin(0.0) * np.sin(d) * np.cos(la), np.cos(d) - np.sin(la) * np.sin(la2))
        self.wpt = (np.degrees(la2), np.degrees(lo2))
        self.reached = 0
        return self._obs()

    def _obs(self):
        bs, t = self.bs, self.bs.traf
        dist, cb, sb, xd, yd = [], [], [], [], []
        self.ac_hdg = t.hdg[0]
        for i in range(self.K):
            j = i + 1
            qdr, dis = bs.tools.geo.kwikqdrdist(t.lat[0], t.lon[0], t.lat[j], t.lon[j])
            dist.append(dis * 1.852)
            brg = _wrap180(self.ac_hdg - qdr)
            cb.append(np.cos(np.deg2rad(brg))); sb.append(np.sin(np.deg2rad(brg)))
            hd = t.hdg[0] - t.hdg[j]
            xd.append(-np.cos(np.deg2rad(hd)) * t.gs[j])
            yd.append(t.gs[0] - np.sin(np.deg2rad(hd)) * t.gs[j])
        qdr, dis = bs.tools.geo.kwikqdrdist(t.lat[0], t.lon[0], self.wpt[0], self.wpt[1])
        self.wdist = dis * 1.852
        self.drift = _wrap180(self.ac_hdg - qdr)
        return np.concatenate([np.array(dist) / 150, cb, sb, np.array(xd) / 150, np.array(yd) / 150,
                               [self.wdist / 150, np.cos(np.deg2rad(self.drift)), np.sin(np.deg2rad(self.drift))]])

    def step(self, action):
        bs, t = self.bs, self.bs.traf
        bs.stack.stack(f"HDG KL001 {self.ac_hdg + action[0] * 45}")
        for _ in range(10):
            bs.sim.step()
        obs = self._obs()
        reach = 0
        if self.wdist < 5 and self.reached != 1:
            self.reached = 1; reach = 1
        dr = abs(np.deg2rad(self.drift))
        self.drifts.append(dr)
        intr = 0
        for i in range(self.K):
            _, dis = bs.tools.geo.kwikqdrdist(t.lat[0], t.lon[0], t.lat[i + 1], t.lon[i + 1])
            if dis < 5:
                self.total_intrusions += 1; intr += -1
        rew = reach + dr * -0.1 + intr
        self.total_reward += rew
        return obs, rew, int(self.reached == 1), False, {'total_reward': self.total_reward,
                                                         'total_intrusions': self.total_intrusions,
                                                         'average_drift': float(np.mean(self.drifts))}


class VerticalNp:
    """VerticalCREnv (vertical_cr_env.py:113-305): 5 creconfs intruders with altitude offsets, the agent commands a
    vertical speed; 30 sim steps of 1 s per action.  Same global np.random draw order as the reference."""
    K = 5

    def __init__(self):
        self.bs = _bs()
        self.bs.init(mode='sim', detached=True)
        self.bs.stack.stack('DT 1;FF')
        self.total_reward = 0.0; self.total_intrusions = 0; self.final_altitude = 0

    def reset(self):
        bs = self.bs
        bs.traf.reset()
        self.total_reward = 0.0; self.total_intrusions = 0; self.final_altitude = 0
        alt_init = np.random.randint(2000, 4000)
        self.target_alt = alt_init + np.random.randint(-500, 500)
        bs.traf.cre('KL001', actype="A320", acalt=alt_init, acspd=150)
        bs.traf.swvnav[0] = False
        altitude, spd = bs.traf.alt[0], bs.traf.gs[0]
        for i in range(self.K):                                     # _generate_conflicts, vertical_cr_env.py:182-199
            dpsi = np.random.randint(45, 315)
            cpa = np.random.randint(0, 5)
            tlosh = np.random.randint(100, int((200 * 0.9) * 1000 / spd))
            average_tod = (200 * 1000 / spd) - 2 * self.target_alt / 12.5
            if tlosh > average_tod:
                dH = np.random.randint(int(-altitude + 500), int((self.target_alt - altitude) + 100))
            else:
                dH = np.random.randint(int((self.target_alt - altitude) - 500), int((self.target_alt - altitude) + 500))
            bs.traf.creconfs(acid=f'{i}', actype="A320", targetidx=0, dpsi=dpsi, dcpa=cpa, tlosh=tlosh, dH=dH,
                             tlosv=100000000000.)
            bs.traf.alt[i + 1] = bs.traf.alt[0] + dH
            bs.traf.ap.selaltcmd(i + 1, bs.traf.alt[0] + dH, 0)
        return self._obs()

    def _obs(self):
        bs, t = self.bs, self.bs.traf
        dist, cb, sb, da, xd, yd, zd = [], [], [], [], [], [], []
        self.altitude, self.vz = t.alt[0], t.vs[0]
        for i in range(self.K):
            j = i + 1
            qdr, dis = bs.tools.geo.kwikqdrdist(t.lat[0], t.lon[0], t.lat[j], t.lon[j])
            dist.append(dis * 1.852)
            da.append(t.alt[j] - self.altitude); zd.append(t.vs[j] - self.vz)
            brg = _wrap180(t.hdg[0] - qdr)
            cb.append(np.cos(np.deg2rad(brg))); sb.append(np.sin(np.deg2rad(brg)))
            hd = t.hdg[0] - t.hdg[j]
            xd.append(-np.cos(np.deg2rad(hd)) * t.gs[j])
            yd.append(t.gs[0] - np.sin(np.deg2rad(hd)) * t.gs[j])
        self.runway_distance = 200 - bs.tools.geo.kwikdist(52, 4, t.lat[0], t.lon[0]) * 1.852
        return np.concatenate([[(self.altitude - 1500) / 3000], [self.vz / 5], [(self.target_alt - 1500) / 3000],
                               [(self.runway_distance - 100) / 200], np.array(dist) / 200, cb, sb, np.array(da) / 3000,
                               np.array(xd) / 150, np.array(yd) / 150, zd])

    def step(self, action):
        bs, t = self.bs, self.bs.traf
        a = action[0] * 12.5
        t.selalt[0] = 1000000 if a >= 0 else 0
        t.selvs[0] = a
        for _ in range(30):
            bs.sim.step()
        obs = self._obs()
        intr = 0
        for i in range(self.K):
            _, dis = bs.tools.geo.kwikqdrdist(t.lat[0], t.lon[0], t.lat[i + 1], t.lon[i + 1])
            if dis < 5 and abs(t.alt[0] - t.alt[i + 1]) < 1000 * 0.3048:
                self.total_intrusions += 1; intr += -50
        done = 0
        if self.runway_distance > 0 and self.altitude > 0:
            pen = abs(self.target_alt - self.altitude) * (-5 / 3000)
        elif self.altitude <= 0:
            pen = -100; self.final_altitude = -100; done = 1
        else:
            pen = self.altitude * (-50 / 3000); self.final_altitude = self.altitude; done = 1
        rew = pen + intr
        self.total_reward += rew
        return obs, rew, done, False, {'total_reward': self.total_reward, 'total_intrusions': self.total_intrusions,
                                       'final_altitude': self.final_altitude}


def vertical_1000_steps(seed=0, steps=1000):
    """BASELINE.json configs[0]: one VerticalCREnv, random actions, `steps` env steps on ONE core.
    Returns (env-steps/s, episodes, seconds)."""
    import time
    np.random.seed(seed)
    env = VerticalNp()
    env.reset()
    env.step(np.zeros(1))
    env.reset()
    t0 = time.perf_counter()
    eps = 0
    for s in range(steps):
        obs, r, term, trunc, info = env.step(np.random.uniform(-1, 1, 1))
        if term:
            env.reset(); eps += 1
    dt = time.perf_counter() - t0
    return steps / dt, eps, dt


def detect_rates(sizes=(20, 50, 500), budget_s=2.0):
    """The numpy state-based detection alone (refshim bluesky.traffic.asas.detect, all N x N pairs -- BlueSky's
    StateBased.detect): ordered pair-evals/s on one core for each traffic size."""
    import time
    _bs()
    from bluesky.traffic import asas
    out = {}
    for n in sizes:
        rng = np.random.default_rng(n)
        lat, lon = 52 + rng.uniform(-2, 2, n), 4 + rng.uniform(-3, 3, n)
        alt, trk = rng.uniform(3000, 11000, n), rng.uniform(0, 360, n)
        gs, vs = rng.uniform(120, 250, n), rng.choice([-5.0, 0.0, 5.0], n)
        args = (lat, lon, alt, trk, gs, vs, np.full(n, 5 * 1852.0), np.full(n, 1000 * 0.3048), np.full(n, 300.0))
        asas.detect(*args)
        t0 = time.perf_counter(); it = 0
        while time.perf_counter() - t0 < budget_s:
            asas.detect(*args); it += 1
        out[n] = it * n * (n - 1) / (time.perf_counter() - t0)
    return out


_REF = os.path.join(os.path.dirname(os.path.abspath(__file__)), '_ref')


def ref_available():
    """True when tools/make_oracle_ref.sh has put the reference's own package under oracle/_ref (git-ignored)."""
    return os.path.isfile(os.path.join(_REF, 'bluesky_gym', 'envs', 'horizontal_cr_env.py'))


def ref_horizontal(n_intruders, asas):
    """The UNMODIFIED reference class bluesky_gym.envs.horizontal_cr_env.HorizontalCREnv (oracle/_ref) on top of
    oracle/refshim, at the workload's aircraft count: NUM_INTRUDERS is a module constant the class reads at run time
    (horizontal_cr_env.py:17,51-55,129,167,263,319) and is set on the imported module, the source is untouched; 'ASAS ON'
    is the stack command the north-star adds (the reference leaves detection off)."""
    bs = _bs()
    for p in (_REF,):
        if p not in sys.path:
            sys.path.insert(0, p)
    import warnings
    warnings.filterwarnings('ignore', category=RuntimeWarning)      # the reference's mean of an empty drift list
    import bluesky_gym.envs.horizontal_cr_env as H
    H.NUM_INTRUDERS = int(n_intruders)
    env = H.HorizontalCREnv(render_mode=None)
    if asas:
        bs.stack.stack('ASAS ON')
    return env


def _worker(args):
    """One process = one simulator singleton (the reference's SubprocVecEnv scaling style,
    scripts/multi_processing_example.py:51-53): run `steps` env steps, return (steps, seconds)."""
    import time
    seed, n_intruders, asas, steps = args
    np.random.seed(seed)
    env = HorizontalNp(n_intruders, asas)
    env.reset()
    env.step(np.zeros(1))                         # warm caches / first-call overheads
    t0 = time.perf_counter()
    done_steps = 0
    for s in range(steps):
        obs, r, term, trunc, info = env.step(np.random.uniform(-1, 1, 1))
        done_steps += 1
        if term:
            env.reset()
    return done_steps, time.perf_counter() - t0


def timed_sample(n_intruders, asas, steps_per_proc, procs, seed=0):
    """Aggregate env-steps/s of `procs` independent single-env processes (numpy BlueSky-style path)."""
    import multiprocessing as mp
    import time
    ctx = mp.get_context('fork')
    t0 = time.perf_counter()
    with ctx.Pool(procs) as pool:
        res = pool.map(_worker, [(seed + p, n_intruders, asas, steps_per_proc) for p in range(procs)])
    wall = time.perf_counter() - t0
    total = sum(r[0] for r in res)
    busy = max(r[1] for r in res)
    return total / busy, total, wall


def _pool_worker(conn, seed, n_intruders, asas, use_ref=False):
    import time
    np.random.seed(seed)
    env = ref_horizontal(n_intruders, asas) if use_ref else HorizontalNp(n_intruders, asas)
    env.reset()
    env.step(np.zeros(1))
    conn.send('ready')
    while True:
        n = conn.recv()
        if n is None:
            break
        t0 = time.perf_counter()
        for _ in range(n):
            obs, r, term, trunc, info = env.step(np.random.uniform(-1, 1, 1))
            if term:
                env.reset()
        conn.send(time.perf_counter() - t0)
    conn.close()


class WorkerPool:
    """`procs` persistent single-env processes (one simulator singleton each: the reference's SubprocVecEnv scaling
    style, scripts/multi_processing_example.py:51-53).  step(n): every process advances its env by n env-steps; returns
    the wall-clock seconds of the slowest one including the pipe round trip."""

    def __init__(self, procs, n_intruders, asas, seed=0, use_ref=False):
        import multiprocessing as mp
        ctx = mp.get_context('fork')
        self.conns, self.procs = [], []
        for p in range(procs):
            a, b = ctx.Pipe()
            pr = ctx.Process(target=_pool_worker, args=(b, seed + p, n_intruders, asas, use_ref), daemon=True)
            pr.start()
            self.conns.append(a); self.procs.append(pr)
        for c in self.conns:
            c.recv()

    def step(self, n):
        import time
        t0 = time.perf_counter()
        for c in self.conns:
            c.send(n)
        for c in self.conns:
            c.recv()
        return time.perf_counter() - t0

    def close(self):
        for c in self.conns:
            c.send(None)
        for p in self.procs:
            p.join(timeout=5)
