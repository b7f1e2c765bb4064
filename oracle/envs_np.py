"""ORACLE (test infrastructure, not product code).

Python/numpy port of the reference env layer on top of oracle/refshim's BlueSky stand-in, generalised to
K intruders -- the reference's own execution style (one process-global simulator per process, per-step
Python loops over intruders, numpy on tiny arrays).  /root/reference does not exist on the GPU box, so
this port is what bench.py's `--impl reference` arm and `cpu_baseline` time there.  It restates:
  HorizontalCREnv  /root/reference/bluesky_gym/envs/horizontal_cr_env.py:82-275
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
