"""e2e breakdown of the host-buffer step (C2 shape): copies to user arrays vs pinned views."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from bluesky_gym_hallucination_b200 import core as bc
kind, E, N = 0, 4096, 20
c = bc.BsgCore(kind, E, N, precision='fp32', asas_on=True, max_episode_steps=300, fixed_nac=True)
c.reset(seed=1); c.set_autoreset(True, seed=2)
a = np.random.default_rng(0).uniform(-1, 1, (E, 1)).astype(np.float32)
obs = np.zeros((E, c.obs_dim), np.float32); rew = np.zeros(E, np.float32); term = np.zeros(E, np.int32)
trunc = np.zeros(E, np.int32); info = np.zeros((E, 4), np.float32)
def t(fn, n=300):
    for _ in range(20): fn()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    return (time.perf_counter() - t0) / n * 1e6
print("step_host (copies to user arrays): %.1f us" % t(lambda: c.step_host(a, obs, rew, term, trunc, info)))
print("step_host_views (pinned views)   : %.1f us" % t(lambda: c.step_host_views(a)))
ad = torch.as_tensor(a, device='cuda')
def dev():
    c.step(ad); torch.cuda.synchronize()
print("device step + sync               : %.1f us" % t(dev))
