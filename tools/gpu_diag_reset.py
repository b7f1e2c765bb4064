"""Diagnostic: fraction of envs ending per step under random actions (auto-reset on), per env kind."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bluesky_gym_hallucination_b200 import core as bc
for kind, N, reso, fixed in ((0, 20, 0, 1), (2, 50, 1, 1), (2, 50, 1, 0), (1, 6, 0, 1)):
    E = 8192
    c = bc.BsgCore(kind, E, N, precision='fp32', asas_on=True, reso_on=bool(reso), reso_off_own=True,
                   max_episode_steps={0: 300, 1: 300, 2: 200}[kind], fixed_nac=bool(fixed))
    c.reset(seed=1); c.set_autoreset(True, seed=2)
    tot = []
    for s in range(300):
        a = torch.rand((E, c.act_dim), device='cuda') * 2 - 1
        c.step(a)
        tot.append(int((c.terminated | c.truncated).sum()))
    print(f"kind {kind} N {N} fixed {fixed}: done/step first 10 {tot[:10]} mean(last 200) {sum(tot[100:]) / 200 / E:.4f} of envs; "
          f"live aircraft mean {c.live_aircraft().mean():.1f}")
    c.close()
