import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from oracle import oracle as orc
from bluesky_gym_hallucination_b200 import core, scenario as sc
kind, N, E = 0, 20, 512
ac, reso, nac, ev = sc.horizontal(E, N, seed=7)
cfg = orc.make_cfg(kind, N, asas_on=1, reso_on=1, reso_off_own=1)
rng = np.random.default_rng(3)
c32 = core.BsgCore(kind, E, N, precision='fp32', asas_on=True, reso_on=True, reso_off_own=True, cd_outputs=True)
F = orc.F
for s in range(3):
    act = rng.uniform(-1, 1, (E, 1))
    pre = (ac.copy(), reso.copy(), ev.copy())
    o_obs, o_rew, o_term, o_trunc, o_cd = orc.step(cfg, ac, reso, nac, ev, act, nthreads=8)
    c32.set_state(pre[0], pre[1], nac, pre[2])
    c32.step(torch.as_tensor(act, dtype=torch.float32, device='cuda')); torch.cuda.synchronize()
    ac1, reso1, _, ev1 = c32.get_state()
    obs = c32.obs.cpu().numpy()
    err = np.abs(obs - o_obs) / np.maximum(np.abs(o_obs), 1)
    bad = np.argwhere(err > 1e-3)
    print('step', s, 'bad', len(bad))
    for e, k in bad[:4]:
        i = (k - 57) % 19 + 1 if k >= 57 else k % 19 + 1
        print(' env', e, 'obs idx', k, 'ac', i, 'gpu', obs[e, k], 'ref', o_obs[e, k])
        for f in ['hdg', 'tas', 'asas_active', 'asas_trk', 'asas_tas', 'aptrk', 'selspd', 'ax', 'phase']:
            print('    ', f, 'gpu', ac1[e, F[f], i], 'ref', ac[e, F[f], i], 'pre', pre[0][e, F[f], i])
        print('     reso gpu', hex(int(reso1[e, i])), 'ref', hex(int(reso[e, i])), 'pre', hex(int(pre[1][e, i])))
