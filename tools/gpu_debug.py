import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import numpy as np, torch
from oracle import oracle as orc
from bluesky_gym_hallucination_b200 import core, scenario as sc
kind, N, E = int(sys.argv[1]), int(sys.argv[2]), 512
gen = {0: sc.horizontal, 1: sc.vertical, 2: sc.sector}[kind]
ac, reso, nac, ev = gen(E, N, seed=7)
cfg = orc.make_cfg(kind, N, asas_on=1, reso_on=1, reso_off_own=1)
rng = np.random.default_rng(3); na = 2 if kind == 2 else 1
c32 = core.BsgCore(kind, E, N, precision='fp32', asas_on=True, reso_on=True, reso_off_own=True, cd_outputs=True)
print('tpb env', os.environ.get('BSG_FORCE_TPB'))
for s in range(3):
    act = rng.uniform(-1, 1, (E, na))
    pre = (ac.copy(), reso.copy(), ev.copy())
    o_obs, o_rew, o_term, o_trunc, o_cd = orc.step(cfg, ac, reso, nac, ev, act, nthreads=8)
    c32.set_state(pre[0], pre[1], nac, pre[2])
    c32.step(torch.as_tensor(act, dtype=torch.float32, device='cuda')); torch.cuda.synchronize()
    ac1, reso1, _, ev1 = c32.get_state()
    obs = c32.obs.cpu().numpy()
    d = np.abs(obs - o_obs) / np.maximum(np.abs(o_obs), 1)
    badenv = np.where(np.nanmax(d, 1) > 1e-4)[0]
    print('step', s, 'bad envs', len(badenv), badenv[:10], 'nan envs', np.isnan(o_obs).any(1).sum(), np.isnan(obs).any(1).sum())
    for e in badenv[:3]:
        print(' env', e, 'nac', nac[e])
        for f in ['lat', 'lon', 'hdg', 'tas', 'asas_active', 'asas_trk', 'asas_tas', 'aptrk']:
            dd = np.abs(ac1[e, orc.F[f]] - ac[e, orc.F[f]])
            j = np.argmax(dd)
            print('   ', f, 'maxdiff', dd.max(), 'ac', j, 'gpu', ac1[e, orc.F[f], j], 'ref', ac[e, orc.F[f], j], 'pre', pre[0][e, orc.F[f], j])
        print('    reso gpu', [hex(int(x)) for x in reso1[e][:8]], 'ref', [hex(int(x)) for x in reso[e][:8]], 'pre', [hex(int(x)) for x in pre[1][e][:8]])
        print('    nconf gpu', c32.cd['nconf'][e].cpu().numpy(), 'ref', o_cd['nconf'][e])
