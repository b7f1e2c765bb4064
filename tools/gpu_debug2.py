import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from oracle import oracle as orc
from bluesky_gym_hallucination_b200 import core, scenario as sc
kind, N, E = 0, 20, 512
ac, reso, nac, ev = sc.horizontal(E, N, seed=7)
cfg = orc.make_cfg(kind, N, substeps=1, asas_on=1, reso_on=1, reso_off_own=1)
rng = np.random.default_rng(3)
c32 = core.BsgCore(kind, E, N, precision='fp32', substeps=1, asas_on=True, reso_on=True, reso_off_own=True, cd_outputs=True)
F = orc.F
for s in range(30):
    act = rng.uniform(-1, 1, (E, 1))
    pre = (ac.copy(), reso.copy(), ev.copy())
    o_obs, o_rew, o_term, o_trunc, o_cd = orc.step(cfg, ac, reso, nac, ev, act, nthreads=8)
    c32.set_state(pre[0], pre[1], nac, pre[2])
    c32.step(torch.as_tensor(act, dtype=torch.float32, device='cuda')); torch.cuda.synchronize()
    ac1, reso1, _, ev1 = c32.get_state()
    d = np.abs(ac1[:, F['asas_trk']] - ac[:, F['asas_trk']])
    d = np.minimum(d, 360 - d)
    bad = np.argwhere(d > 0.01)
    nconf_g = c32.cd['nconf'].cpu().numpy()
    print('substep', s, 'bad asas_trk', len(bad), 'nconf mismatch rows', int((nconf_g != o_cd['nconf']).sum()), 'nan ref/gpu', np.isnan(ac).any((1,2)).sum(), np.isnan(ac1).any((1,2)).sum())
    for e, i in bad[:2]:
        print('  env', e, 'ac', i, 'gpu trk', ac1[e, F['asas_trk'], i], 'ref', ac[e, F['asas_trk'], i], 'pre', pre[0][e, F['asas_trk'], i],
              'nconf gpu', nconf_g[e, i], 'ref', o_cd['nconf'][e, i], 'near gpu', int(c32.cd['near_idx'][e, i]), 'ref', o_cd['near_idx'][e, i],
              'active gpu/ref', ac1[e, F['asas_active'], i], ac[e, F['asas_active'], i], 'env anyconf ref', o_cd['nconf'][e].sum(), 'gpu', nconf_g[e].sum())
        j = int(c32.cd['near_idx'][e, i])
        if j >= 0:
            p = pre[0][e]
            dd = orc.detect_dense(p[F['lat']], p[F['lon']], p[F['alt']], p[F['hdg']], p[F['tas']], p[F['vs']])
            print('     pair', i, j, 'ref tcpa', dd['tcpa'][i, j], 'dcpa', np.sqrt(dd['dcpa2'][i, j]), 'sw', dd['swconfl'][i, j], 'margin', dd['margin_conf'][i, j],
                  'gpu tcpa', float(c32.cd['near_tcpa'][e, i]), 'dcpa', float(c32.cd['near_dcpa'][e, i]))
            print('     hdg', p[F['hdg'], i], p[F['hdg'], j], 'tas', p[F['tas'], i], p[F['tas'], j], 'dlat', p[F['lat'], j] - p[F['lat'], i], 'dlon', p[F['lon'], j] - p[F['lon'], i])
    if len(bad) > 0 and s > 3:
        break
