import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np, torch
from bluesky_gym_hallucination_b200 import core, scenario as sc
from test_gpu_kernels import _pack
lat, lon, alt, trk, gs, vs = sc.raw_traffic(2, 1200, seed=11, box_nm=300.0)
rec = _pack(lat, lon, alt, trk, gs, vs)
full = core.cd_sweep(rec); sym = core.cd_sweep_sym(rec); torch.cuda.synchronize()
for k in ('inconf', 'nconf', 'nlos', 'near_idx', 'near_tcpa', 'near_dcpa', 'near_tin', 'tcpamax'):
    a, b = full[k].cpu().numpy(), sym[k].cpu().numpy()
    bad = np.argwhere(a != b)
    print(k, len(bad))
    for e, i in bad[:5]:
        print('   ', e, i, a[e, i], b[e, i], 'nconf', full['nconf'][e, i].item(), 'idx', full['near_idx'][e, i].item(), sym['near_idx'][e, i].item(),
              'tcpa', full['near_tcpa'][e,i].item(), sym['near_tcpa'][e,i].item())
