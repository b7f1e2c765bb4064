import sys, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), 'tests'))
import numpy as np, torch
from conftest import load_golden
from oracle import oracle as orc
from test_oracle_golden import golden_cfg
from test_gpu_parity import make_core, run_step, relerr, STATE_FIELDS, KIND, EPS32, RT64, obs_tol32
orc.lib()
for name in ['V_mvpv', 'V_mvphv', 'H_mvphv', 'S_mvpv']:
    g = load_golden(name); kind = name[0]; mode = 2 if 'mvphv' in name else 1
    T, _, n = g['ac0'].shape; na = 2 if kind == 'S' else 1
    nac = g['nac'].astype(np.int32); cfg = golden_cfg(orc, name, g)
    act = np.ascontiguousarray(g['act'][:, :na])
    o_ac, o_reso, o_ev = g['ac0'].copy(), g['reso0'].copy(), g['ev0'].copy()
    o_obs, o_rew, o_term, o_trunc, o_cd, tr = orc.step_trace(cfg, o_ac, o_reso, nac, o_ev, act, EPS32, nthreads=4)
    margin = tr['res_margin'].min(1); live = g['term'] == 0
    idx = [orc.F[k] for k in STATE_FIELDS]
    for prec, eps in (('fp64', 1e-9), ('fp32', EPS32)):
        core = make_core(KIND[kind], T, n, prec, 1, reso_mode=mode)
        obs, rew, term, trunc, info, ac1, reso1, ev1, cd = run_step(core, g['ac0'], g['reso0'], nac, g['ev0'], act)
        drop = (margin < eps) & live
        tol = RT64 * 200 if prec == 'fp64' else None
        def rows_ok(a, b):
            if prec == 'fp64':
                return (np.abs(a - b) / np.maximum(np.abs(b), 1.0) < tol).all(1) | (np.isnan(a) & np.isnan(b)).all(1)
            return (np.abs(a - b) <= obs_tol32(kind, b)).all(1)
        m_gold = rows_ok(obs, g['obs1']) & (reso1 == g['reso1']).all(1)
        m_orc = rows_ok(obs, o_obs) & (reso1 == o_reso).all(1)
        orc_gold = (np.abs(o_obs - g['obs1']) / np.maximum(np.abs(g['obs1']), 1.0) < RT64 * 200).all(1)
        print(f"{name} {prec}: T={T} live={int(live.sum())} dropped={int(drop.sum())}; dropped matching golden {int((drop & m_gold).sum())}, "
              f"dropped matching C oracle {int((drop & m_orc).sum())}, dropped matching neither {int((drop & ~m_gold & ~m_orc).sum())}; "
              f"kept not matching golden {int((~drop & live & ~m_gold).sum())}; oracle==golden on dropped {int((drop & orc_gold).sum())}")
        core.close()
