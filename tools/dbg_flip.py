"""Debug aid: where does the fp32 step leave the oracle?  Re-creates the batch of tests/test_gpu_parity.py
test_batched_step_vs_oracle(kind, N), goes to the pre-state of step S, then advances env ENV one substep at a time in RAW
mode (traffic only: kinematics + CD + MVP + resumenav) with the oracle and with the fp32 kernels, printing the first
substeps at which aircraft differ.   python tools/dbg_flip.py kind N step env [env ...]"""
import sys, os
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle as orc
from bluesky_gym_hallucination_b200 import scenario as sc, _lib as L
from bluesky_gym_hallucination_b200.core import BsgCore

kind, N, S = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
envs = [int(x) for x in sys.argv[4:]]
E = 512
gen = {0: sc.horizontal, 1: sc.vertical, 2: sc.sector}[kind]
ac, reso, nac, ev = gen(E, N, seed=7)
cfg = orc.make_cfg(kind, N, asas_on=1, reso_on=1, reso_off_own=1)
rng = np.random.default_rng(3)
na = 2 if kind == 2 else 1
for s in range(S + 1):
    act = rng.uniform(-1, 1, (E, na))
    if s == S: break
    orc.step(cfg, ac, reso, nac, ev, act, nthreads=8)
F = orc.F
K, dt = cfg.substeps, cfg.simdt
# one env-layer step first on both sides so the action is taken in, but with ZERO substeps?  Not available: RAW from the
# pre-state instead (the agent keeps its old autopilot targets; the flips under study are between intruders)
rcfg = orc.make_cfg(orc.KIND_RAW, N, substeps=1, simdt=dt, asas_on=1, reso_on=1, reso_off_own=1)
o_ac, o_reso = ac[envs].copy(), reso[envs].copy()
n_e = len(envs)
core = BsgCore(L.KIND_RAW, n_e, N, precision='fp32', substeps=1, simdt=dt, asas_on=True, reso_on=True, reso_off_own=True, cd_outputs=True)
core.set_state(o_ac, o_reso, nac[envs], ev[envs])
def pair_terms(acs, i, n, rpz=5 * 1852.0, fach=1.05, dtlook=300.0):
    """fp64 CD + MVP terms of aircraft i against everybody (oracle formulas, oracle/c/bsgo_cd.c, bsgo_asas.c)."""
    lat, lon, hdg, tas = acs[F['lat'], :n], acs[F['lon'], :n], acs[F['hdg'], :n], acs[F['tas'], :n]
    u, v = tas * np.sin(np.radians(hdg)), tas * np.cos(np.radians(hdg))
    out = []
    for j in range(n):
        if j == i: continue
        cav = np.cos(np.radians(0.5 * (lat[i] + lat[j])))
        dx = 6371000.0 * np.radians(lon[j] - lon[i]) * cav; dy = 6371000.0 * np.radians(lat[j] - lat[i])
        dist = np.hypot(dx, dy)
        du, dv = u[i] - u[j], v[i] - v[j]                # ownship minus intruder (statebased.py)
        dv2 = max(du * du + dv * dv, 1e-6)
        tcpa = -(du * (-dx) + dv * (-dy)) / dv2 if False else (du * dx + dv * dy) / dv2
        dcpa2 = abs(dist * dist - tcpa * tcpa * dv2)
        R2 = rpz * rpz
        if dcpa2 >= R2: continue
        dtin = np.sqrt((R2 - dcpa2) / dv2)
        tin, tout = tcpa - dtin, tcpa + dtin
        if not (tout > 0 and tin < dtlook): continue
        vrx, vry = -du, -dv                              # intruder minus ownship
        dcx, dcy = dx + vrx * tcpa, dy + vry * tcpa
        dabsH = np.hypot(dcx, dcy)
        out.append((j, dist, tcpa, np.sqrt(dcpa2), dabsH, np.hypot(vrx, vry), tin))
    return out

names = ['hdg', 'tas', 'asas_trk', 'asas_tas', 'asas_active', 'lat', 'lon']
for k in range(K):
    o_prev = o_ac.copy()
    _, _, _, _, ocd, tr = orc.step_trace(rcfg, o_ac, o_reso, nac[envs].copy(), ev[envs].copy(), None, 2e-4, nthreads=1)
    core.step(None); torch.cuda.synchronize()
    g_ac, g_reso, _, _ = core.get_state()
    gcd = {k2: v.cpu().numpy() for k2, v in core.cd.items()}
    for q, e in enumerate(envs):
        n = int(nac[e])
        dtrk = np.abs((g_ac[q, F['asas_trk'], :n] - o_ac[q, F['asas_trk'], :n] + 180) % 360 - 180)
        dhdg = np.abs((g_ac[q, F['hdg'], :n] - o_ac[q, F['hdg'], :n] + 180) % 360 - 180)
        dtas = np.abs(g_ac[q, F['asas_tas'], :n] - o_ac[q, F['asas_tas'], :n])
        dact = g_ac[q, F['asas_active'], :n] != o_ac[q, F['asas_active'], :n]
        drow = g_reso[q, :n] != o_reso[q, :n]
        dnc = gcd['nconf'][q, :n] != ocd['nconf'][q, :n]
        bad = (dtrk > 0.02) | (dtas > 0.02) | dact | drow | dnc | (dhdg > 0.02)
        print(f"env {e} substep {k}: differing aircraft {np.where(bad)[0].tolist()}")
        for j in np.where(bad)[0][:8]:
            print(f"    ac {j}: dtrk {dtrk[j]:.4f} dhdg {dhdg[j]:.4f} dtas {dtas[j]:.4f} active o/g {o_ac[q, F['asas_active'], j]:.0f}/{g_ac[q, F['asas_active'], j]:.0f} "
                  f"row o/g {int(o_reso[q, j]):#x}/{int(g_reso[q, j]):#x} nconf o/g {ocd['nconf'][q, j]}/{gcd['nconf'][q, j]} "
                  f"margin {tr['res_margin'][q, j]:.3e} kappa {tr['kappa'][q, j]:.3e} near_conf {tr['near_conf'][q, j]}")
            print(f"        asas_trk o/g {o_ac[q, F['asas_trk'], j]:.5f}/{g_ac[q, F['asas_trk'], j]:.5f} asas_tas o/g {o_ac[q, F['asas_tas'], j]:.5f}/{g_ac[q, F['asas_tas'], j]:.5f} tas {o_prev[q, F['tas'], j]:.3f} hdg {o_prev[q, F['hdg'], j]:.4f}")
            for t in pair_terms(o_prev[q], j, n):
                print('        conflict with %d: dist %.2f tcpa %.4f dcpa %.3f dabsH %.3f vrel %.3f tin %.3f' % t)
core.close()
