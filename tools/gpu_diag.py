"""First-contact diagnostics on the GPU box: prints parity errors instead of asserting."""
import sys, os, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import numpy as np, torch
from oracle import oracle as orc
from bluesky_gym_hallucination_b200 import core, scenario as sc, _lib as L
from conftest import load_golden

def rel(a, b): 
    a = np.asarray(a, float); b = np.asarray(b, float)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1.0))) if a.size else 0.0

print(torch.cuda.get_device_name(0))
print('peak fp32 TF', core.peak_fp32(0, 4096), 'fp64', core.peak_fp64(0, 2048))
KIND = {'H': 0, 'V': 1, 'S': 2}
for name in ['H', 'V', 'S', 'H_asas', 'S_asas']:
    g = load_golden(name); kind = name[0]; asas = 'asas' in name
    T, _, n = g['ac0'].shape; na = 2 if kind == 'S' else 1
    for prec in ('fp64', 'fp32'):
        c = core.BsgCore(KIND[kind], T, n, precision=prec, asas_on=asas, reso_on=asas, reso_off_own=asas, cd_outputs=True)
        nac = g['nac'].astype(np.int32)
        c.set_state(g['ac0'], g['reso0'], nac, g['ev0'])
        e0 = rel(c.observe().cpu().numpy(), g['obs0'])
        c.step(torch.as_tensor(np.ascontiguousarray(g['act'][:, :na]), dtype=c.rdtype, device=c.device))
        torch.cuda.synchronize()
        ac1, reso1, _, ev1 = c.get_state()
        live = g['term'] == 0
        obs = c.obs.cpu().numpy(); rew = c.reward.cpu().numpy()
        per = {k: rel(ac1[live][:, orc.F[k]], g['ac1'][live][:, orc.F[k]]) for k in orc.FIELDS}
        print(name, prec, 'obs0', f'{e0:.2e}', 'obs1', f'{rel(obs, g["obs1"]):.2e}', 'rew', f'{rel(rew, g["rew"]):.2e}',
              'term!=', int((c.terminated.cpu().numpy() != g['term']).sum()), 'trunc!=', int((c.truncated.cpu().numpy() != g['trunc']).sum()),
              'reso!=', int((reso1[live] != g['reso1'][live]).sum()))
        print('   state', {k: f'{v:.1e}' for k, v in per.items() if v > 0})
        c.close()
# timing of the C2 / C3 shapes
for kind, gen, E, N, asas, reso in [(0, sc.horizontal, 4096, 20, 1, 0), (2, sc.sector, 8192, 50, 1, 1)]:
    ac, r, nac, ev = gen(E, N, seed=1)
    c = core.BsgCore(kind, E, N, precision='fp32', asas_on=bool(asas), reso_on=bool(reso), reso_off_own=True)
    c.set_state(ac, r, nac, ev)
    act = torch.rand((E, c.act_dim), device='cuda') * 2 - 1
    for _ in range(3): c.step(act)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): c.step(act)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    sub = {0: 10, 2: 5}[kind]
    print(f'kind {kind} E {E} N {N}: {ms:.3f} ms/step, {E / ms * 1e3:.3e} env-steps/s, {E * sub * N * (N - 1) / ms * 1e3:.3e} pair-evals/s')
    c.close()
# giant sweep timing
lat, lon, alt, trk, gs, vs = sc.raw_traffic(16, 10000, seed=3, box_nm=1000.0)
dev = 'cuda'
rec = core.cd_pack(torch.as_tensor(lat, device=dev), torch.as_tensor(lon, device=dev), torch.as_tensor(alt, dtype=torch.float32, device=dev),
                   torch.as_tensor(trk, dtype=torch.float32, device=dev), torch.as_tensor(gs, dtype=torch.float32, device=dev),
                   torch.as_tensor(vs, dtype=torch.float32, device=dev), 52.0, 4.0)
out = core.cd_sweep(rec)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): core.cd_sweep(rec, out=out)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(f'giant sweep 16x10000: {ms:.3f} ms, {16 * 1e4 * (1e4 - 1) / ms * 1e3:.3e} pair-evals/s, conflicts rows {int(out["inconf"].sum())}')
