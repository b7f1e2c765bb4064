"""Where the step kernel's time goes: event-timed launches (L2 flushed) of the C2 / C3 shapes with the CD sweep
on/off and different substep counts; the slope per substep separates fixed cost (launch, loads, env layer)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bluesky_gym_hallucination_b200 import core as bc

NOFLUSH = os.environ.get('NOFLUSH') == '1'

def timeit(fn, n=200):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    for _ in range(10): fn()
    for a, b in ev:
        if not NOFLUSH: flush.zero_()
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    t = sorted(a.elapsed_time(b) for a, b in ev)
    return 1e3 * t[len(t) // 2]

which = sys.argv[1] if len(sys.argv) > 1 else 'c2'
kind, E, N, reso = (0, 4096, 20, 0) if which == 'c2' else (2, 65536, 50, 1)
for asas in (0, 1):
    for sub in (1, 5, 10, 20):
        c = bc.BsgCore(kind, E, N, precision='fp32', asas_on=bool(asas), reso_on=bool(reso and asas), reso_off_own=True,
                       substeps=sub, max_episode_steps=0, cd_outputs=True, fixed_nac=True)
        c.reset(seed=1)
        act = torch.rand((E, c.act_dim), device='cuda') * 2 - 1
        t_obs = timeit(lambda: c.observe(), 50)
        t = timeit(lambda: c.step(act), 100 if which == 'c2' else 30)
        print(f"{which} asas {asas} substeps {sub:2d}: step {t:8.1f} us   observe-only {t_obs:6.1f} us")
        c.close()
