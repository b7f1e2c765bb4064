"""Phase timeline of step_kernel from a -DBSG_TRACE build (BSG_LIB_PATH=build/ab/T_trace.so):
per-CTA %globaltimer stamps at entry / after loads / after substeps / after state store / after env layer."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from bluesky_gym_hallucination_b200 import core as bc
which = sys.argv[1] if len(sys.argv) > 1 else 'c2'
kind, E, N, reso = (0, 4096, 20, 0) if which == 'c2' else (2, 65536, 50, 1)
for mode in ('observe', 'step'):
    c = bc.BsgCore(kind, E, N, precision='fp32', asas_on=True, reso_on=bool(reso), reso_off_own=True,
                   max_episode_steps=0, cd_outputs=True, fixed_nac=True)
    c.reset(seed=1); c.set_autoreset(True, seed=3)
    act = torch.rand((E, c.act_dim), device='cuda') * 2 - 1
    flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
    fn = (lambda: c.observe()) if mode == 'observe' else (lambda: c.step(act))
    for _ in range(20): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    flush.zero_(); e0.record(); fn(); e1.record(); torch.cuda.synchronize()
    buf = np.zeros(8 * 16384, np.uint64)
    c.lib.bsg_debug_trace.argtypes = [C.c_void_p, C.c_int]
    assert c.lib.bsg_debug_trace(buf.ctypes.data_as(C.c_void_p), buf.size) == 0
    ncta = int((buf.reshape(16384, 8)[:, 0] != 0).sum())      # CTAs that left a stamp (the array is cleared per launch below)
    t = buf.reshape(16384, 8)[:ncta, :5].astype(np.int64)
    t0 = t[:, 0].min()
    print(f"{which} {mode}: events {1e3 * e0.elapsed_time(e1):.1f} us; in-kernel span {(t[:, 4].max() - t0) / 1e3:.1f} us; CTA start spread "
          f"{(t[:, 0].max() - t0) / 1e3:.1f} us")
    d = np.diff(t, axis=1) / 1e3
    for nm, col in zip(('loads+barrier', 'substeps', 'state store', 'env layer'), d.T):
        print(f"   {nm:14s} per CTA: median {np.median(col):7.2f}  p95 {np.percentile(col, 95):7.2f}  max {col.max():7.2f} us")
    full = buf.reshape(16384, 8)[:ncta].astype(np.int64)
    if full[:, 7].any():
        print(f"   prologue: entry -> first batch of loads consumed (reset test) median {np.median(full[:, 7] - full[:, 0]) / 1e3:.2f} us, "
              f"-> state derived, scalars staged, barrier {np.median(full[:, 1] - full[:, 7]) / 1e3:.2f} us")
    if full[:, 5].any() and mode == 'step':     # detection-only kernels: first chunk's integration / sweeps
        print(f"   first chunk: integration + barrier median {np.median(full[:, 5] - full[:, 1]) / 1e3:.2f} us, "
              f"sweeps median {np.median(full[:, 6] - full[:, 5]) / 1e3:.2f} us; rest of the substeps {np.median(full[:, 2] - full[:, 6]) / 1e3:.2f} us")
    print(f"   CTA lifetime median {np.median(t[:, 4] - t[:, 0]) / 1e3:.2f} us, max {(t[:, 4] - t[:, 0]).max() / 1e3:.2f} us")
    c.close()
