"""Write-only / read-only / copy HBM bandwidth on this GPU (torch ops: fill_, sum, copy_), for the ND-map roofline:
the ND-map kernel only WRITES (1.13 GB per build), and MEASURED_PEAKS.json's hbm_gbs is a COPY figure (read + write bytes)."""
import torch
dev = torch.device('cuda')
n = 1 << 30
a = torch.empty(n, dtype=torch.uint8, device=dev); b = torch.empty(n, dtype=torch.uint8, device=dev)
def t(f, reps=10):
    f(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); f(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best
ms = t(lambda: a.zero_()); print(f"write-only  memset 1 GiB: {ms:.3f} ms  {n / ms / 1e6:.1f} GB/s")
ms = t(lambda: a.fill_(7)); print(f"write-only  fill_  1 GiB: {ms:.3f} ms  {n / ms / 1e6:.1f} GB/s")
a32 = a.view(torch.int32)
ms = t(lambda: a32.sum()); print(f"read-only   sum    1 GiB: {ms:.3f} ms  {n / ms / 1e6:.1f} GB/s")
ms = t(lambda: b.copy_(a)); print(f"copy        1 GiB -> 1 GiB: {ms:.3f} ms  {2 * n / ms / 1e6:.1f} GB/s (read + write)")
