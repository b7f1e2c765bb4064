"""Multi-GPU tests of the giant-airspace path (run where the box has >= 2 GPUs: `gpurun --gpus N -- pytest -m gpu`).
One process per GPU (torchrun), NCCL for the plumbing, the symmetric sweep over NVLink peer memory checked bitwise
against the ordered single-GPU sweep, per rank, over several steps of CHANGING traffic."""
import os
import subprocess
import sys
import pytest
import torch
from conftest import ROOT

pytestmark = pytest.mark.gpu


def _torchrun(script, n, timeout=900):
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', f'--nproc-per-node={n}', '--master-addr', '127.0.0.1',
           '--master-port', '29531', os.path.join(ROOT, script)]
    return subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=timeout)


def test_peer_sweep_matches_ordered_sweep_across_gpus():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs on the box")
    n = 8 if n >= 8 else (4 if n >= 4 else 2)
    r = _torchrun('tools/check_peer_sweep.py', n)
    print(r.stdout[-3000:])
    assert r.returncode == 0, r.stderr[-3000:]
    assert 'PEER SWEEP CHECK OK' in r.stdout, r.stdout[-3000:]
