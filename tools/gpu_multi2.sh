#!/bin/bash
# 2-GPU pass: multi-GPU test, peer-sweep check output, default bench at N=2 (c2 weak + c3 strong + c5 peer)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q -s -p no:cacheprovider > gpurun_out/pytest_multi_2gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_multi_2gpu.log; tail -5 gpurun_out/pytest_multi_2gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/check_peer_sweep.py > gpurun_out/check_peer_sweep_2gpu.log 2>&1; echo "check rc=$?"; grep -E "PEER SWEEP CHECK|Error|error" gpurun_out/check_peer_sweep_2gpu.log | tail -5
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 200 --warmup 5 > gpurun_out/bench_default_2gpu.json 2> gpurun_out/bench_default_2gpu.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_default_2gpu.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_default_2gpu.json').read().strip().splitlines()[-1])
print('c2', d['n_gpus'], d['ms_per_step'], d['value'], d['roofline']['frac'], 'e2e', d['e2e']['value'])
for k in ('c3','c5'):
    x=d.get(k); print(k, x and (x['n_gpus'], x['scaling'], x['ms_per_step'], x['value'], x['roofline']['frac'], x.get('e2e',{}).get('value')))
PY
