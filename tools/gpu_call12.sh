#!/bin/bash
# 8-GPU pass: peer-sweep check, multi-GPU test, default bench at N = 8 and 4 (c2 weak + c3 strong + c5 peer)
mkdir -p gpurun_out
N=${1:-8}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/check_peer_sweep.py > gpurun_out/check_peer_sweep_${N}gpu.log 2>&1; echo "check rc=$?"; grep -E "PEER SWEEP CHECK|Error|error" gpurun_out/check_peer_sweep_${N}gpu.log | tail -5
for n in $N 4; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $n --steps 200 --warmup 5 > gpurun_out/bench_default_${n}gpu.json 2> gpurun_out/bench_default_${n}gpu.err; echo "bench $n rc=$?"; tail -2 gpurun_out/bench_default_${n}gpu.err
python - $n <<'PY'
import json, sys
n=sys.argv[1]
d=json.loads(open(f'gpurun_out/bench_default_{n}gpu.json').read().strip().splitlines()[-1])
print('c2', d['n_gpus'], d['ms_per_step'], d['value'], d['roofline']['frac'], 'e2e', d['e2e']['value'])
for k in ('c3','c5'):
    x=d.get(k); print(k, x and (x['n_gpus'], x['scaling'], x['ms_per_step'], x['value'], x['roofline']['frac'], x.get('e2e',{}).get('value')))
PY
done
