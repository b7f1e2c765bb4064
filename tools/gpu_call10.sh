#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
grep -E "passed|failed|Error|error" gpurun_out/pytest_gpu.log | cut -c1-600 | tail -8
timeout 900 python bench.py --steps 200 --warmup 5 > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_default.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_default.json').read().strip().splitlines()[-1])
print('c2', d['ms_per_step'], d['value'], d['roofline']['frac'], 'e2e', d['e2e']['value'], d['clocks'])
for k in ('c3','c5','v'):
    x=d.get(k); print(k, x and (x['ms_per_step'], x['value'], x['roofline']['frac'], x.get('e2e',{}).get('value')))
print('cpu', d.get('cpu_baseline',{}).get('value'), d.get('cpu_baseline_c_openmp',{}).get('value'), 'c1', d.get('c1',{}).get('value'), d.get('numpy_detect',{}).get('by_aircraft'))
print('traffic_state', d['traffic_state'])
PY
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 | cut -c1-400
