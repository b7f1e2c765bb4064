#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
grep -E "^\[|passed|failed|Error|error" gpurun_out/pytest_gpu.log | cut -c1-1500 | tail -30
for w in c3; do
  for rep in 1 2; do
  timeout 300 python bench.py --workload $w --steps 200 --warmup 10 --no-cpu 2>/dev/null | tail -1 | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print("list:", d["ms_per_step"], d["roofline"]["frac"])'
  BSG_FORCE_WALK=1 timeout 300 python bench.py --workload $w --steps 200 --warmup 10 --no-cpu 2>/dev/null | tail -1 | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print("walk:", d["ms_per_step"], d["roofline"]["frac"])'
  done
done
timeout 600 ncu --set full --import-source on --clock-control none -k regex:step_kernel -s 12 -c 1 -o gpurun_out/step_c3_r2 -f python bench.py --workload c3 --steps 4 --warmup 10 --no-cpu > gpurun_out/ncu_c3.log 2>&1; tail -2 gpurun_out/ncu_c3.log
