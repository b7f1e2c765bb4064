#!/bin/bash
# GPU pass 3: chunked detection-only step kernel (kinematics ahead of one barrier) A/B, parity tests, ncu of c2/c3
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
grep -E "^\[|passed|failed|Error|error" gpurun_out/pytest_gpu.log | cut -c1-1500 | tail -40
for nb in 2 3 5; do
  BSG_FORCE_NBUF=$nb timeout 300 python bench.py --workload c2 --steps 200 --warmup 10 --no-cpu > gpurun_out/ab_nbuf${nb}_c2.json 2> gpurun_out/ab_nbuf${nb}_c2.err
  echo "nbuf=$nb c2: $(tail -1 gpurun_out/ab_nbuf${nb}_c2.json | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["roofline"])')"
done
timeout 300 python bench.py --workload c3 --steps 100 --warmup 10 --no-cpu > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; tail -1 gpurun_out/bench_c3.json | cut -c1-600
# ncu (after the plain commands above exited): c2 full set of the step kernel
timeout 600 ncu --set full --import-source on --clock-control none -k regex:step_kernel -s 12 -c 1 -o gpurun_out/step_c2_r2 -f python bench.py --workload c2 --steps 4 --warmup 10 --no-cpu > gpurun_out/ncu_c2.log 2>&1; tail -2 gpurun_out/ncu_c2.log
timeout 600 ncu --set full --import-source on --clock-control none -k regex:step_kernel -s 12 -c 1 -o gpurun_out/step_c3_r2 -f python bench.py --workload c3 --steps 4 --warmup 10 --no-cpu > gpurun_out/ncu_c3.log 2>&1; tail -2 gpurun_out/ncu_c3.log
