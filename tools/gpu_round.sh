#!/bin/bash
# one-GPU round: GPU tests, then (each only after its plain run exited 0) ncu captures named by $1 = list of c2 c3 c4 c5
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
grep -E "^\[reset|passed|failed|Error|error" gpurun_out/pytest_gpu.log | cut -c1-1500 | tail -12
for w in $1; do
  timeout 300 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err || { echo "bench $w failed"; continue; }
  tail -1 gpurun_out/bench_$w.json | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["config"]["workload"], d["ms_per_step"], d["roofline"]["frac"])'
  case $w in c4) kn=ndmap_kernel;; c5) kn=cd_sym_sweep_kernel;; *) kn=step_kernel;; esac
  timeout 600 ncu --set full --import-source on --clock-control none -k regex:$kn -s 8 -c 1 -o gpurun_out/ncu_${w}_r2 -f python bench.py --workload $w --steps 4 --warmup 5 --no-cpu > gpurun_out/ncu_$w.log 2>&1; tail -1 gpurun_out/ncu_$w.log
done
# launch list of the default bench command (kernel share of the step; cold-cache serialised times)
if [ -n "$2" ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_default_r2.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/ncu_launches.log 2>&1; tail -1 gpurun_out/ncu_launches.log | cut -c1-200
fi
