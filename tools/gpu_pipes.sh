#!/bin/bash
# per-pipe executed-instruction counts of the hot kernels (one ncu metrics pass each, after the plain run exited 0)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
grep -E "passed|failed|Error" gpurun_out/pytest_gpu.log | cut -c1-300 | tail -6
for w in c2 c3 c4 c5; do
  timeout 300 python bench.py --workload $w --steps 50 --warmup 5 --no-cpu > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err || { echo "bench $w failed"; continue; }
  tail -1 gpurun_out/bench_$w.json | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["config"]["workload"], d["ms_per_step"], d["roofline"]["frac"])'
  case $w in c4) kn=ndmap_kernel;; c5) kn=cd_sym_sweep_kernel;; *) kn=step_kernel;; esac
  timeout 600 ncu --metrics sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_fmaheavy.sum,sm__inst_executed_pipe_fmalite.sum,sm__inst_executed_pipe_fp64.sum,sm__inst_executed_pipe_xu.sum,sm__inst_executed_pipe_lsu.sum,sm__inst_executed_pipe_cbu.sum,sm__inst_executed_pipe_adu.sum,sm__inst_executed_pipe_uniform.sum,sm__inst_executed_pipe_fp16.sum,sm__inst_executed_pipe_tex.sum,sm__inst_executed.sum,smsp__inst_executed.sum,sm__cycles_active.avg,gpu__time_duration.sum --clock-control none -k regex:$kn -s 8 -c 1 --csv --log-file gpurun_out/pipes_${w}_r2.csv python bench.py --workload $w --steps 4 --warmup 5 --no-cpu > gpurun_out/ncu_pipes_$w.log 2>&1; grep -c pipe gpurun_out/pipes_${w}_r2.csv
done
