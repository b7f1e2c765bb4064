#!/bin/bash
# One GPU pass: parity tests, smoke, bench lines (C2 default, reference arm, C3), ncu launch list + full capture of the C2 step kernel.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench c2 rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
timeout 600 python bench.py --workload c3 --steps 200 --warmup 5 --no-cpu > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench c3 rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 5 -c 1 -o gpurun_out/step_c2 -f python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/ncu_full.log 2>&1; echo "ncu full c2 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 5 -c 1 -o gpurun_out/step_c3 -f python bench.py --workload c3 --steps 10 --warmup 3 --no-cpu > gpurun_out/ncu_full_c3.log 2>&1; echo "ncu full c3 rc=$?"
cat gpurun_out/bench_c2.json | tail -1 | cut -c1-600
