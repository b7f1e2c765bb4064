#!/bin/bash
# GPU pass 2: pipe micro-benchmarks, ND-map v2 (tests + bench), parity tests with the oracle trace
mkdir -p gpurun_out
./build/ubench_fma > gpurun_out/ubench_fma.log 2>&1; cat gpurun_out/ubench_fma.log
timeout 1500 python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
grep -E "^\[|passed|failed|Error|error" gpurun_out/pytest_gpu.log | cut -c1-700 | tail -40
timeout 300 python bench.py --workload c4 --steps 50 --warmup 5 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; tail -1 gpurun_out/bench_c4.json | cut -c1-900
