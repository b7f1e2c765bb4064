#!/bin/bash
# GPU pass 1 of round 2: parity tests with the packed sweep, smoke, A/B of the packed-trip interleave on C2 / C3
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt
timeout 1500 python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
L=bluesky_gym_hallucination_b200/libbsgcore.so
timeout 900 tools/ab.sh c2 1000 $L build/lib_pg2.so build/lib_pg4.so build/lib_pg5.so > gpurun_out/ab_pg_c2.log 2>&1; cat gpurun_out/ab_pg_c2.log
timeout 900 tools/ab.sh c3 100 $L build/lib_pg2.so build/lib_pg4.so build/lib_pg5.so > gpurun_out/ab_pg_c3.log 2>&1; cat gpurun_out/ab_pg_c3.log
