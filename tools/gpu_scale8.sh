#!/bin/bash
# 8-GPU bench lines (C2, C3 env-sharded; C5 giant airspace over NVLink peer memory)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $TR bench.py --gpus 8 --no-cpu > gpurun_out/bench_c2_8gpu.json 2> gpurun_out/bench_c2_8gpu.err; echo "c2 rc=$?"
timeout 600 $TR bench.py --gpus 8 --workload c3 --steps 200 --warmup 5 --no-cpu > gpurun_out/bench_c3_8gpu.json 2> gpurun_out/bench_c3_8gpu.err; echo "c3 rc=$?"
timeout 600 $TR bench.py --gpus 8 --workload c5 --steps 200 --warmup 5 --no-cpu > gpurun_out/bench_c5_8gpu.json 2> gpurun_out/bench_c5_8gpu.err; echo "c5 rc=$?"
for f in c2 c3 c5; do tail -1 gpurun_out/bench_${f}_8gpu.json | cut -c1-160; done
