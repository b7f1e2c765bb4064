#!/bin/bash
# N-GPU bench lines: tools/gpu_scale.sh N   (C2, C3 env-sharded; C5 giant airspace over NVLink peer memory)
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $TR bench.py --gpus $N --no-cpu > gpurun_out/bench_c2_${N}gpu.json 2> gpurun_out/bench_c2_${N}gpu.err; echo "c2 rc=$?"
timeout 600 $TR bench.py --gpus $N --workload c3 --steps 200 --warmup 5 --no-cpu > gpurun_out/bench_c3_${N}gpu.json 2> gpurun_out/bench_c3_${N}gpu.err; echo "c3 rc=$?"
timeout 600 $TR bench.py --gpus $N --workload c5 --steps 200 --warmup 5 --no-cpu > gpurun_out/bench_c5_${N}gpu.json 2> gpurun_out/bench_c5_${N}gpu.err; echo "c5 rc=$?"
for f in c2 c3 c5; do tail -1 gpurun_out/bench_${f}_${N}gpu.json | cut -c1-160; done
