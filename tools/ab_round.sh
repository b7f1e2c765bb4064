mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
bash tools/ab.sh c3 100 build/ab/base.so build/ab/split.so 2>&1 | tee gpurun_out/ab_split.log
bash tools/ab.sh c2 500 build/ab/base.so build/ab/split.so 2>&1 | tee -a gpurun_out/ab_split.log
bash tools/ab_env.sh c2 500 BSG_HOST_DIRECT=0 BSG_HOST_DIRECT=1 2>&1 | tee gpurun_out/ab_direct.log
bash tools/ab_env.sh c3 100 BSG_HOST_DIRECT=0 BSG_HOST_DIRECT=1 2>&1 | tee -a gpurun_out/ab_direct.log
