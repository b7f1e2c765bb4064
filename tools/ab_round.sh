mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
bash tools/ab.sh c2 500 build/ab/split.so build/ab/sign.so 2>&1 | tee gpurun_out/ab_sign.log
bash tools/ab.sh c3 100 build/ab/split.so build/ab/sign.so 2>&1 | tee -a gpurun_out/ab_sign.log
