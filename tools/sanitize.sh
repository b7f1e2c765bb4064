#!/bin/bash
# compute-sanitizer passes over the GPU tests that exercise shared-memory atomics, warp-shuffle reductions, the pair work
# list / slot rows of the resolution phase, the ND-map and the reset paths (VERDICT r2 "Sanitizer evidence").
# Run on the GPU box: tools/sanitize.sh  ->  gpurun_out/sanitizer_r3_{memcheck,racecheck,synccheck}.log
# (copy the summaries to profiles/).  Each pass is bounded by `timeout`.
K='batched_step_vs_oracle or pair_work_list or ndmap_bit_exact or cd_sweep_symmetric or no_alloc or autoreset'
KS='pair_work_list or mvp or batched_step_vs_oracle'
mkdir -p gpurun_out
for tool in memcheck racecheck; do
  timeout ${SAN_TIMEOUT:-900} compute-sanitizer --tool $tool --print-limit 30 --log-file gpurun_out/sanitizer_r3_$tool.log \
      python -m pytest tests -q -m gpu -x -p no:cacheprovider -k "$K" > gpurun_out/sanitizer_r3_${tool}_pytest.log 2>&1
  echo "$tool: pytest rc=$?" >> gpurun_out/sanitizer_r3_${tool}_pytest.log
done
timeout ${SAN_TIMEOUT:-900} compute-sanitizer --tool synccheck --print-limit 30 --log-file gpurun_out/sanitizer_r3_synccheck.log \
    python -m pytest tests -q -m gpu -x -p no:cacheprovider -k "$KS" > gpurun_out/sanitizer_r3_synccheck_pytest.log 2>&1
echo "synccheck: pytest rc=$?" >> gpurun_out/sanitizer_r3_synccheck_pytest.log
tail -n 4 gpurun_out/sanitizer_r3_*.log
