#!/bin/bash
# A/B of kernel variants on one box: tools/ab.sh <workload> <steps> lib1.so lib2.so ...  (each run twice, interleaved)
wl=$1; steps=$2; shift 2
for rep in 1 2; do for lib in "$@"; do
  BSG_LIB_PATH=$PWD/$lib python bench.py --workload $wl --steps $steps --warmup 10 --no-cpu 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$lib', '$wl', 'ms_per_step %.5f kernel_ms %.5f frac %.3f e2e %.3e launches %d' % (d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['e2e']['value'], d['gpu_launches']))"
done; done
