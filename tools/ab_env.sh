#!/bin/bash
# A/B of launch-shape / kernel switches selected by environment variables, on one box, interleaved twice:
#   tools/ab_env.sh <workload> <steps> "VAR=val VAR2=val" "VAR=val" ...   ("" = defaults)
wl=$1; steps=$2; shift 2
for rep in 1 2; do for envs in "$@"; do
  env $envs python bench.py --workload $wl --steps $steps --warmup 10 --no-cpu 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('[$envs]', '$wl', 'ms_per_step %.5f kernel_ms %.5f frac %.3f e2e %.3e launches %d' % (d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['e2e']['value'], d['gpu_launches']))"
done; done
