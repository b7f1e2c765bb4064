#!/bin/bash
# A/B of launch shapes with one library: tools/ab_env.sh <workload> <steps> "ENV=.. ENV=.." "ENV=.." ...
wl=$1; steps=$2; shift 2
for rep in 1 2; do for cfg in "$@"; do
  env $cfg python bench.py --workload $wl --steps $steps --warmup 10 --no-cpu 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$cfg', '$wl', 'ms_per_step %.5f frac %.3f e2e %.3e' % (d['ms_per_step'], d['roofline']['frac'], d['e2e']['value']))"
done; done
