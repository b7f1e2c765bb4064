#!/bin/bash
# launch-size sweep of the C2 kernel (envs per launch -> us per step, roofline fraction): profiles/envsweep_c2_r2.log
for e in 148 592 888 1776 2664 3552 4096 4440 8192 16384; do
python bench.py --workload c2 --envs $e --steps 300 --warmup 10 --no-cpu 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('envs $e', 'ms_per_step %.5f frac %.3f' % (d['ms_per_step'], d['roofline']['frac']))"
done
