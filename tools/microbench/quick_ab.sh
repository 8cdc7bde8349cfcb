#!/bin/bash
# quick A/B of the shipped geometries: default, persistent, observer on, 4 096 and 262 144 envs (BlindGrasping)
run() {
  env $1 python bench.py --steps 2000 --warmup 100 --no-cpu-baseline $2 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        print('$1 $2', '| step_us', round(d['ms_per_step'] * 1e3, 2), 'post', round(d['kernels_us']['post_physics'], 2), 'pre', round(d['kernels_us']['pre_physics'], 2), r['kernel'], 'frac', round(r['frac'], 3), 'hot post', round(d['l2_hot']['post_us'], 2), 'hot step', round(d['l2_hot']['step_us'], 2))
"
}
run "X=1" ""
run "DEXSTEP_POST_GEOMETRY=single" ""
run "X=1" "--observer 1"
run "X=1" "--workload blindgrasping_4096"
run "X=1" "--workload basetask_4096"
run "X=1" "--envs 65536"
run "X=1" "--envs 262144"
