#!/bin/bash
# lean vs observer-on control step (16 384 envs BlindGrasping), default and persistent geometry
for v in ":--observer 0" ":--observer 1" "DEXSTEP_POST_GEOMETRY=persistent:--observer 0" "DEXSTEP_POST_GEOMETRY=persistent:--observer 1" ":--observer 1 --workload blindgrasping_4096" ":--observer 0 --workload basetask_4096" ":--observer 1 --workload basetask_4096"; do
  envs=${v%%:*}; args=${v#*:}
  env $envs python bench.py --steps 2000 --warmup 100 --no-cpu-baseline $args 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        print('$envs $args', '| step_us', round(d['ms_per_step'] * 1e3, 2), 'post', round(d['kernels_us']['post_physics'], 2), 'pre', round(d['kernels_us']['pre_physics'], 2), 'floor', round(d['kernels_us']['empty_kernel_launch_floor'], 2), r['kernel'], 'hot step', round(d['l2_hot']['step_us'], 2))
"
done
