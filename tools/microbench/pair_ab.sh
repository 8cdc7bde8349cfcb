#!/bin/bash
# pair geometry (two tiles per pass, one CTA per SM) against the persistent kernel
run() {
  env $1 python bench.py --steps 2000 --warmup 100 --no-cpu-baseline $2 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        print('$1 $2', '| step_us', round(d['ms_per_step'] * 1e3, 2), 'post', round(d['kernels_us']['post_physics'], 2), r['kernel'], 'frac', round(r['frac'], 3), 'hot post', round(d['l2_hot']['post_us'], 2), 'hot step', round(d['l2_hot']['step_us'], 2))
"
}
for w in "" "--envs 65536" "--envs 262144" "--workload blindgrasping_4096" "--envs 8192" "--envs 32768"; do
  run "DEXSTEP_POST_GEOMETRY=pair" "$w"
  run "DEXSTEP_POST_GEOMETRY=nopair" "$w"
done
