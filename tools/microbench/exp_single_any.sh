#!/bin/bash
# EXPERIMENT (reverted, profiles/r02_variants.md): the single-pass kernel as a multi-wave grid at large sizes.  Needs the one-line
# patch that lifts its residency cap when EXP_SINGLE_ANY is set (launch_post2_single: ok = tiles + fused > 2 * SMs); without it
# the runs below fall back to the persistent kernel.
run() {
  env $1 python bench.py --steps 1000 --warmup 50 --no-cpu-baseline $2 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        print('$1 $2', '| step_us', round(d['ms_per_step'] * 1e3, 2), 'post', round(d['kernels_us']['post_physics'], 2), r['kernel'], 'frac', round(r['frac'], 3))
"
}
for w in "--envs 32768" "--envs 65536" "--envs 262144"; do
  run "EXP_SINGLE_ANY=1 DEXSTEP_POST_GEOMETRY=single" "$w"
done
