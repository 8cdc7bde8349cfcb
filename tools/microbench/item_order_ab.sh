#!/bin/bash
# EXPERIMENT (rejected and reverted, profiles/r02_variants.md): a balanced item-to-warp table for phase A (BlindGrasping) against
# the shipped round-robin assignment.  Expects the two builds in dexrobot_isaac_b200/_exp/ (libdexstep_lpt.so, libdexstep_rr.so);
# swaps them into place, alternating, and leaves the shipped one behind.
run() {
  cp dexrobot_isaac_b200/_exp/libdexstep_$1.so dexrobot_isaac_b200/libdexstep.so
  python bench.py --steps 2000 --warmup 100 --no-cpu-baseline $2 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        print('$1 $2', '| step_us', round(d['ms_per_step'] * 1e3, 2), 'post', round(d['kernels_us']['post_physics'], 2), 'frac', round(r['frac'], 3), 'hot post', round(d['l2_hot']['post_us'], 2), 'hot step', round(d['l2_hot']['step_us'], 2))
"
}
for w in "" "--envs 65536" "--envs 262144" "--workload blindgrasping_4096"; do
  run lpt "$w"
  run rr "$w"
done
run lpt "--single-pass"
run rr "--single-pass"
cp dexrobot_isaac_b200/_exp/libdexstep_lpt.so dexrobot_isaac_b200/libdexstep.so
