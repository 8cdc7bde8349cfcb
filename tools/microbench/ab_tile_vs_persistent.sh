#!/bin/bash
# tile kernel (round 1, table driven) against the persistent kernel, alternating, three rounds per size (BlindGrasping unless noted)
run() {
  python bench.py --steps 2000 --warmup 100 --no-cpu-baseline $1 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        print('$1', '| step_us', round(d['ms_per_step'] * 1e3, 2), 'post', round(d['kernels_us']['post_physics'], 2), 'hot step', round(d['l2_hot']['step_us'], 2))
"
}
for i in 1 2 3; do
  run ""
  run "--tile-kernel"
done
for w in "--workload blindgrasping_4096" "--workload basetask_4096" "--envs 8192" "--envs 32768" "--envs 65536"; do
  run "$w"
  run "$w --tile-kernel"
done
