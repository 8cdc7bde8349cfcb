export DEXSTEP_POST_GEOMETRY=persistent
timeout 900 python -m pytest tests/test_cuda_v2.py -m gpu -q -x -k "bitwise or graph_replay or is_the_kernel" 2>&1 | tail -15
for v in "DEXSTEP_POST_TMA=1" "DEXSTEP_POST_TMA=0"; do
  env $v python bench.py --steps 2000 --warmup 100 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        print('$v', '| step_us', round(d['ms_per_step'] * 1e3, 2), 'post', round(d['kernels_us']['post_physics'], 2), r['kernel'], 'frac', round(r['frac'],3), 'hot post', round(d['l2_hot']['post_us'], 2), 'hot step', round(d['l2_hot']['step_us'], 2))
"
done
