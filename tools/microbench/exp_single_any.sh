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
