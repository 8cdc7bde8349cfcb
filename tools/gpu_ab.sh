#!/bin/bash
# usage: tools/gpu_ab.sh "ENV=VAL ..." "ENV=VAL ..." ...  -> the short bench once per environment setting, interleaved
# twice on the same box (box-to-box variance is larger than most single optimisations)
CMD=""
for rep in 1 2; do
for V in "$@"; do
  CMD="$CMD echo \"== $V (rep $rep)\"; (env $V timeout 300 python bench.py --steps 3000 --warmup 200 --no-cpu-baseline 2>&1 | tail -1 | python -c \"
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('value', d['value'], 'ms/step', d['ms_per_step'], 'kernels', d['kernels_us'])
    else: print(l.strip())
\");"
done
done
/usr/local/graft/bin/gpurun --timeout 900 -- "$CMD" 2>&1 | tail -30
