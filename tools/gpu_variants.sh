#!/bin/bash
# usage: tools/gpu_variants.sh "FLAGS1" "FLAGS2" ...   -> rebuilds libdexstep.so on the GPU box with each set of extra
# nvcc flags (experiment macros; extra bench.py arguments in $BENCH_ARGS) and prints the bench kernel times; the default build is restored locally afterwards.
CMD=""
for F in "$@"; do
CMD="$CMD echo \"== [$F]\"; DEXSTEP_NVCC_FLAGS=\"$F\" python -m dexrobot_isaac_b200.build --force > /dev/null && timeout 300 python bench.py --steps 1500 --warmup 100 --no-cpu-baseline $BENCH_ARGS 2>&1 | tail -1 | python -c \"
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('step_us', round(1e3*d['ms_per_step'],2), 'kernels', {k:(round(v,2) if v else v) for k,v in d['kernels_us'].items()})
    else: print(l.strip()[:300])
\";"
done
/usr/local/graft/bin/gpurun --timeout 1500 -- "$CMD" 2>&1 | grep -E "^==|step_us|rror|status"
