#!/bin/bash
# usage: tools/gpu_quick.sh TAG [envs]  -> gpu parity tests, a short bench (no CPU leg), the per-CTA phase timeline
TAG=${1:-q}
N=${2:-16384}
/usr/local/graft/bin/gpurun --timeout 1200 -- "
(timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo \"pytest rc=\$?\" >> gpurun_out/pytest_gpu_$TAG.log); tail -6 gpurun_out/pytest_gpu_$TAG.log
(timeout 300 python bench.py --steps 2000 --warmup 100 --no-cpu-baseline > gpurun_out/bench_$TAG.log 2>&1; echo \"bench rc=\$?\" >> gpurun_out/bench_$TAG.log); tail -2 gpurun_out/bench_$TAG.log | python -c \"
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('value', d['value'], 'ms/step', d['ms_per_step'], 'kernels', d['kernels_us'], 'frac', d['roofline']['frac'], 'e2e', d['e2e']['value'])
    else: print(l.strip())
\"
(timeout 300 python bench.py --steps 600 --warmup 50 --no-cpu-baseline --envs 131072 > gpurun_out/bench_${TAG}_128k.log 2>&1); tail -1 gpurun_out/bench_${TAG}_128k.log | python -c \"
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('128k: value', d['value'], 'kernels', d['kernels_us'], 'frac', d['roofline']['frac'])
\"
(timeout 300 python tools/timeline.py $N > gpurun_out/timeline_$TAG.log 2>&1); tail -25 gpurun_out/timeline_$TAG.log
" 2>&1 | tail -50
