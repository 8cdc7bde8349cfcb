#!/bin/bash
# usage: tools/gpu_round.sh TAG [sweep]
# Full evidence pass on one B200: gpu parity tests, smoke, bench (both arms), ncu launch list of the bench command,
# ncu --set full of the post- and pre-physics kernels.  Every ncu pass runs only after the same command exited 0
# without ncu.  Outputs land in gpurun_out/ (TAG names them); summarise into profiles/ with tools/ncu_summary.py.
TAG=${1:-x}
SWEEP=${2:+--sweep}
B="python bench.py --steps 200 --warmup 20 --no-cpu-baseline"
/usr/local/graft/bin/gpurun --timeout 2400 -- "
(timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo \"pytest rc=\$?\" >> gpurun_out/pytest_gpu_$TAG.log); tail -8 gpurun_out/pytest_gpu_$TAG.log
(timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke_$TAG.log 2>&1; echo \"smoke rc=\$?\" >> gpurun_out/smoke_$TAG.log); tail -4 gpurun_out/smoke_$TAG.log
(timeout 900 python bench.py $SWEEP > gpurun_out/bench_$TAG.log 2>&1; echo \"bench rc=\$?\" >> gpurun_out/bench_$TAG.log); tail -2 gpurun_out/bench_$TAG.log | cut -c1-3000
(timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/bench_ref_$TAG.log 2>&1; echo \"ref rc=\$?\" >> gpurun_out/bench_ref_$TAG.log); tail -2 gpurun_out/bench_ref_$TAG.log | cut -c1-1200
for W in basetask_4096 blindgrasping_4096; do (timeout 300 python bench.py --workload \$W --no-cpu-baseline > gpurun_out/bench_\${W}_$TAG.log 2>&1); tail -1 gpurun_out/bench_\${W}_$TAG.log | cut -c1-1500; done
$B > gpurun_out/plain_$TAG.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_$TAG.csv $B > gpurun_out/ncu_${TAG}_list.log 2>&1
$B > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:post_physics -s 20 -c 1 -f -o gpurun_out/prof_${TAG}_post $B > gpurun_out/ncu_${TAG}_post.log 2>&1
$B > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:pre_physics -s 20 -c 1 -f -o gpurun_out/prof_${TAG}_pre $B > gpurun_out/ncu_${TAG}_pre.log 2>&1
ls -la gpurun_out | tail -20
" 2>&1 | tail -60
