#!/bin/bash
# Everything the round's numbers come from, in one GPU call (one B200):   bash tools/round_profile.sh r02g
# bench lines of every configuration BASELINE.md quotes -> gpurun_out/<tag>_*.json, then (only after the plain runs
# exited 0) the ncu launch list and one --set full capture per hot kernel.  Numbers printed under ncu are never bench values.
tag=${1:-r02g}
o=gpurun_out
set -x
python bench.py --sweep > $o/${tag}_bench.json 2> $o/${tag}_bench.err || exit 1
python bench.py --impl reference --steps 20 --warmup 3 > $o/${tag}_bench_reference.json 2>> $o/${tag}_bench.err
python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $o/${tag}_bench_driver_protocol.json 2>> $o/${tag}_bench.err
for w in blindgrasping_4096 basetask_4096; do
  python bench.py --workload $w --no-cpu-baseline > $o/${tag}_bench_$w.json 2>> $o/${tag}_bench.err
done
python bench.py --observer 1 --no-cpu-baseline > $o/${tag}_bench_observer.json 2>> $o/${tag}_bench.err
python bench.py --obs-norm 1 --no-cpu-baseline > $o/${tag}_bench_obs_norm.json 2>> $o/${tag}_bench.err
python bench.py --nb-hand 37 --no-cpu-baseline > $o/${tag}_bench_nb37.json 2>> $o/${tag}_bench.err
DEXSTEP_EARLY_SIM_LOADS=0 python bench.py --no-cpu-baseline > $o/${tag}_bench_no_early_sim_loads.json 2>> $o/${tag}_bench.err
python bench.py --single-pass --no-cpu-baseline > $o/${tag}_bench_single_pass.json 2>> $o/${tag}_bench.err
python bench.py --tile-kernel --no-cpu-baseline > $o/${tag}_bench_tile_kernel.json 2>> $o/${tag}_bench.err
# ---- ncu: launch list (gpu__time_duration) of the default command, then full captures ----
# (only this library's kernels: the set-up of the bench launches thousands of torch fill / copy kernels first)
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'physics|epilogue|reset_kernel|launch_floor' -c 600 --csv \
    --log-file $o/${tag}_launches.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $o/${tag}_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:post_physics -s 30 -c 1 -o $o/${tag}_post -f \
    python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $o/${tag}_ncu_post.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:pre_physics -s 30 -c 1 -o $o/${tag}_pre -f \
    python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $o/${tag}_ncu_pre.log 2>&1
DEXSTEP_POST_GEOMETRY=single ncu --set full --clock-control none --import-source on -k regex:post_physics -s 30 -c 1 \
    -o $o/${tag}_post_single_pass -f python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $o/${tag}_ncu_post_single_pass.log 2>&1
ls -la $o | tail -30
