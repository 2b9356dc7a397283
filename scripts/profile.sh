#!/bin/bash
# ncu evidence for the bench command (run under gpurun, one GPU).
#   1. launch list of `python bench.py` (per-launch device time, serialised)
#   2. full capture of the dominant kernel
set -x
TAG=${TAG:-r01}
mkdir -p gpurun_out
python bench.py $ARGS > gpurun_out/bench_plain_$TAG.json 2> gpurun_out/bench_plain_$TAG.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/launches_$TAG.csv python bench.py $ARGS > gpurun_out/ncu_launch_$TAG.log 2>&1
python bench.py $ARGS --steps 3 --warmup 2 --no-cpu > gpurun_out/plain2_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:${KERNEL:-k_step_kinematic} -s ${SKIP:-3} -c ${COUNT:-2} \
    -o gpurun_out/prof_${TAG} python bench.py $ARGS --steps 3 --warmup 2 --no-cpu > gpurun_out/ncu_full_$TAG.log 2>&1
tail -2 gpurun_out/ncu_full_$TAG.log
cat gpurun_out/bench_plain_$TAG.json
