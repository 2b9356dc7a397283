#!/bin/bash
# round-2 ncu evidence (one GPU): launch list of the default bench command, full captures of the step kernels
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_default_r02.json 2> gpurun_out/bench_default_r02.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r02.csv python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/ncu_launch_r02.log 2>&1
for w in kinematic diffusive dynamic; do
  K=k_step_kinematic; [ $w != kinematic ] && K="k_phase_a|k_phase_b"
  python bench.py --wave $w --steps 20 --warmup 5 --no-cpu --no-grid-rain --no-parity > gpurun_out/bench_${w}_r02.json 2> gpurun_out/bench_${w}_r02.err
  ncu --set full --clock-control none --import-source on -k "regex:$K" -s 6 -c 2 -o gpurun_out/prof_${w}_r02 python bench.py --wave $w --steps 3 --warmup 2 --no-cpu --no-grid-rain --no-parity > gpurun_out/ncu_full_${w}_r02.log 2>&1
  tail -1 gpurun_out/ncu_full_${w}_r02.log
done
cat gpurun_out/bench_default_r02.json | cut -c1-300
