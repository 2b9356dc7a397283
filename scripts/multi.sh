#!/bin/bash
# multi-GPU checks and bench lines on one box: scripts/multi.sh N
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
mkdir -p gpurun_out
if [ "$N" = "2" ]; then python -m pytest tests/test_gpu_multi.py -q 2>&1 | tail -3; fi
$TR bench.py --gpus $N --steps 20 --warmup 5 --no-grid-rain > gpurun_out/r2_weak_kin_${N}gpu.json 2> gpurun_out/r2_weak_kin_${N}gpu.err; tail -c 1500 gpurun_out/r2_weak_kin_${N}gpu.json; echo
$TR bench.py --gpus $N --steps 20 --warmup 5 --no-grid-rain --no-cpu --scaling strong --wave dynamic --size 16384 > gpurun_out/r2_strong_dyn_${N}gpu.json 2> gpurun_out/r2_strong_dyn_${N}gpu.err; tail -c 600 gpurun_out/r2_strong_dyn_${N}gpu.json; echo
$TR bench.py --gpus $N --steps 20 --warmup 5 --no-grid-rain --no-cpu --no-parity --scaling strong --wave diffusive --size 8192 --with-area > gpurun_out/r2_strong_dif_area_${N}gpu.json 2> gpurun_out/r2_strong_dif_area_${N}gpu.err; tail -c 600 gpurun_out/r2_strong_dif_area_${N}gpu.json; echo
