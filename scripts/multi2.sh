#!/bin/bash
# round-2 multi-GPU lines: scripts/multi2.sh N [coupled_size]
N=${1:-2}; CS=${2:-8192}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
mkdir -p gpurun_out
$TR bench.py --gpus $N --steps 20 --warmup 5 --no-grid-rain > gpurun_out/r2_default_${N}gpu.json 2> gpurun_out/r2_default_${N}gpu.err; python -c "
import json; l=json.load(open('gpurun_out/r2_default_${N}gpu.json')); print('default N=$N ms/step', l['ms_per_step'], 'value %.4g' % l['value'], 'parity ok', l['parity']['ok'], l['parity']['seconds'], 's; strong', l['strong_scaling_dynamic'])" || tail -5 gpurun_out/r2_default_${N}gpu.err
$TR bench.py --gpus $N --coupled --scaling strong --size $CS --steps 60 --warmup 31 > gpurun_out/r2_coupled_${N}gpu.json 2> gpurun_out/r2_coupled_${N}gpu.err; python -c "
import json; l=json.load(open('gpurun_out/r2_coupled_${N}gpu.json')); print('coupled N=$N size $CS ms/step', l['ms_per_step'], 'value %.4g' % l['value'], 'e2e %.4g' % l['e2e']['value'])" || tail -5 gpurun_out/r2_coupled_${N}gpu.err
