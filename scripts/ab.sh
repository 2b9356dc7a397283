#!/bin/bash
# A/B of the step kernels at one size: scripts/ab.sh SIZE "ENVPREFIX1" "ENVPREFIX2" ...
SIZE=${1:-8192}; shift
for w in kinematic diffusive dynamic; do
  for e in "$@"; do
    env $e python bench.py --wave $w --size $SIZE --no-cpu --no-grid-rain --no-parity --steps 20 --warmup 5 2> gpurun_out/ab.err | python -c "import sys,json; l=json.loads(sys.stdin.read()); print('$w [$e]', round(l['ms_per_step'],4), 'ms/step', l['roofline']['layout_bytes_per_cell'], 'B/cell', l['clocks']['sm_mhz'], 'MHz')" || tail -5 gpurun_out/ab.err
  done
done
