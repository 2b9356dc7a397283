#!/bin/bash
# final round-2 check on one GPU: tests, smoke, the default bench line and the three waves
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_default_r02.json 2> gpurun_out/bench_default_r02.err; tail -c 400 gpurun_out/bench_default_r02.json; echo
for w in kinematic diffusive dynamic; do
  python bench.py --wave $w --steps 20 --warmup 5 --no-cpu --no-grid-rain --no-parity > gpurun_out/bench_${w}_r02.json 2> gpurun_out/bench_${w}_r02.err
  python -c "import json; l=json.load(open('gpurun_out/bench_${w}_r02.json')); r=l['roofline']; print('$w', round(l['ms_per_step'],3), 'ms frac', round(r['frac'],3), 'frac_actual', round(r['frac_actual'],3))"
done
python bench.py --impl reference --steps 10 --warmup 2 > gpurun_out/bench_reference_r02.json 2>/dev/null; cut -c1-200 gpurun_out/bench_reference_r02.json
