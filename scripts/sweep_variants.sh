#!/bin/bash
# Time the kinematic step kernel for each built library variant
# (topoflow_old_b200/libtopoflow_b200_*.so); prints "variant kernel_ms frac".
SIZE=${SIZE:-8192}
for lib in topoflow_old_b200/libtopoflow_b200*.so; do
  name=$(basename $lib .so)
  out=$(TF_B200_LIB=$PWD/$lib python bench.py --size $SIZE --steps ${STEPS:-10} --warmup 3 --no-cpu ${WAVE:+--wave $WAVE} 2>/dev/null)
  echo "$name $(echo "$out" | python -c "import json,sys; j=json.loads(sys.stdin.read()); print('%.4f ms  frac %.3f  e2e %.3e' % (j['roofline']['kernel_ms'], j['roofline']['frac'], j['e2e']['value']))")"
done
