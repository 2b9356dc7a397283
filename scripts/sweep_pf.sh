#!/bin/bash
# Look-ahead distance sweep (TF_B200_PF_ROWS; "auto" = from the grid width).
#   SIZE=8192 WAVE=kinematic PFS="auto 0 8 16 24 32" ./scripts/sweep_pf.sh
SIZE=${SIZE:-8192}
for pf in ${PFS:-auto 0 8 16 24 32 48}; do
  if [ "$pf" = auto ]; then unset TF_B200_PF_ROWS; else export TF_B200_PF_ROWS=$pf; fi
  out=$(python bench.py --size $SIZE --steps 10 --warmup 3 --no-cpu ${WAVE:+--wave $WAVE} 2>/dev/null)
  echo "size $SIZE wave ${WAVE:-kinematic} pf_rows $pf: $(echo "$out" | python -c "import json,sys; j=json.loads(sys.stdin.read()); print('%.4f ms  frac %.3f  clk %s' % (j['roofline']['kernel_ms'], j['roofline']['frac'], j['clocks']['sm_mhz']))")"
done
