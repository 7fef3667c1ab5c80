#!/bin/bash
# A/B of the small-batch conv tiling (OPB200_CONV_SMALL) on ONE box: full hot path, batches 1..8 and 44.
echo "batch  small  frames/s  e2e frames/s  ms/step  conv_ms  conv TFLOP/s"
for b in 1 2 4 8 44; do
  for s in 0 1; do
    steps=$(( b < 8 ? 200 : 50 ))
    OPB200_CONV_SMALL=$s timeout 120 python bench.py --batch $b --steps $steps --warmup 10 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); r=d['roofline']
print('%5d  %5s  %8.1f  %12.1f  %7.3f  %7.3f  %12.1f' % ($b, '$s', d['value'], d['e2e']['value'], d['ms_per_step'], r['launch_ms'], r['achieved']))"
  done
done
