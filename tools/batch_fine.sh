#!/bin/bash
# Fine batch sweep around the default on one B200 (wave quantisation of the 46x82 layers: 74 CTA pairs x 4 tiles), interleaved
# twice so that clock drift shows. usage: tools/batch_fine.sh "44 47 57"
bs=${1:-"44 47 57"}
echo "batch  frames/s  ms/step  conv_ms  conv TFLOP/s  sm_mhz"
for r in 1 2; do
 for b in $bs; do
  timeout 120 python bench.py --batch $b --steps 60 --warmup 10 --no-cpu --no-engine --no-latency 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); r=d['roofline']
print('%5d  %8.1f  %7.3f  %7.3f  %12.1f  %s' % ($b, d['value'], d['ms_per_step'], r['launch_ms'], r['achieved'], d['clocks']['sm_mhz']))"
 done
done
