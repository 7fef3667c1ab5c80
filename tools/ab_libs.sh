#!/bin/bash
# Same-box A/B of two builds of libopb200.so: alternates `prev` (openpose_b200/lib/libopb200_prev.so) and `new` at several batches.
# usage: tools/ab_libs.sh [rounds] [batches...]
rounds=${1:-2}; shift; batches=${@:-"1 4 16 44"}
echo "batch  lib   frames/s  ms/step  conv_ms  conv TFLOP/s  sm_mhz"
for r in $(seq $rounds); do
 for b in $batches; do
  for lib in prev new; do
    if [ $lib = prev ]; then export OPB200_LIB_PATH=$PWD/openpose_b200/lib/libopb200_prev.so; else unset OPB200_LIB_PATH; fi
    steps=$(( b < 8 ? 200 : 40 ))
    timeout 120 python bench.py --batch $b --steps $steps --warmup 10 --no-cpu --no-engine --no-latency 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); r=d['roofline']
print('%5d  %-4s  %8.1f  %7.3f  %7.3f  %12.1f  %s' % ($b, '$lib', d['value'], d['ms_per_step'], r['launch_ms'], r['achieved'], d['clocks']['sm_mhz']))"
  done
 done
done
