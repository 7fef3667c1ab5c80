#!/bin/bash
# Batch sweep of the full hot path (net -> resize -> nms -> connector) on one B200: frames/s, ms per step (= latency of a
# batch), conv fraction of the sustained bf16 peak. usage: tools/batch_sweep.sh [out.jsonl]
out=${1:-gpurun_out/batch_sweep.jsonl}
: > "$out"
for b in 1 2 4 8 16 32 44; do
    steps=$(( b < 8 ? 200 : 50 ))
    python bench.py --batch $b --steps $steps --warmup 10 --no-cpu >> "$out" 2>> gpurun_out/batch_sweep.err
done
python - "$out" <<'PY'
import json, sys
print("batch  frames/s  e2e frames/s  ms/step  conv_ms  conv TFLOP/s  frac")
for line in open(sys.argv[1]):
    d = json.loads(line)
    r = d["roofline"]
    print("%5d  %8.1f  %12.1f  %7.3f  %7.3f  %12.1f  %.3f" % (d["config"]["batch_per_gpu"], d["value"], d["e2e"]["value"],
          d["ms_per_step"], r["launch_ms"], r["achieved"], r["frac"]))
PY
