#!/bin/bash
# Scaling on one multi-GPU box: the driver's launch line for bench.py at N = 1, 2, 4 (8) GPUs, then the in-library engine in
# ONE process over the same devices. usage: tools/gpu_scale.sh "1 2 4" [out.jsonl]
ns=${1:-"1 2"}; out=${2:-gpurun_out/scale.jsonl}
: > "$out"
for n in $ns; do
  if [ "$n" = 1 ]; then
    timeout 300 python bench.py --gpus 1 --steps 60 --warmup 10 --no-cpu --no-latency >> "$out" 2>> gpurun_out/scale.err
  else
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
        bench.py --gpus $n --steps 60 --warmup 10 --no-cpu --no-latency >> "$out" 2>> gpurun_out/scale.err
  fi
done
for n in $ns; do
  PYTHONPATH=. timeout 300 python tools/engine_multi_gpu.py $n 60 >> "$out" 2>> gpurun_out/scale.err
done
python - "$out" <<'PY'
import json, sys
for line in open(sys.argv[1]):
    line = line.strip()
    if not line.startswith("{"):
        continue
    d = json.loads(line)
    if d.get("tool"):
        print("engine, one process, %d GPUs: %.0f frames/s" % (d["n_gpus"], d["frames_per_s"]))
    else:
        print("bench.py %d GPUs: value %.0f frames/s, e2e %.0f frames/s" % (d["n_gpus"], d["value"], d["e2e"]["value"]))
PY
