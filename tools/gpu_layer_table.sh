#!/bin/bash
# Per-layer table of one net forward at a given batch: ncu launch list + plan dump -> tools/layer_table.py.
# usage: tools/gpu_layer_table.sh <batch> <tag>   (writes gpurun_out/launches_<tag>.csv, gpurun_out/layers_<tag>.txt)
b=${1:-1}; tag=${2:-b$b}
python - <<'PY' > gpurun_out/plan.txt
from openpose_b200 import net, modelgen
proto, cm, _, _ = modelgen.build_files("body_25", 1234)
print(net.describe(proto, cm))
PY
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$tag.csv \
    python bench.py --stages net --batch $b --steps 1 --warmup 1 --no-cpu --no-latency --no-engine > gpurun_out/ncu_$tag.log 2>&1
python tools/layer_table.py gpurun_out/launches_$tag.csv gpurun_out/plan.txt $b 368 656 200 > gpurun_out/layers_$tag.txt 2>&1
head -3 gpurun_out/layers_$tag.txt
