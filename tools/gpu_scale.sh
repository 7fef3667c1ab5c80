# 1 -> 8 GPU scaling of the bench line on one box, with the driver's launch line
mkdir -p gpurun_out; : > gpurun_out/scale.jsonl
python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu 2>/dev/null | tail -1 >> gpurun_out/scale.jsonl
for n in ${SCALE_NS:-2 4 8}; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29520+n)) bench.py --gpus $n --steps 10 --warmup 3 --no-cpu 2>/dev/null | tail -1 >> gpurun_out/scale.jsonl
done
python -c "
import json
for l in open('gpurun_out/scale.jsonl'):
    r = json.loads(l); print(r['n_gpus'], round(r['value'],1), round(r['e2e']['value'],1), r['clocks']['sm_mhz'])"
