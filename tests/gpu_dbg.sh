for r in 1 2 3; do for m in 0 1; do
OPB200_CONV_FLAT=$m python bench.py --stages net --no-cpu --steps 40 --warmup 5 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('flat=$m', round(d['value']), d['stages_ms'], d['clocks']['sm_mhz'])"
done; done
