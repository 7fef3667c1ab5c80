for b in 44 45 46 47 48 52; do
python bench.py --stages net --no-cpu --steps 20 --warmup 3 --batch $b 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('batch=$b', round(d['value']), d['stages_ms'])"
done
