python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --no-cpu --steps 20 --warmup 3 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['e2e']['value']), d['stages_ms'])"
