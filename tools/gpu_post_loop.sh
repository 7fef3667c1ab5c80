python -m pytest tests/test_post_gpu.py tests/test_extractor_gpu.py -x -q > gpurun_out/tests.log 2>&1; echo tests=$? > gpurun_out/rc.log
python bench.py --no-cpu --steps 10 --warmup 3 > gpurun_out/bench_full.log 2>&1; echo bench=$? >> gpurun_out/rc.log
tail -3 gpurun_out/tests.log; cat gpurun_out/rc.log; tail -1 gpurun_out/bench_full.log | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e']['value'], d['stages_ms'])"
