python -m pytest tests -m gpu -x -q > gpurun_out/tests.log 2>&1; echo tests=$? > gpurun_out/rc.log
python bench.py --no-cpu --steps 10 --warmup 3 > gpurun_out/bench_full.log 2>&1; echo bench=$? >> gpurun_out/rc.log
tail -3 gpurun_out/tests.log; cat gpurun_out/rc.log; tail -1 gpurun_out/bench_full.log | cut -c1-2400
