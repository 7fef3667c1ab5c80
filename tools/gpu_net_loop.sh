python -m pytest tests/test_net_gpu.py -x -q > gpurun_out/tests.log 2>&1; echo tests=$? > gpurun_out/rc.log
python bench.py --stages net --no-cpu --steps 10 --warmup 3 > gpurun_out/bench_net.log 2>&1; echo bench=$? >> gpurun_out/rc.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_net.csv python bench.py --stages net --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu.log 2>&1; echo ncu=$? >> gpurun_out/rc.log
tail -3 gpurun_out/tests.log; cat gpurun_out/rc.log; tail -1 gpurun_out/bench_net.log | cut -c1-1800
