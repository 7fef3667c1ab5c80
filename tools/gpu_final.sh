# Round-end capture: tests, smoke, both bench arms, ncu launch lists (never a bench value under ncu).
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/tests.log 2>&1; echo tests=$? > gpurun_out/rc.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo smoke=$? >> gpurun_out/rc.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/ref.log 2>&1; echo ref=$? >> gpurun_out/rc.log
python bench.py > gpurun_out/bench.log 2>&1; echo bench=$? >> gpurun_out/rc.log
python bench.py --stages post --no-cpu > gpurun_out/bench_post.log 2>&1; echo post=$? >> gpurun_out/rc.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu.log 2>&1; echo ncu=$? >> gpurun_out/rc.log
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 228 --csv --log-file gpurun_out/launches_net_dram.csv python bench.py --stages net --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu2.log 2>&1; echo ncu2=$? >> gpurun_out/rc.log
tail -3 gpurun_out/tests.log; cat gpurun_out/rc.log; tail -1 gpurun_out/bench.log | cut -c1-1800
PYTHONPATH=. python tools/bench_kernels.py > gpurun_out/post_kernels.log 2>&1; echo kernels=$? >> gpurun_out/rc.log
PYTHONPATH=. python tools/bench_models.py 10 > gpurun_out/models.log 2>&1; echo models=$? >> gpurun_out/rc.log
for p in 8 32 100; do python bench.py --stages post --no-cpu --people $p --steps 20 --warmup 3 2>&1 | tail -1 > gpurun_out/bench_post_p$p.log; done
cat gpurun_out/rc.log
