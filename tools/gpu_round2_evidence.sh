# Round-2 evidence capture on one B200 (never a bench value under ncu): launch lists, DRAM traffic of the net, the
# per-kernel tools, one --set full capture of a trunk-stage conv launch.
mkdir -p gpurun_out
export PYTHONPATH=.
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/r02_full_launches_b44.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-latency --no-engine > gpurun_out/ncu1.log 2>&1; echo ncu1=$? > gpurun_out/rc.log
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 228 --csv --log-file gpurun_out/r02_net_launches_dram_b44.csv python bench.py --stages net --steps 1 --warmup 1 --no-cpu --no-latency --no-engine > gpurun_out/ncu2.log 2>&1; echo ncu2=$? >> gpurun_out/rc.log
timeout 300 ncu --set full --clock-control none --import-source on -k regex:convIgemm -s 170 -c 1 -f -o gpurun_out/r02_conv_mid python bench.py --stages net --steps 1 --warmup 1 --no-cpu --no-latency --no-engine > gpurun_out/ncu3.log 2>&1; echo ncu3=$? >> gpurun_out/rc.log
timeout 300 python tools/bench_kernels.py > gpurun_out/r02_post_kernels.txt 2>&1; echo kernels=$? >> gpurun_out/rc.log
timeout 300 python tools/bench_models.py 10 > gpurun_out/r02_models_net.jsonl 2>&1; echo models=$? >> gpurun_out/rc.log
for p in 8 32 100; do timeout 200 python bench.py --stages post --no-cpu --people $p --steps 20 --warmup 3 2>&1 | tail -1 > gpurun_out/r02_bench_post_p$p.json; done
timeout 100 python tools/bench_nms_dense.py 44 > gpurun_out/r02_nms_dense.txt 2>&1
cat gpurun_out/rc.log
