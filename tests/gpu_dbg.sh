python -m pytest tests/test_net_gpu.py -x -q 2>&1 | tail -2
for m in 0 1; do
OPB200_CONV_SERPENTINE=$m python bench.py --stages net --no-cpu --steps 10 --warmup 3 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('serp=$m', d['value'], d['stages_ms'])"
OPB200_CONV_SERPENTINE=$m ncu --metrics gpu__time_duration.sum,lts__t_sector_hit_rate.pct,dram__bytes_read.sum --clock-control none -c 114 --csv --log-file gpurun_out/launches_serp$m.csv python bench.py --stages net --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu_serp$m.log 2>&1
done
