python -m pytest tests/test_net_gpu.py -x -q 2>&1 | tail -2
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:firstConv -c 3 --csv --log-file gpurun_out/fc.csv python bench.py --stages net --steps 1 --warmup 1 --no-cpu > /dev/null 2>&1
grep firstConv gpurun_out/fc.csv | awk -F'","' '{print $NF}'
