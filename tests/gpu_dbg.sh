mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_net_gpu.py -x -q > gpurun_out/tests_net.log 2>&1; echo tests=$?; tail -2 gpurun_out/tests_net.log
run() { timeout 300 python bench.py --stages net --no-cpu --steps 20 --warmup 5 2>&1 | tail -1 | python -c "
import sys, json
r = json.loads(sys.stdin.read()); print('$1', round(r['ms_per_step'],3), round(r['roofline']['achieved'],1), r['clocks']['sm_mhz'])"; }
run staging2
OPB200_CONV_STAGING=1 run staging1
run staging2
OPB200_CONV_STAGING=1 run staging1
for t in 2:128:3:3 6:512:1:1 1:64:3:1; do
  OPB200_NO_GRAPH=1 OPB200_CONV_TRACE=$t timeout 300 python bench.py --stages net --no-cpu --steps 1 --warmup 1 2>&1 | grep -A 6 "conv trace"
done
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 228 --csv --log-file gpurun_out/launches_net_st2.csv python bench.py --stages net --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu3.log 2>&1; echo ncu3=$?
