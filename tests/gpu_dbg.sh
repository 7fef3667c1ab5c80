mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_net_gpu.py -x -q > gpurun_out/tests_net.log 2>&1; echo tests=$?; tail -2 gpurun_out/tests_net.log
run() { timeout 300 python bench.py --stages net --no-cpu --steps 20 --warmup 5 2>&1 | tail -1 | python -c "
import sys, json
r = json.loads(sys.stdin.read()); print('$1', round(r['ms_per_step'],3), round(r['roofline']['achieved'],1), r['clocks']['sm_mhz'])"; }
run warm
OPB200_CONV_DEBUG=8 run nowarm
run warm
OPB200_CONV_DEBUG=8 run nowarm
OPB200_NO_GRAPH=1 OPB200_CONV_TRACE=2:128:3:3 timeout 300 python bench.py --stages net --no-cpu --steps 1 --warmup 1 2>&1 | grep -A 2 "conv trace"
OPB200_CONV_DEBUG=8 OPB200_NO_GRAPH=1 OPB200_CONV_TRACE=2:128:3:3 timeout 300 python bench.py --stages net --no-cpu --steps 1 --warmup 1 2>&1 | grep -A 2 "conv trace"
