for m in 4 8 10 12; do
OPB200_BSTAGES=$m ncu --metrics gpu__time_duration.sum --clock-control none -c 114 --csv --log-file gpurun_out/launches_bs$m.csv python bench.py --stages net --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu_bs$m.log 2>&1
done
ls gpurun_out
