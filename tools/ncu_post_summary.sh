# ncu --set full of the post-processing kernels at the headline shape (one launch each, after warm-up) -> gpurun_out/r02_post_*.ncu-rep
export PYTHONPATH=.
timeout 200 ncu --set full --clock-control none -k regex:"nmsKernel|resize8Stream|pafCandKernel|pafMatchKernel|assembleKernel" -s 35 -c 5 -f -o gpurun_out/r02_post_kernels python bench.py --stages post --steps 2 --warmup 8 --no-cpu > gpurun_out/ncu_post.log 2>&1; echo post=$?
timeout 200 ncu --set full --clock-control none -k regex:resizeMergeStream -s 20 -c 1 -f -o gpurun_out/r02_resize_ms python tools/bench_resize_merge4.py > gpurun_out/ncu_ms.log 2>&1; echo ms=$?
