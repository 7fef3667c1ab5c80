python -m pytest tests/test_post_gpu.py tests/test_extractor_gpu.py tests/test_shim_gpu.py -x -q 2>&1 | tail -3
PYTHONPATH=. python tests/bench_kernels.py 2>&1 | grep -E "4 scales|resize x8"
