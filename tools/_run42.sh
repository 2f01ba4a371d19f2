timeout 600 python -m pytest tests/test_ops_gpu.py tests/test_gemm_gpu.py -m gpu -x -q 2>&1 | tail -2 > gpurun_out/r2y_pytest.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-kernels > gpurun_out/r2y_bench.json 2> gpurun_out/r2y_bench.err
true
