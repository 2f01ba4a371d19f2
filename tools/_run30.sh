timeout 600 python -m pytest tests/test_gemm_gpu.py -m gpu -x -q 2>&1 | tail -3 > gpurun_out/r2j_pytest.log
timeout 300 python tools/bench_ops.py conv linear > gpurun_out/r2j_ops.txt 2>&1
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-kernels > gpurun_out/r2j_bench.json 2> gpurun_out/r2j_bench.err
true
