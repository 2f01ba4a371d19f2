timeout 600 python -m pytest tests/test_ops_gpu.py tests/test_train_step_gpu.py -m gpu -x -q 2>&1 | tail -3 > gpurun_out/r2x_pytest.log
timeout 200 python tools/bench_ops.py norm > gpurun_out/r2x_ops.txt 2>&1
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-kernels > gpurun_out/r2x_bench.json 2> gpurun_out/r2x_bench.err
true
