timeout 600 python -m pytest tests/test_ops_gpu.py tests/test_train_step_gpu.py -m gpu -x -q 2>&1 | tail -4 > gpurun_out/r1t_pytest.log
timeout 300 python tools/bench_ops.py norm > gpurun_out/r1t_ops.txt 2>&1
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-kernels > gpurun_out/r1t_bench.json 2> gpurun_out/r1t_bench.err
true
