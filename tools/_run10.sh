timeout 600 python -m pytest tests/test_train_step_gpu.py tests/test_golden_gpu.py -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r1l_pytest.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-kernels > gpurun_out/r1l_bench.json 2> gpurun_out/r1l_bench.err
SDB_SIDE_STREAM=0 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-kernels > gpurun_out/r1l_bench_noside.json 2> gpurun_out/r1l_bench_noside.err
true
