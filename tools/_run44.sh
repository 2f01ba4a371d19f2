timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 > gpurun_out/r3a_pytest.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-kernels > gpurun_out/r3a_bench.json 2> gpurun_out/r3a_bench.err
true
