timeout 300 python -m pytest tests/test_ops_gpu.py -m gpu -q -k "layernorm or ln or norm" 2>&1 | tail -2 > gpurun_out/r2q_pytest.log
timeout 200 python tools/bench_ops.py norm > gpurun_out/r2q_ops.txt 2>&1
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-kernels > gpurun_out/r2q_bench.json 2> gpurun_out/r2q_bench.err
true
