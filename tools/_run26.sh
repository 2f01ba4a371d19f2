timeout 300 python -m pytest tests/test_ops_gpu.py -m gpu -q -k attention 2>&1 | tail -15 > gpurun_out/r2c_pytest.log
timeout 300 python tools/bench_ops.py attn > gpurun_out/r2c_ops.txt 2>&1
true
