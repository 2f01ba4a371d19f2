timeout 600 python -m pytest tests/test_ops_gpu.py tests/test_train_step_gpu.py -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r1h_pytest.log
timeout 300 python tools/bench_ops.py attn > gpurun_out/r1h_ops_attn.txt 2>&1
true
