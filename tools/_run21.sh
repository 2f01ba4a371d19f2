timeout 300 python -m pytest tests/test_ops_gpu.py -m gpu -q -k attention 2>&1 | tail -25 > gpurun_out/r1y_pytest.log
timeout 300 python tools/bench_ops.py attn > gpurun_out/r1y_ops.txt 2>&1
SDB_ATTN_ALIAS=1 timeout 300 python -m pytest tests/test_ops_gpu.py -m gpu -q -k attention 2>&1 | tail -5 > gpurun_out/r1y_pytest_alias.log
SDB_ATTN_ALIAS=1 timeout 300 python tools/bench_ops.py attn > gpurun_out/r1y_ops_alias.txt 2>&1
true
