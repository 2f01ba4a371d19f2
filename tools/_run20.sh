timeout 600 python -m pytest tests/test_ops_gpu.py -m gpu -x -q -k attention 2>&1 | tail -4 > gpurun_out/r1x_pytest.log
timeout 300 python tools/bench_ops.py attn > gpurun_out/r1x_ops.txt 2>&1
SDB_ATTN_PBUF=2 timeout 300 python tools/bench_ops.py attn > gpurun_out/r1x_ops_pbuf2.txt 2>&1
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-kernels > gpurun_out/r1x_bench.json 2> gpurun_out/r1x_bench.err
true
