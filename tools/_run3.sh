set -x
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r1g_pytest.log
timeout 300 python tools/bench_ops.py attn > gpurun_out/r1g_ops_attn.txt 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r1g_bench.json 2> gpurun_out/r1g_bench.err
true
