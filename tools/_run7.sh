timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -12 > gpurun_out/r1j_pytest.log
timeout 300 python tools/bench_ops.py conv linear > gpurun_out/r1j_ops.txt 2>&1
SDB_MT=1 timeout 300 python tools/bench_ops.py conv linear > gpurun_out/r1j_ops_mt1.txt 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-kernels > gpurun_out/r1j_bench.json 2> gpurun_out/r1j_bench.err
true
