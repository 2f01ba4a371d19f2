timeout 600 python -m pytest tests/test_gemm_gpu.py -m gpu -x -q 2>&1 | tail -4 > gpurun_out/r2v_pytest.log
timeout 200 python tools/bench_ops.py conv linear > gpurun_out/r2v_ops.txt 2>&1
SDB_SPLITK=0 timeout 200 python tools/bench_ops.py conv linear > gpurun_out/r2v_ops_nosplit.txt 2>&1
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-kernels > gpurun_out/r2v_bench.json 2> gpurun_out/r2v_bench.err
true
