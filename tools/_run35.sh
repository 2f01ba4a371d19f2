timeout 300 python -m pytest tests/test_gemm_gpu.py -m gpu -x -q 2>&1 | tail -2 > gpurun_out/r2o_pytest.log
timeout 200 python tools/bench_ops.py conv linear > gpurun_out/r2o_ops.txt 2>&1
SDB_CLUSTER=0 timeout 200 python tools/bench_ops.py conv linear > gpurun_out/r2o_ops_nocl.txt 2>&1
SDB_MT=1 SDB_BN=160 timeout 200 python tools/bench_ops.py conv linear > gpurun_out/r2o_ops_160.txt 2>&1
SDB_MT=1 SDB_BN=160 SDB_CLUSTER=0 timeout 200 python tools/bench_ops.py conv linear > gpurun_out/r2o_ops_160_nocl.txt 2>&1
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-kernels > gpurun_out/r2o_bench.json 2> gpurun_out/r2o_bench.err
true
