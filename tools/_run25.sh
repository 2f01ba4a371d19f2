timeout 300 python -m pytest tests/test_ops_gpu.py -m gpu -q -k attention 2>&1 | tail -3 > gpurun_out/r2b_pytest.log
timeout 300 python tools/bench_ops.py attn > gpurun_out/r2b_ops.txt 2>&1
for dc in 3000 2000 1200; do
SDB_DELIVERY_CLKS=$dc timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-kernels > gpurun_out/r2b_bench_$dc.json 2> gpurun_out/r2b_bench.err
done
true
