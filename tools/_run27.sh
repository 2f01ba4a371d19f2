timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 > gpurun_out/r2d_pytest.log
for dc in 1200 800; do
SDB_DELIVERY_CLKS=$dc timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-kernels > gpurun_out/r2d_bench_$dc.json 2> gpurun_out/r2d_bench.err
done
true
