timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r1k_pytest.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r1k_bench.json 2> gpurun_out/r1k_bench.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file gpurun_out/r1k_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu --no-kernels > gpurun_out/r1k_ncu.log 2>&1
true
