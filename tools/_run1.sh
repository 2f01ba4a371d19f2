set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r1e_pytest.log
python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r1e_bench.json 2> gpurun_out/r1e_bench.err
python tools/bench_ops.py norm > gpurun_out/r1e_ops_norm.txt 2>&1
ncu --set full --clock-control none --import-source on -k regex:lion8 -c 1 -s 2 -o gpurun_out/r1e_lion -f python tools/prof_one.py lion > gpurun_out/ncu_lion.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gn_ -c 12 -o gpurun_out/r1e_gn -f python tools/prof_one.py gn > gpurun_out/ncu_gn.log 2>&1
true
