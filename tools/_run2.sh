set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r1f_pytest.log
python tools/bench_ops.py conv linear > gpurun_out/r1f_ops_base.txt 2>&1
for d in 1 2 4 3 6; do SDB_GEMM_DBG=$d python tools/bench_ops.py conv > gpurun_out/r1f_ops_dbg$d.txt 2>&1; done
for bn in 80 128 144 192 256; do SDB_BN=$bn python tools/bench_ops.py conv > gpurun_out/r1f_ops_bn$bn.txt 2>&1; done
python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r1f_bench.json 2> gpurun_out/r1f_bench.err
ncu --set full --clock-control none --import-source on -k regex:lion8 -c 1 -s 2 -o gpurun_out/r1f_lion -f python tools/prof_one.py lion > gpurun_out/ncu_lion.log 2>&1
true
