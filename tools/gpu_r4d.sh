#!/bin/bash
# Round-2 GPU call D: wgrad tile order (m tile / tap fastest) and 256-row wgrad tiles everywhere (experiment).
O=gpurun_out/r4d
mkdir -p $O
timeout 300 python -m pytest tests/test_gemm_gpu.py -q -x > $O/pytest_gemm.log 2>&1; tail -2 $O/pytest_gemm.log
timeout 300 python tools/bench_ops.py conv linear > $O/ops_default.txt 2>&1
SDB_EXPERIMENTS=1 SDB_WGRAD_MT2=1 timeout 300 python tools/bench_ops.py conv linear > $O/ops_wgrad_mt2.txt 2>&1
grep wgrad $O/ops_default.txt | awk '{print $1,$2,$3,$4,$5}' > /tmp/a; grep wgrad $O/ops_wgrad_mt2.txt | awk '{print $5}' > /tmp/b; paste /tmp/a /tmp/b
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu --no-kernels > $O/bench_sd15.json 2> $O/bench_sd15.err; cut -c1-200 $O/bench_sd15.json
SDB_EXPERIMENTS=1 SDB_WGRAD_MT2=1 timeout 300 python bench.py --allow-experiments --steps 20 --warmup 5 --no-cpu --no-kernels > $O/bench_sd15_mt2.json 2> $O/bench_sd15_mt2.err; cut -c1-200 $O/bench_sd15_mt2.json
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu --no-kernels > $O/bench_sd15_b.json 2> $O/bench_sd15_b.err; cut -c1-200 $O/bench_sd15_b.json
