#!/bin/bash
# Round-2 GPU call B: new default GEMM (lean producer + two epilogue sets + deterministic slab split-K): full suite,
# bench, red-vs-slab comparison, ncu launch list of one step.
O=gpurun_out/r4b
mkdir -p $O
echo "== gpu tests"; timeout 1500 python -m pytest tests -q -m gpu -s > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log; tail -4 $O/pytest_gpu.log
grep -A3 "SD-1.5 B=2, 2 steps" $O/pytest_gpu.log
echo "== bench sd15"; timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench_sd15.json 2> $O/bench_sd15.err; echo "rc=$?"; cut -c1-300 $O/bench_sd15.json
echo "== bench sd15 (red split-K)"; SDB_EXPERIMENTS=1 SDB_SPLITK_RED=1 timeout 300 python bench.py --allow-experiments --steps 20 --warmup 5 --no-cpu --no-kernels > $O/bench_sd15_red.json 2> $O/bench_sd15_red.err; cut -c1-200 $O/bench_sd15_red.json
echo "== bench sd15 again (slab)"; timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu --no-kernels > $O/bench_sd15_b.json 2> $O/bench_sd15_b.err; cut -c1-200 $O/bench_sd15_b.json
echo "== smoke"; timeout 300 python __graft_entry__.py --smoke > $O/smoke.log 2>&1; tail -2 $O/smoke.log
echo "== ncu launch list"; timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 5600 --csv --log-file $O/launches_sd15_b8.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-kernels > $O/ncu.log 2>&1; echo "rc=$?"; wc -l $O/launches_sd15_b8.csv
echo done
