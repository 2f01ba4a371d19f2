#!/bin/bash
# Final-build evidence of the round: full GPU suite, smoke(), default bench line (with cpu_baseline), the reference arm,
# the complete-step and sd21 / lion configs, and the ncu launch list of the bench command.
O=gpurun_out/r4u
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/smoke.log
timeout 900 python bench.py > $O/bench_sd15_b8.json 2> $O/bench_sd15_b8.err; echo "bench rc=$?"; cut -c1-700 $O/bench_sd15_b8.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference_arm.json 2> $O/bench_reference_arm.err; echo "ref rc=$?"; cut -c1-400 $O/bench_reference_arm.json
timeout 300 python bench.py --config sd15_full --steps 10 --warmup 3 --no-cpu --no-kernels > $O/bench_sd15_full_step.json 2> $O/bench_full.err; grep -o '"ms_per_step": [0-9.]*' $O/bench_sd15_full_step.json
timeout 300 python bench.py --config sd21 --steps 10 --warmup 3 --no-cpu --no-kernels > $O/bench_sd21_b4.json 2> $O/bench_sd21.err; grep -o '"ms_per_step": [0-9.]*' $O/bench_sd21_b4.json
timeout 300 python bench.py --config lion --steps 20 --warmup 5 --no-cpu > $O/bench_lion.json 2> $O/bench_lion.err; grep -o '"ms_per_step": [0-9.]*' $O/bench_lion.json
timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu --no-kernels > $O/plain.json 2> $O/plain.err || { echo "plain bench failed"; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 5600 --csv --log-file $O/launches_sd15_b8.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-kernels > $O/ncu.log 2>&1; echo "ncu rc=$?"
wc -l $O/launches_sd15_b8.csv; du -sh $O
