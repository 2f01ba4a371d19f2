#!/bin/bash
# Round-2 GPU call A: full GPU test-suite, full-size parity tests, bench lines (sd15 / lion / sd21 / reference arm),
# and the GEMM experiment builds (-DSDB_GEMM_FAST_PRODUCER / -DSDB_GEMM_EPI2): parity + timing.
O=gpurun_out/r4a
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
nproc > $O/nproc.txt; free -g >> $O/nproc.txt
echo "== quick gpu tests"; timeout 900 python -m pytest tests -q -m gpu --deselect tests/test_train_step_full_gpu.py -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log; tail -5 $O/pytest_gpu.log
echo "== full-size parity"; timeout 1500 python -m pytest tests/test_train_step_full_gpu.py -q -s > $O/pytest_full.log 2>&1; echo "rc=$?" >> $O/pytest_full.log; tail -5 $O/pytest_full.log
echo "== bench sd15"; timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench_sd15.json 2> $O/bench_sd15.err; echo "rc=$?"; cut -c1-400 $O/bench_sd15.json
echo "== bench lion"; timeout 300 python bench.py --config lion --steps 20 --warmup 5 > $O/bench_lion.json 2> $O/bench_lion.err; echo "rc=$?"; cut -c1-300 $O/bench_lion.json
echo "== bench sd21"; timeout 600 python bench.py --config sd21 --steps 10 --warmup 3 --no-cpu > $O/bench_sd21.json 2> $O/bench_sd21.err; echo "rc=$?"; cut -c1-300 $O/bench_sd21.json
echo "== reference arm"; timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "rc=$?"; cut -c1-300 $O/bench_ref.json
echo "== bench_ops default"; timeout 300 python tools/bench_ops.py conv linear > $O/ops_default.txt 2>&1
for tag in fastprod epi2 fp_epi2; do
  export SDB_EXPERIMENTS=1 SDB_LIB_PATH=$PWD/stable_diffusion_training_b200/libsdb200_$tag.so
  echo "== variant $tag"
  timeout 300 python -m pytest tests/test_gemm_gpu.py -q -x > $O/pytest_gemm_$tag.log 2>&1; echo "rc=$?" >> $O/pytest_gemm_$tag.log; tail -2 $O/pytest_gemm_$tag.log
  timeout 300 python tools/bench_ops.py conv linear > $O/ops_$tag.txt 2>&1
  timeout 300 python bench.py --allow-experiments --steps 10 --warmup 3 --no-cpu --no-kernels > $O/bench_$tag.json 2> $O/bench_$tag.err; cut -c1-200 $O/bench_$tag.json
  unset SDB_EXPERIMENTS SDB_LIB_PATH
done
timeout 200 python tools/bench_ops.py attn norm > $O/ops_attn_norm.txt 2>&1
echo done
