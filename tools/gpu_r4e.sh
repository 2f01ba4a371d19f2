#!/bin/bash
# Round-2 GPU call E (2 GPUs): data-parallel correctness (tools/dp_check.py) and the one-graph (direct NCCL, captured
# collectives) vs segmented (torch.distributed collectives between graph segments) step time.
O=gpurun_out/r4e
mkdir -p $O
echo "== 1-GPU tests (time-embedding gradients on the side stream)"; timeout 900 python -m pytest tests/test_train_step_gpu.py tests/test_flags_gpu.py tests/test_train_step_full_gpu.py -q -x > $O/pytest_1gpu.log 2>&1; tail -2 $O/pytest_1gpu.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
echo "== dp_check tiny"; timeout 600 $TR --master-port 29511 tools/dp_check.py --config tiny --batch 2 --steps 3 --out $O/dp_check_tiny.json > $O/dp_check_tiny.log 2>&1; echo "rc=$?"; tail -3 $O/dp_check_tiny.log | cut -c1-600
echo "== dp_check sd15"; timeout 900 $TR --master-port 29512 tools/dp_check.py --config sd15 --batch 2 --steps 3 --out $O/dp_check_sd15.json > $O/dp_check_sd15.log 2>&1; echo "rc=$?"; tail -3 $O/dp_check_sd15.log | cut -c1-900
echo "== bench n2 one graph"; timeout 600 $TR --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu --no-kernels > $O/bench_n2.json 2> $O/bench_n2.err; echo "rc=$?"; cut -c1-250 $O/bench_n2.json; tail -3 $O/bench_n2.err
echo "== bench n2 segments (torch collectives)"; SDB_DP_NCCL=torch timeout 600 $TR --master-port 29514 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu --no-kernels --allow-experiments > $O/bench_n2_seg.json 2> $O/bench_n2_seg.err; echo "rc=$?"; cut -c1-250 $O/bench_n2_seg.json
echo "== bench n1"; timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu --no-kernels > $O/bench_n1.json 2> $O/bench_n1.err; cut -c1-200 $O/bench_n1.json
