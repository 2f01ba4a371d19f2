#!/bin/bash
# Round-2 GPU call F (2 GPUs): peer-memory (copy-engine) gradient all-reduce vs NCCL: correctness + step time.
O=gpurun_out/r4f
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
echo "== colsum / ops tests (1 GPU)"; timeout 600 python -m pytest tests/test_ops_gpu.py tests/test_train_step_gpu.py -q -x > $O/pytest_1gpu.log 2>&1; tail -2 $O/pytest_1gpu.log
echo "== dp_check tiny (peer)"; timeout 600 $TR --master-port 29511 tools/dp_check.py --config tiny --batch 2 --steps 3 --out $O/dp_check_tiny.json > $O/dp_check_tiny.log 2>&1; echo "rc=$?"; tail -2 $O/dp_check_tiny.log | cut -c1-400
echo "== dp_check sd15 (peer)"; timeout 900 $TR --master-port 29512 tools/dp_check.py --config sd15 --batch 2 --steps 3 --out $O/dp_check_sd15.json > $O/dp_check_sd15.log 2>&1; echo "rc=$?"; tail -2 $O/dp_check_sd15.log | cut -c1-400
echo "== bench n2 peer"; timeout 600 $TR --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu --no-kernels > $O/bench_n2_peer.json 2> $O/bench_n2_peer.err; echo "rc=$?"; cut -c1-200 $O/bench_n2_peer.json; tail -2 $O/bench_n2_peer.err | cut -c1-300
echo "== bench n2 nccl"; SDB_DP_ALLREDUCE=nccl timeout 600 $TR --master-port 29514 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu --no-kernels --allow-experiments > $O/bench_n2_nccl.json 2> $O/bench_n2_nccl.err; echo "rc=$?"; cut -c1-200 $O/bench_n2_nccl.json
echo "== bench n1"; timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu --no-kernels > $O/bench_n1.json 2> $O/bench_n1.err; cut -c1-200 $O/bench_n1.json
