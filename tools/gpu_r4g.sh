#!/bin/bash
# Round-2 GPU call G (8 GPUs): peer-memory all-reduce at 8 ranks — correctness (dp_check tiny), SD-1.5 step time vs
# NCCL, bucket size, SD-2.x config 4 and the Lion-alone config 5 at 8 GPUs, N = 1 on the same box.
O=gpurun_out/r4g
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
echo "== dp_check tiny n8"; timeout 600 $TR --nproc-per-node 8 --master-port 29511 tools/dp_check.py --config tiny --batch 2 --steps 3 --out $O/dp_check_tiny_n8.json > $O/dp_check_tiny_n8.log 2>&1; echo "rc=$?"
echo "== bench n8 peer"; timeout 600 $TR --nproc-per-node 8 --master-port 29512 bench.py --gpus 8 --steps 20 --warmup 5 --no-cpu --no-kernels > $O/bench_n8_peer.json 2> $O/bench_n8_peer.err; echo "rc=$?"; cut -c1-200 $O/bench_n8_peer.json; tail -2 $O/bench_n8_peer.err | cut -c1-300
echo "== bench n8 nccl"; SDB_DP_ALLREDUCE=nccl timeout 600 $TR --nproc-per-node 8 --master-port 29513 bench.py --gpus 8 --steps 20 --warmup 5 --no-cpu --no-kernels --allow-experiments > $O/bench_n8_nccl.json 2> $O/bench_n8_nccl.err; echo "rc=$?"; cut -c1-200 $O/bench_n8_nccl.json
echo "== bench n8 peer 128 MB buckets"; SDB_DP_BUCKET_MB=128 timeout 600 $TR --nproc-per-node 8 --master-port 29514 bench.py --gpus 8 --steps 20 --warmup 5 --no-cpu --no-kernels --allow-experiments > $O/bench_n8_peer_b128.json 2> $O/bench_n8_peer_b128.err; echo "rc=$?"; cut -c1-200 $O/bench_n8_peer_b128.json
echo "== bench n4 peer"; timeout 600 $TR --nproc-per-node 4 --master-port 29515 bench.py --gpus 4 --steps 20 --warmup 5 --no-cpu --no-kernels > $O/bench_n4_peer.json 2> $O/bench_n4_peer.err; echo "rc=$?"; cut -c1-200 $O/bench_n4_peer.json
echo "== bench n1"; timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu --no-kernels > $O/bench_n1.json 2> $O/bench_n1.err; cut -c1-200 $O/bench_n1.json
echo "== sd21 n8"; timeout 600 $TR --nproc-per-node 8 --master-port 29516 bench.py --config sd21 --gpus 8 --steps 10 --warmup 3 --no-cpu --no-kernels > $O/bench_sd21_n8.json 2> $O/bench_sd21_n8.err; echo "rc=$?"; cut -c1-200 $O/bench_sd21_n8.json
echo "== lion n8"; timeout 600 $TR --nproc-per-node 8 --master-port 29517 bench.py --config lion --gpus 8 --steps 20 --warmup 5 --no-cpu > $O/bench_lion_n8.json 2> $O/bench_lion_n8.err; echo "rc=$?"; cut -c1-200 $O/bench_lion_n8.json
