#!/bin/bash
# Round-2 GPU call H (8 GPUs): peer all-reduce sweep — bucket size x (staged | direct peer loads), tail buckets 32 MB.
O=gpurun_out/r4h
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
port=29520
for cfg in "0 256" "1 256" "1 128" "0 512"; do
  set -- $cfg; direct=$1; mb=$2
  port=$((port+1))
  echo "== bench n8 peer direct=$direct bucket=$mb"
  SDB_DP_PEER_DIRECT=$direct SDB_DP_BUCKET_MB=$mb timeout 300 $TR --master-port $port bench.py --gpus 8 --steps 20 --warmup 5 --no-cpu --no-kernels --allow-experiments > $O/bench_n8_d${direct}_b${mb}.json 2> $O/bench_n8_d${direct}_b${mb}.err; echo "rc=$?"; python -c "import json;d=json.load(open('$O/bench_n8_d${direct}_b${mb}.json'));print(d['ms_per_step'],d['value'])"
done
echo "== bench n1"; timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu --no-kernels > $O/bench_n1.json 2> $O/bench_n1.err; python -c "import json;d=json.load(open('$O/bench_n1.json'));print(d['ms_per_step'],d['value'])"
