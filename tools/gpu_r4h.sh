#!/bin/bash
# Round-2 GPU call H (8 GPUs): peer all-reduce sweep — bucket size x (staged | direct peer loads), tail buckets 32 MB.
O=gpurun_out/r4h
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
port=29520
echo "== dp_check tiny n8 direct"; SDB_DP_PEER_DIRECT=1 timeout 600 $TR --master-port $port tools/dp_check.py --config tiny --batch 2 --steps 3 --out $O/dp_check_tiny_n8_direct.json > $O/dp_check_tiny_n8_direct.log 2>&1; echo "rc=$?"
for direct in 0 1; do
  for mb in 128 256 512; do
    port=$((port+1))
    echo "== bench n8 peer direct=$direct bucket=$mb"
    SDB_DP_PEER_DIRECT=$direct SDB_DP_BUCKET_MB=$mb timeout 300 $TR --master-port $port bench.py --gpus 8 --steps 20 --warmup 5 --no-cpu --no-kernels --allow-experiments > $O/bench_n8_d${direct}_b${mb}.json 2> $O/bench_n8_d${direct}_b${mb}.err; echo "rc=$?"; python -c "import json;d=json.load(open('$O/bench_n8_d${direct}_b${mb}.json'));print(d['ms_per_step'],d['value'])"
  done
done
echo "== bench n1"; timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu --no-kernels > $O/bench_n1.json 2> $O/bench_n1.err; python -c "import json;d=json.load(open('$O/bench_n1.json'));print(d['ms_per_step'],d['value'])"
