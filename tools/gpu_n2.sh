#!/bin/bash
# 2 GPUs: NVLS (multicast) gradient exchange vs the copy-engine exchange — correctness (dp_check) and step time
O=gpurun_out/r4w
mkdir -p $O
python - <<'PY'
import torch
from torch._C._distributed_c10d import _SymmetricMemory as S
print("multicast support:", S.has_multicast_support(torch.device("cuda").type, 0))
PY
nvidia-smi topo -m 2>/dev/null | head -6
SDB_DP_ALLREDUCE=nvls timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 tools/dp_check.py --config tiny --batch 2 --steps 3 --out $O/dp_check_tiny_n2_nvls.json > $O/dp_check_nvls.log 2>&1; echo "dp nvls rc=$?"; grep -o '"exchange": "[a-z]*"' $O/dp_check_tiny_n2_nvls.json | sort | uniq -c; tail -3 $O/dp_check_nvls.log | cut -c1-300
SDB_DP_ALLREDUCE=nvls timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29562 tools/dp_check.py --config sd15 --batch 2 --steps 2 --out $O/dp_check_sd15_n2_nvls.json > $O/dp_check_sd15_nvls.log 2>&1; echo "dp sd15 nvls rc=$?"
for m in nvls peer nvls peer; do
  SDB_DP_ALLREDUCE=$m timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29563 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu --no-kernels > $O/bench_n2_$m.json 2> $O/bench_n2_$m.err
  echo "$m $(grep -o '"ms_per_step": [0-9.]*' $O/bench_n2_$m.json)"; grep -i "sdb dp\|error" $O/bench_n2_$m.err | head -3
done
for c in 32 64; do
  SDB_DP_NVLS_CTAS=$c SDB_DP_ALLREDUCE=nvls timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29564 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu --no-kernels > $O/bench_n2_nvls_c$c.json 2> $O/bench_n2_nvls_c$c.err
  echo "nvls ctas $c $(grep -o '"ms_per_step": [0-9.]*' $O/bench_n2_nvls_c$c.json)"
done
