for cfg in "2 64" "4 32"; do set -- $cfg
SDB_DP_CUT=$1 SDB_DP_BUCKET_MB=$2 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu --no-kernels > gpurun_out/r1q_n8_cut$1_b$2.json 2> gpurun_out/r1q_n8_cut$1_b$2.err
done
true
