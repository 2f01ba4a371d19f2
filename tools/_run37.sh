run() { tag=$1; shift; env "$@" timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu --no-kernels > gpurun_out/r2s_n2_$tag.json 2> gpurun_out/r2s_n2_$tag.err; }
run base A=1
run sms140_cta8 SDB_GEMM_SMS=140 NCCL_MAX_CTAS=8
run sms132_cta16 SDB_GEMM_SMS=132 NCCL_MAX_CTAS=16
true
