timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu --no-kernels > gpurun_out/r1u_bench_n8.json 2> gpurun_out/r1u_bench_n8.err
true
