#!/bin/bash
# Round-2 GPU call L: the reference's complete step (VAE + CLIP + UNet) as a bench line; final suite + default bench.
O=gpurun_out/r4l
mkdir -p $O
echo "== bench sd15_full"; timeout 600 python bench.py --config sd15_full --steps 10 --warmup 3 --no-kernels > $O/bench_sd15_full.json 2> $O/bench_sd15_full.err; echo "rc=$?"; cut -c1-300 $O/bench_sd15_full.json; tail -3 $O/bench_sd15_full.err | cut -c1-300
echo "== bench sd15"; timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu --no-kernels > $O/bench_sd15.json 2> $O/bench_sd15.err; cut -c1-200 $O/bench_sd15.json
