NVCC_EXTRA=-DSDB_ATTN_TIMELINE python -m stable_diffusion_training_b200._build --force > gpurun_out/tl_build.log 2>&1
timeout 300 python tools/attn_timeline.py > gpurun_out/attn_timeline2.txt 2>&1
true
