#!/bin/bash
mkdir -p gpurun_out/r4m
timeout 600 python tools/time_full_parts.py > gpurun_out/r4m/parts.txt 2>&1; cat gpurun_out/r4m/parts.txt | tail -20
timeout 600 python -m pytest tests/test_vae_gpu.py -q -s > gpurun_out/r4m/pytest_vae.log 2>&1; echo "rc=$?"; grep -E "VAE encode|passed|failed|^E " gpurun_out/r4m/pytest_vae.log | head
