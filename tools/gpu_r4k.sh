#!/bin/bash
O=gpurun_out/r4k
mkdir -p $O
timeout 900 python -m pytest tests/test_vae_gpu.py -q -s > $O/pytest_vae.log 2>&1; echo "rc=$?"; grep -E "VAE encode|passed|failed|Error|^E " $O/pytest_vae.log | head -30
