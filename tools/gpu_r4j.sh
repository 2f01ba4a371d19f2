#!/bin/bash
O=gpurun_out/r4j
mkdir -p $O
timeout 900 python -m pytest tests/test_clip_gpu.py -q -s > $O/pytest_clip.log 2>&1; echo "rc=$?"; grep -E "CLIP text|joint step|passed|failed|Error" $O/pytest_clip.log | head -20
