#!/bin/bash
# Round-2 GPU call I: text encoder (8(f)-2), DDIM + checkpoint I/O (8(f)-4), then the whole GPU suite.
O=gpurun_out/r4i
mkdir -p $O
echo "== clip"; timeout 900 python -m pytest tests/test_clip_gpu.py -q -s -x > $O/pytest_clip.log 2>&1; echo "rc=$?"; grep -E "CLIP text|joint step|passed|failed|Error" $O/pytest_clip.log | head -20
echo "== ddim / checkpoint"; timeout 600 python -m pytest tests/test_golden_gpu.py tests/test_checkpoint.py -q -m gpu > $O/pytest_f4.log 2>&1; echo "rc=$?"; tail -3 $O/pytest_f4.log
echo "== whole suite"; timeout 1500 python -m pytest tests -q -m gpu > $O/pytest_gpu.log 2>&1; echo "rc=$?"; tail -3 $O/pytest_gpu.log
echo "== bench"; timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu --no-kernels > $O/bench_sd15.json 2> $O/bench_sd15.err; cut -c1-200 $O/bench_sd15.json
