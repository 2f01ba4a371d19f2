#!/bin/bash
mkdir -p gpurun_out/r4p
timeout 900 python -m pytest tests/test_pipeline_gpu.py tests/test_clip_gpu.py tests/test_train_step_gpu.py tests/test_flags_gpu.py -q -s -x > gpurun_out/r4p/pytest_pipeline.log 2>&1; echo "rc=$?"
grep -E "VAE decode|pipeline tiny|passed|failed|^E |Error" gpurun_out/r4p/pytest_pipeline.log | head -30
timeout 900 python tools/time_sampler.py --steps 20 --batches 1,4,8 > gpurun_out/r4p/sampler.txt 2>&1; echo "rc=$?"; tail -14 gpurun_out/r4p/sampler.txt
