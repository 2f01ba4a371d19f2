#!/bin/bash
mkdir -p gpurun_out/r4q
timeout 600 python -m pytest tests/test_clip_gpu.py -q -x > gpurun_out/r4q/pytest_clip.log 2>&1; echo "clip rc=$?"; tail -3 gpurun_out/r4q/pytest_clip.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 tools/dp_check.py --config tiny --batch 2 --steps 3 --out gpurun_out/r4q/dp_check_tiny_joint_n2.json > gpurun_out/r4q/dp_check.log 2>&1; echo "dp rc=$?"
tail -5 gpurun_out/r4q/dp_check.log | cut -c1-1500
python - <<'PY'
import json
try:
    r=json.load(open('gpurun_out/r4q/dp_check_tiny_joint_n2.json'))
    print('ok', r['ok'], 'joint', json.dumps(r.get('joint'))[:1500])
except Exception as e: print('no report', e)
PY
