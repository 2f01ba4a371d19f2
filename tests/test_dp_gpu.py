"""Data-parallel train_step equivalence on hardware (needs >= 2 GPUs; skipped on a 1-GPU box): tools/dp_check.py under
torchrun — N ranks x B == 1 rank x N*B, ranks bit-identical after every step in eager and segmented-graph mode,
sharded in-step sampling == the global draw. The committed record of the full-size run is profiles/*_dp_check*.json."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_two_ranks_equal_one_rank_on_the_concatenated_batch(cuda):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tools", "dp_check.py"), "--config", "tiny",
           "--batch", "2", "--steps", "3"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    rep = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert rep["ok"], rep
