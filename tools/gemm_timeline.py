"""Timeline of the operand ring of one GEMM CTA (clock64 at the hand-off points). Build the library with the macro first:
   NVCC_EXTRA=-DSDB_GEMM_TIMELINE python -m stable_diffusion_training_b200._build --force ; python tools/gemm_timeline.py
Answers what paces the main loop (48 B/clk per SM against the 110 B/clk tools/ingest_rate.cu
reaches without MMAs): per K block it prints when the producer found the stage free, when its TMA was issued, when the
stage had landed (issuer woke up) and when the MMAs were issued — i.e. the fetch latency, the MMA residency and the
time a free stage waited for the producer."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from stable_diffusion_training_b200 import ops, _lib  # noqa: E402

dev = torch.device("cuda:0")


def rnd(*shape, scale=1.0):
    return (torch.randn(*shape, device=dev) * scale).to(torch.bfloat16)


what = sys.argv[1] if len(sys.argv) > 1 else "conv320"
if what == "conv320":
    x, w = rnd(8, 64, 64, 320), rnd(3, 3, 320, 320, scale=0.02)
elif what == "conv1280":
    x, w = rnd(8, 16, 16, 1280), rnd(3, 3, 1280, 1280, scale=0.02)
else:
    raise SystemExit("usage: gemm_timeline.py conv320|conv1280")
bias = torch.zeros(w.shape[3], device=dev)
for _ in range(3):
    y = ops.conv3x3(x, w, bias=bias)
torch.cuda.synchronize()
lib = C.CDLL(_lib.LIB_PATH)
buf = np.zeros(3 * 512 * 2, dtype=np.int64)
rc = lib.sdb_gemm_timeline(buf.ctypes.data_as(C.c_void_p))
assert rc == 0, rc
t = buf.reshape(3, 512, 2)
n = int((t[0, :, 1] > 0).sum())
t0 = t[0, 0, 0]
print(f"{what}: {n} K blocks recorded for CTA 0 (cycles relative to the first)")
print("   k | stage free | TMA issued | landed (issuer 0) | MMAs issued | fetch latency | issue | period")
prev = None
for k in range(min(n, 64)):
    free, issued = t[0, k] - t0
    landed, mma = t[1, k] - t0
    per = "" if prev is None else f"{landed - prev:6d}"
    prev = landed
    print(f"{k:4d} | {free:10d} | {issued:10d} | {landed:17d} | {mma:11d} | {landed - issued:13d} | {mma - landed:5d} | {per}")
if n > 16:
    lat = (t[1, 8:n, 0] - t[0, 8:n, 1]).astype(np.float64)
    per = np.diff(t[1, 8:n, 0]).astype(np.float64)
    print(f"steady state: fetch latency {lat.mean():.0f} clk (min {lat.min():.0f}, max {lat.max():.0f}), "
          f"period {per.mean():.0f} clk per K block")
