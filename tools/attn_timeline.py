"""Timeline of one backward-attention CTA (clock64 at the hand-off points). Build the library with the macro first:
   NVCC_EXTRA=-DSDB_ATTN_TIMELINE python -m stable_diffusion_training_b200._build --force ; python tools/attn_timeline.py"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from stable_diffusion_training_b200 import ops, _lib  # noqa: E402

dev = torch.device("cuda:0")
B, H, S, d = 8, 8, 4096, 40
q, k, v, dO = ((torch.randn(B, S, H * d, device=dev)).to(torch.bfloat16) for _ in range(4))
o, lse = ops.attention_fwd(q, k, v, H)
for _ in range(2):
    ops.attention_bwd(q, k, v, o, dO, lse, H)
torch.cuda.synchronize()
lib = C.CDLL(_lib.LIB_PATH)
buf = np.zeros(8 * 64 * 4, dtype=np.int64)
rc = lib.sdb_attn_timeline(buf.ctypes.data_as(C.c_void_p))
assert rc == 0, rc
t = buf.reshape(8, 64, 4)
t0 = t[t > 0].min()
names = ["S issuer   [start, operands ready, issued]", "dP issuer", "acc1 issuer [start, P/dS ready, issued]", "acc2 issuer",
         "compute g0 [barrier, S/dP done, first slice computed and acc(i-2) done, arrived]", "compute g1"]
# the LAST launch was mode 1 (rows = queries): roles 3 (acc2) unused there
for r, nm in enumerate(names):
    print(nm)
    for i in range(0, 16):
        row = t[r, i]
        if row.max() > 0:
            print(f"  block {i + 8:3d}: " + " ".join(f"{(x - t0) if x > 0 else -1:8d}" for x in row))
