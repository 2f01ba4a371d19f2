"""Run ONE operator a few times (for `ncu --set full -k regex:<kernel>`): python tools/prof_one.py attn|conv|wgrad|lion"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from stable_diffusion_training_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
what = sys.argv[1]


def rnd(*shape, scale=1.0):
    return (torch.randn(*shape, device=dev) * scale).to(torch.bfloat16)


if what == "attn":
    B, H, S, d = 8, 8, 4096, 40
    q, k, v, dO = (rnd(B, S, H * d) for _ in range(4))
    for _ in range(3):
        o, lse = ops.attention_fwd(q, k, v, H)
        ops.attention_bwd(q, k, v, o, dO, lse, H)
elif what in ("conv", "wgrad"):
    x, w, dy = rnd(8, 64, 64, 320), rnd(3, 3, 320, 320, scale=0.02), rnd(8, 64, 64, 320)
    for _ in range(3):
        ops.conv3x3(x, w, bias=torch.zeros(320, device=dev))
        ops.conv3x3_dgrad(dy, w)
        ops.conv3x3_wgrad(x, dy)
elif what == "lion":
    from stable_diffusion_training_b200 import _lib
    n = 1 << 28
    g = torch.randn(n, device=dev) * 3e-5
    p = torch.randn(n, device=dev) * 0.02
    sh = torch.empty(n, device=dev, dtype=torch.bfloat16)
    codes = torch.full((n,), 3, dtype=torch.int8, device=dev)
    inv = torch.ones(n // 16, device=dev)
    for _ in range(3):
        _lib.check(_lib.lib().sdb_lion8_update(g.data_ptr(), p.data_ptr(), p.data_ptr(), sh.data_ptr(), codes.data_ptr(),
                                               inv.data_ptr(), n, 16, 0.9, 0.99, 1e-6 / 7, 0.07, 1, 0, 1.0, 0, 0, 0.0,
                                               _lib.stream_ptr()), "lion8")
elif what == "gn":
    for (B, HW, C) in [(8, 4096, 320), (8, 1024, 640)]:
        x, dy = rnd(B, HW, C), rnd(B, HW, C)
        g, b_ = torch.ones(C, device=dev), torch.zeros(C, device=dev)
        dg, db = torch.empty(C, device=dev), torch.empty(C, device=dev)
        for _ in range(2):
            y, st = ops.groupnorm_fwd(x, g, b_, silu=True)
            ops.groupnorm_bwd(x, dy, st, g, b_, dg, db, silu=True)
torch.cuda.synchronize()
print("ok")
