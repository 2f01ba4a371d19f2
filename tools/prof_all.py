"""One pass over the dominant kernels for `ncu --set full` (each op runs twice: read the second launch of each kernel):
    ncu --set full --clock-control none --import-source on -k regex:'gemm_tc|attn_|gn_fused|lion8|ln_|geglu' -c 80 \
        -o gpurun_out/prof python tools/prof_all.py
Shapes are the SD-1.5 / B = 8 ones of the benchmark (BASELINE.json configs[1])."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from stable_diffusion_training_b200 import _lib, ops  # noqa: E402

dev = torch.device("cuda:0")
which = set(sys.argv[1:]) or {"attn", "conv", "linear", "norm", "lion"}


def rnd(*shape, scale=1.0):
    return (torch.randn(*shape, device=dev) * scale).to(torch.bfloat16)


for rep in range(2):
    if "attn" in which:
        for (B, H, Sq, Skv, d) in [(8, 8, 4096, 4096, 40), (8, 8, 4096, 77, 40), (8, 8, 1024, 1024, 80)]:
            q, k, v, dO = rnd(B, Sq, H * d), rnd(B, Skv, H * d), rnd(B, Skv, H * d), rnd(B, Sq, H * d)
            o, lse = ops.attention_fwd(q, k, v, H)
            ops.attention_bwd(q, k, v, o, dO, lse, H)
    if "conv" in which:
        for (Hh, cin, cout) in [(64, 320, 320), (16, 1280, 1280), (8, 1280, 1280)]:
            x, w, dy = rnd(8, Hh, Hh, cin), rnd(3, 3, cin, cout, scale=0.02), rnd(8, Hh, Hh, cout)
            ops.conv3x3(x, w, bias=torch.zeros(cout, device=dev))
            ops.conv3x3_dgrad(dy, w)
            ops.conv3x3_wgrad(x, dy)
    if "linear" in which:
        for (M, K, N) in [(32768, 320, 2560), (2048, 1280, 10240)]:
            x, w, dy = rnd(M, K), rnd(K, N, scale=0.02), rnd(M, N)
            ops.linear(x, w, bias=torch.zeros(N, device=dev))
            ops.linear_dgrad(dy, w)
            ops.linear_wgrad(x, dy)
    if "norm" in which:
        B, HW, C = 8, 4096, 320
        x, dy = rnd(B, HW, C), rnd(B, HW, C)
        g, b_ = torch.ones(C, device=dev), torch.zeros(C, device=dev)
        dg, db = torch.empty(C, device=dev), torch.empty(C, device=dev)
        y, st = ops.groupnorm_fwd(x, g, b_, silu=True)
        ops.groupnorm_bwd(x, dy, st, g, b_, dg, db, silu=True)
        x2 = rnd(32768, 320)
        y2, st2 = ops.layernorm_fwd(x2, g, b_)
        ops.layernorm_bwd(x2, rnd(32768, 320), st2, g, dg, db)
        u = rnd(32768, 2560)
        ops.geglu_bwd(u, ops.geglu_fwd(u))
    if "lion" in which:
        n = 1 << 28
        gg = torch.randn(n, device=dev) * 3e-5
        p = torch.randn(n, device=dev) * 0.02
        sh = torch.empty(n, device=dev, dtype=torch.bfloat16)
        codes = torch.full((n,), 3, dtype=torch.int8, device=dev)
        inv = torch.ones(n // 16, device=dev)
        _lib.check(_lib.lib().sdb_lion8_update(gg.data_ptr(), p.data_ptr(), p.data_ptr(), sh.data_ptr(), codes.data_ptr(),
                                               inv.data_ptr(), n, 16, 0.9, 0.99, 1e-6 / 7, 0.07, 1, 0, 1.0, 0, 0, 0.0,
                                               _lib.stream_ptr()), "lion8")
        del gg, p, sh, codes, inv
torch.cuda.synchronize()
print("ok")
