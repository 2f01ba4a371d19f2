"""Per-operator microbenchmarks on one B200 (CUDA events, L2 flushed between iterations). Development tool:
    python tools/bench_ops.py [attn] [conv] [linear] [norm]
Prints one line per case: ms, achieved TFLOP/s (or GB/s) and the fraction of the measured peak."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from stable_diffusion_training_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
P = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) \
    else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(fn, iters=8):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def rnd(*shape, scale=1.0):
    return (torch.randn(*shape, device=dev) * scale).to(torch.bfloat16)


def report(name, ms, flops=None, nbytes=None):
    if flops is not None:
        tf = flops / ms / 1e9
        print(f"{name:58s} {ms:8.4f} ms  {tf:8.1f} TFLOP/s  {tf / P['bf16_tflops']:6.1%} of measured bf16 burst", flush=True)
    else:
        gb = nbytes / ms / 1e6
        print(f"{name:58s} {ms:8.4f} ms  {gb:8.1f} GB/s     {gb / P['hbm_gbs']:6.1%} of measured HBM", flush=True)


def attn():
    for (B, H, Sq, Skv, d) in [(8, 8, 4096, 4096, 40), (8, 8, 1024, 1024, 80), (8, 8, 256, 256, 160), (8, 8, 4096, 77, 40),
                               (8, 8, 1024, 77, 80), (4, 5, 9216, 9216, 64)]:
        C = H * d
        q, k, v, dO = rnd(B, Sq, C), rnd(B, Skv, C), rnd(B, Skv, C), rnd(B, Sq, C)
        o, lse = ops.attention_fwd(q, k, v, H)
        fl = 4.0 * B * H * Sq * Skv * d
        report(f"attn fwd B{B} H{H} Sq{Sq} Skv{Skv} d{d}", timed(lambda: ops.attention_fwd(q, k, v, H, out=o)), fl)
        dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
        report(f"attn bwd B{B} H{H} Sq{Sq} Skv{Skv} d{d}",
               timed(lambda: ops.attention_bwd(q, k, v, o, dO, lse, H, dq=dq, dk=dk, dv=dv)), 2.5 * fl)


def conv():
    for (B, Hh, cin, cout) in [(8, 64, 320, 320), (8, 32, 640, 640), (8, 16, 1280, 1280), (8, 8, 1280, 1280),
                               (8, 16, 2560, 1280), (8, 64, 960, 320), (8, 32, 1920, 640)]:
        x, w = rnd(B, Hh, Hh, cin), rnd(3, 3, cin, cout, scale=0.02)
        dy = rnd(B, Hh, Hh, cout)
        bias = torch.zeros(cout, device=dev)
        y, dx = torch.empty_like(dy), torch.empty_like(x)
        dw = torch.empty(3, 3, cin, cout, device=dev)
        fl = 2.0 * B * Hh * Hh * cout * 9 * cin
        report(f"conv3x3 fwd   {cin}->{cout} @{Hh}", timed(lambda: ops.conv3x3(x, w, bias=bias, out=y)), fl)
        report(f"conv3x3 dgrad {cin}->{cout} @{Hh}", timed(lambda: ops.conv3x3_dgrad(dy, w, out=dx)), fl)
        report(f"conv3x3 wgrad {cin}->{cout} @{Hh}", timed(lambda: ops.conv3x3_wgrad(x, dy, out=dw)), fl)


def linear():
    for (M, K, N) in [(32768, 320, 320), (32768, 320, 2560), (32768, 1280, 320), (8192, 640, 5120), (8192, 2560, 640),
                      (2048, 1280, 10240), (2048, 5120, 1280), (2048, 1280, 1280), (616, 768, 320)]:
        x, w, dy = rnd(M, K), rnd(K, N, scale=0.02), rnd(M, N)
        y, dx = torch.empty_like(dy), torch.empty_like(x)
        dw = torch.empty(K, N, device=dev)
        fl = 2.0 * M * K * N
        report(f"linear fwd   [{M},{K}]x[{K},{N}]", timed(lambda: ops.linear(x, w, out=y)), fl)
        report(f"linear dgrad [{M},{K}]x[{K},{N}]", timed(lambda: ops.linear_dgrad(dy, w, out=dx)), fl)
        report(f"linear wgrad [{M},{K}]x[{K},{N}]", timed(lambda: ops.linear_wgrad(x, dy, out=dw)), fl)


def norm():
    for (B, HW, C) in [(8, 4096, 320), (8, 1024, 640), (8, 256, 1280), (8, 4096, 960)]:
        x, dy = rnd(B, HW, C), rnd(B, HW, C)
        g, b_ = torch.ones(C, device=dev), torch.zeros(C, device=dev)
        dg, db = torch.empty(C, device=dev), torch.empty(C, device=dev)
        y, st = ops.groupnorm_fwd(x, g, b_, silu=True)
        n = B * HW * C
        report(f"groupnorm+silu fwd B{B} HW{HW} C{C}", timed(lambda: ops.groupnorm_fwd(x, g, b_, silu=True, out=y)), nbytes=4.0 * n)
        dx = torch.empty_like(x)
        report(f"groupnorm+silu bwd B{B} HW{HW} C{C}",
               timed(lambda: ops.groupnorm_bwd(x, dy, st, g, b_, dg, db, silu=True, out=dx)), nbytes=6.0 * n)
    for (rows, C) in [(32768, 320), (8192, 640), (2048, 1280)]:
        x = rnd(rows, C)
        g, b_ = torch.ones(C, device=dev), torch.zeros(C, device=dev)
        report(f"layernorm fwd [{rows},{C}]", timed(lambda: ops.layernorm_fwd(x, g, b_)), nbytes=4.0 * rows * C)
    for (rows, cols) in [(32768, 320), (32768, 2560), (8192, 5120), (2048, 10240)]:
        x = rnd(rows, cols)
        out = torch.empty(cols, device=dev)
        report(f"colsum [{rows},{cols}]", timed(lambda: ops.colsum(x, out)), nbytes=2.0 * rows * cols)


if __name__ == "__main__":
    which = sys.argv[1:] or ["attn", "conv", "linear", "norm"]
    for w in which:
        {"attn": attn, "conv": conv, "linear": linear, "norm": norm}[w]()
