"""Host-side operator wrappers over the C ABI (include/sdb200.h). Tensors are torch CUDA tensors used purely as
device buffers; every op enqueues hand-written sm_100a kernels on the current stream.

Activation layout: NHWC bf16 (Flax's layout, diffusers FlaxUNet2DConditionModel transposes NCHW->NHWC on entry).
Weights: Flax layouts — Dense kernel (in, out), Conv kernel HWIO — row-major bf16 shadows of the fp32 masters.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import GemmDesc, View4, SDB_GEMM_FWD, SDB_GEMM_WGRAD

BF16 = torch.bfloat16
NUM_SMS = 148


def _view_mat(t: torch.Tensor) -> View4:
    """[rows, cols] matrix (cols contiguous) as a 4-D view."""
    assert t.dim() == 2 and t.stride(1) == 1
    return View4(t.data_ptr(), t.shape[1], t.shape[0], 1, 1, t.stride(0), t.stride(0) * t.shape[0],
                 t.stride(0) * t.shape[0])


def _view_nhwc(t: torch.Tensor) -> View4:
    assert t.dim() == 4 and t.stride(3) == 1
    return View4(t.data_ptr(), t.shape[3], t.shape[2], t.shape[1], t.shape[0], t.stride(2), t.stride(1), t.stride(0))


def _next_pow2(x: int) -> int:
    p = 1
    while p < x:
        p *= 2
    return p


def _pixel_box(W: int, H: int, N: int, pixels: int):
    """Box (bw, bh, bn) holding exactly `pixels` pixels (a power of two), w fastest. Extents that are not powers of
    two are covered by the next power of two: TMA zero-fills the overhang and the epilogue masks it."""
    bw = min(pixels, _next_pow2(W))
    rest = pixels // bw
    bh = min(rest, _next_pow2(H))
    bn = rest // bh
    assert bw * bh * bn == pixels and bn <= 256, (W, H, N, pixels, bw, bh, bn)
    return bw, bh, bn


def _pick_block_n(N: int, need64: bool, m_tiles: int) -> int:
    if N <= 256:
        return N if (not need64 or N % 64 == 0) else min(256, (N + 63) // 64 * 64) if N > 64 else N
    best, best_cost = 256, None
    for bn in (256, 192, 128):
        tiles = (N + bn - 1) // bn
        total = tiles * m_tiles
        waves = (total + NUM_SMS - 1) // NUM_SMS
        cost = waves * bn  # time ~ waves x tile width
        if best_cost is None or cost < best_cost:
            best, best_cost = bn, cost
    return best


def _launch(desc: GemmDesc, impl: str) -> None:
    l = _lib.lib()
    fn = l.sdb_gemm_tc if impl == "tc" else l.sdb_gemm_simt
    _lib.check(fn(C.byref(desc), _lib.stream_ptr()), "sdb_gemm_" + impl)


def _set_taps(desc: GemmDesc, taps) -> None:
    desc.ntaps = len(taps)
    for i, (src, dh, dw, brow) in enumerate(taps):
        desc.tap_src[i] = src
        desc.tap_dh[i] = dh
        desc.tap_dw[i] = dw
        desc.tap_brow[i] = brow


def gemm_fwd(a_views, taps, b: torch.Tensor, b_mn: bool, out: torch.Tensor, pix, K: int, N: int, *, bias=None,
             temb=None, residual=None, impl="tc") -> None:
    """out[pixel, n] = sum_taps A_src[pixel + shift, :K] @ Bop[tap] (+ bias + temb[img] + residual)."""
    W, H, NI = pix
    d = GemmDesc()
    d.mode = SDB_GEMM_FWD
    d.num_src = len(a_views)
    for i, v in enumerate(a_views):
        d.a[i] = v
    d.b = _view_mat(b)
    d.b_mn_major = 1 if b_mn else 0
    _set_taps(d, taps)
    d.outW, d.outH, d.outN = W, H, NI
    d.bw, d.bh, d.bn = _pixel_box(W, H, NI, 128)
    d.N, d.K = N, K
    m_tiles = -(-W // d.bw) * -(-H // d.bh) * -(-NI // d.bn)
    d.block_n = _pick_block_n(N, b_mn, m_tiles)
    d.splits = 1
    out2 = out.view(-1, out.shape[-1])
    d.d = out2.data_ptr()
    d.ldd = out2.stride(0)
    d.bias = _lib.ptr(bias)
    d.temb = _lib.ptr(temb)
    if residual is not None:
        r2 = residual.view(-1, residual.shape[-1])
        d.residual = r2.data_ptr()
        d.ldr = r2.stride(0)
    _launch(d, impl)


def gemm_wgrad(a_views, taps, b_view: View4, out: torch.Tensor, pix, M: int, N: int, *, impl="tc") -> None:
    """out[tap_brow + m, n] = sum_pixels A_src[pixel + shift, m] * B[pixel, n]; out fp32 [rows, N]."""
    W, H, NI = pix
    d = GemmDesc()
    d.mode = SDB_GEMM_WGRAD
    d.num_src = len(a_views)
    for i, v in enumerate(a_views):
        d.a[i] = v
    d.b = b_view
    d.b_mn_major = 1
    _set_taps(d, taps)
    d.outW, d.outH, d.outN = W, H, NI
    d.bw, d.bh, d.bn = _pixel_box(W, H, NI, 64)
    d.M, d.N = M, N
    d.block_n = N if N <= 256 and (N % 64 == 0 or N <= 64) else (256 if N > 256 else (N + 63) // 64 * 64)
    if d.block_n > 256:
        d.block_n = 256
    tiles = len(taps) * -(-M // 128) * -(-N // d.block_n)
    total_pb = -(-W // d.bw) * -(-H // d.bh) * -(-NI // d.bn)
    splits = max(1, min(total_pb, (2 * NUM_SMS) // max(1, tiles)))
    d.splits = splits
    if splits > 1:
        out.zero_()
    d.d = out.data_ptr()
    d.ldd = out.stride(0)
    _launch(d, impl)


# ---------------------------------------------------------------------------------------------------------------
def linear(x: torch.Tensor, w: torch.Tensor, *, bias=None, residual=None, out=None, impl="tc") -> torch.Tensor:
    """y[M,N] = x[M,K] @ w[K,N] (+bias +residual). flax nn.Dense / 1x1 nn.Conv forward."""
    M, K = x.shape
    N = w.shape[1]
    if out is None:
        out = torch.empty((M, N), device=x.device, dtype=BF16)
    gemm_fwd([_view_mat(x)], [(0, 0, 0, 0)], w, True, out, (M, 1, 1), K, N, bias=bias, residual=residual, impl=impl)
    return out


def linear_dgrad(dy: torch.Tensor, w: torch.Tensor, *, residual=None, out=None, impl="tc") -> torch.Tensor:
    """dx[M,K] = dy[M,N] @ w[K,N]^T."""
    M, N = dy.shape
    K = w.shape[0]
    if out is None:
        out = torch.empty((M, K), device=dy.device, dtype=BF16)
    gemm_fwd([_view_mat(dy)], [(0, 0, 0, 0)], w, False, out, (M, 1, 1), N, K, residual=residual, impl=impl)
    return out


def linear_wgrad(x: torch.Tensor, dy: torch.Tensor, *, out=None, impl="tc") -> torch.Tensor:
    """dw[K,N] = x[M,K]^T @ dy[M,N] (fp32)."""
    M, K = x.shape
    N = dy.shape[1]
    if out is None:
        out = torch.empty((K, N), device=x.device, dtype=torch.float32)
    gemm_wgrad([_view_mat(x)], [(0, 0, 0, 0)], _view_mat(dy), out, (M, 1, 1), K, N, impl=impl)
    return out


def _conv_taps(cin_rows: int, flip: bool):
    taps = []
    for kh in range(3):
        for kw in range(3):
            dh, dw = kh - 1, kw - 1
            if flip:
                dh, dw = -dh, -dw
            taps.append((0, dh, dw, (kh * 3 + kw) * cin_rows))
    return taps


def conv3x3(x: torch.Tensor, w: torch.Tensor, *, bias=None, temb=None, residual=None, out=None,
            impl="tc") -> torch.Tensor:
    """Flax nn.Conv 3x3 stride 1 pad 1: x NHWC bf16, w HWIO bf16 -> NHWC bf16."""
    B, H, W, Cin = x.shape
    Cout = w.shape[3]
    if out is None:
        out = torch.empty((B, H, W, Cout), device=x.device, dtype=BF16)
    gemm_fwd([_view_nhwc(x)], _conv_taps(Cin, False), w.view(9 * Cin, Cout), True, out, (W, H, B), Cin, Cout,
             bias=bias, temb=temb, residual=residual, impl=impl)
    return out


def conv3x3_dgrad(dy: torch.Tensor, w: torch.Tensor, *, out=None, impl="tc") -> torch.Tensor:
    """dx = conv_transpose(dy, w): taps flipped, weights read as [Cin rows, Cout cols] per tap (K-major B)."""
    B, H, W, Cout = dy.shape
    Cin = w.shape[2]
    if out is None:
        out = torch.empty((B, H, W, Cin), device=dy.device, dtype=BF16)
    gemm_fwd([_view_nhwc(dy)], _conv_taps(Cin, True), w.view(9 * Cin, Cout), False, out, (W, H, B), Cout, Cin,
             impl=impl)
    return out


def conv3x3_wgrad(x: torch.Tensor, dy: torch.Tensor, *, out=None, impl="tc") -> torch.Tensor:
    """dw HWIO fp32 = sum_pixels x[pixel + tap] (outer) dy[pixel]."""
    B, H, W, Cin = x.shape
    Cout = dy.shape[3]
    if out is None:
        out = torch.empty((3, 3, Cin, Cout), device=x.device, dtype=torch.float32)
    gemm_wgrad([_view_nhwc(x)], _conv_taps(Cin, False), _view_nhwc(dy), out.view(9 * Cin, Cout), (W, H, B), Cin, Cout,
               impl=impl)
    return out
