"""Host-side operator wrappers over the C ABI (include/sdb200.h). Tensors are torch CUDA tensors used purely as
device buffers; every op enqueues hand-written sm_100a kernels on the current stream.

Activation layout: NHWC bf16 (Flax's layout, diffusers FlaxUNet2DConditionModel transposes NCHW->NHWC on entry).
Weights: Flax layouts — Dense kernel (in, out), Conv kernel HWIO — row-major bf16 shadows of the fp32 masters.
"""
from __future__ import annotations

import os

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


_TILE_CACHE: dict = {}


_DELIVERY_CLKS = float(os.environ.get("SDB_DELIVERY_CLKS", "1200"))


def _kblock_clks(mt: int, bn: int, b_mn: bool) -> float:
    """Cycles one 64-deep K block of a (128*mt x bn) tile occupies an SM — measured model (tools/mma_rate.cu and the
    SDB_GEMM_DBG experiments): the main loop runs at the slower of
      * operand delivery: the smem ring holds `stages` K blocks and a refill takes ~1200 clk under load (in-step scan of SDB_DELIVERY_CLKS, profiles/r2b; the L2
        request queues are full: ~13 TB/s chip-wide), so one K block per 3000 / stages clk whatever its size —
        which is why fewer, larger tiles (less re-fetching of A and B) win until the ring gets too shallow;
      * tcgen05.mma with both operands in shared memory: bn/2 + 43 clk per K=16 step from one issuing warp (issued
        back to back from uniform registers); two issuing warps (mt == 2, one per 128-row half) overlap their operand
        fetches: 38.5 + 0.375*bn each."""
    bcols = -(-bn // 64) * 64 if b_mn else bn
    stage_bytes = (128 * mt + bcols) * 128
    stages = max(1, min(8, (225 * 1024) // stage_bytes))
    delivery = _DELIVERY_CLKS / stages
    mma = 4 * (bn / 2 + 43.0) if mt == 1 else 8 * (38.5 + 0.375 * bn)
    return max(delivery, mma)


def _pick_tile(N: int, pix, b_mn: bool, nkb: int):
    """(m_halves, block_n) for the persistent GEMM: simulate the kernel's static round-robin schedule (tile t -> CTA
    t % 148, n-tile = t % tiles_n) for every legal shape and take the one whose busiest CTA finishes first."""
    key = (N, pix, b_mn, nkb)
    if key in _TILE_CACHE:
        return _TILE_CACHE[key]
    W, H, NI = pix
    forced_bn = int(os.environ.get("SDB_BN", "0"))
    forced_mt = int(os.environ.get("SDB_MT", "0"))
    best, best_cost = None, None
    for mt in (1, 2):
        if forced_mt and mt != forced_mt:
            continue
        try:
            bw, bh, bn_ = _pixel_box(W, H, NI, 128 * mt)
        except AssertionError:
            continue
        m_tiles = -(-W // bw) * -(-H // bh) * -(-NI // bn_)
        if mt == 2 and W * H * NI < 256:
            continue
        for bn in range(256, 63, -16):
            if forced_bn and bn != min(forced_bn, (N + 15) // 16 * 16):
                continue
            tiles_n = -(-N // bn)
            if bn * (tiles_n - 1) >= N:
                continue
            last = (N - bn * (tiles_n - 1) + 15) // 16 * 16
            total = tiles_n * m_tiles
            # per-tile fixed cost: pipeline fill; with two halves the accumulators are not double-buffered, so the
            # epilogue of a tile is exposed as well
            fixed = 1500.0 if mt == 1 else 4000.0 + 16.0 * bn
            c_full = nkb * _kblock_clks(mt, bn, b_mn) + fixed
            c_last = nkb * _kblock_clks(mt, last, b_mn) + fixed
            load = [0.0] * min(total, NUM_SMS)
            for t in range(total):
                load[t % NUM_SMS] += c_full if (t % tiles_n) < tiles_n - 1 else c_last
            cost = max(load)
            if best_cost is None or cost < best_cost:
                best, best_cost = (mt, bn), cost
    if best is None:
        best = (1, (N + 15) // 16 * 16)
    _TILE_CACHE[key] = best
    return best


_INGEST_B_PER_CLK = 48.0  # empirical: whole-kernel bytes per cycle and SM (r2g / r2f: 46-48 for every tile shape). Not a
# fabric limit (tools/ingest_rate.cu: 110 B/clk without MMAs): it folds the producer warp's ~600-clk iteration, the exposed
# epilogue and the launch overhead into one number (profiles/r3g_gemm_timeline.md); refit when those change
_SPLITK_CACHE: dict = {}


def _ingest_kblock_clks(mt: int, bn: int, b_mn: bool) -> float:
    bcols = -(-bn // 64) * 64 if b_mn else bn
    stage_bytes = (128 * mt + bcols) * 128
    mma = 4 * (bn / 2 + 43.0) if mt == 1 else 8 * (38.5 + 0.375 * bn)
    return max(stage_bytes / _INGEST_B_PER_CLK, mma)


def _model_time(N, pix, b_mn, nkb, mt, bn, splits):
    """Modelled cycles of a forward GEMM with (mt, bn) tiles and `splits` K ranges per tile, or None if illegal."""
    W, H, NI = pix
    try:
        bw, bh, bn_ = _pixel_box(W, H, NI, 128 * mt)
    except AssertionError:
        return None
    if mt == 2 and W * H * NI < 256:
        return None
    m_tiles = -(-W // bw) * -(-H // bh) * -(-NI // bn_)
    tiles_n = -(-N // bn)
    if bn * (tiles_n - 1) >= N:
        return None
    per = -(-nkb // splits)
    if splits > 1 and per * (splits - 1) >= nkb:
        return None
    fixed = 1500.0 if mt == 1 else 4000.0 + 16.0 * bn
    waves = -(-(m_tiles * tiles_n * splits) // NUM_SMS)
    t = waves * (per * _ingest_kblock_clks(mt, bn, b_mn) + fixed)
    if splits > 1:  # memset + red.add traffic + the finishing pass (two extra launches); fitted to r2v measurements
        t += 24000.0 + W * H * NI * N * (4.0 * splits + 14.0) / 3300.0
    return t


def _pick_splitk(N: int, pix, b_mn: bool, nkb: int, mt0: int, bn0: int):
    """(mt, bn, splits) when splitting K over several CTAs per output tile beats the regular pick by >= 15 % in the
    ingest model — the low-resolution layers (M = 512 / 2048 pixels) have too few 128 x ~144 tiles to afford large
    ones, and small tiles move more operand bytes per FLOP than an SM can take in (~48 B/clk)."""
    key = (N, pix, b_mn, nkb, mt0, bn0)
    if key in _SPLITK_CACHE:
        return _SPLITK_CACHE[key]
    best = None
    force = os.environ.get("SDB_SPLITK_FORCE")
    if os.environ.get("SDB_SPLITK", "1") != "0":
        if force:
            mt, bn, sp = (int(v) for v in force.split(","))
            if _model_time(N, pix, b_mn, nkb, mt, min(bn, (N + 15) // 16 * 16), sp) is not None and sp > 1:
                best = (mt, min(bn, (N + 15) // 16 * 16), sp)
        else:
            t0 = _model_time(N, pix, b_mn, nkb, mt0, bn0, 1)
            best_t = None if t0 is None else 0.85 * t0
            for mt in (1, 2):
                for bn in (128, 160, 192, 256):
                    if bn > (N + 15) // 16 * 16:
                        continue
                    for sp in range(2, 9):
                        t = _model_time(N, pix, b_mn, nkb, mt, bn, sp)
                        if t is not None and best_t is not None and t < best_t:
                            best, best_t = (mt, bn, sp), t
    _SPLITK_CACHE[key] = best
    return best


# experiment switch (SDB_EXPERIMENTS=1 SDB_SPLITK_RED=1): accumulate split-K partial sums with red.global.add into one
# buffer (the round-1 path, schedule-dependent summation order) instead of per-range slabs folded in a fixed order
_SPLITK_RED = _lib.experiments_enabled() and os.environ.get("SDB_SPLITK_RED") == "1"
# SDB_DETERMINISTIC=1 (or ops.DETERMINISTIC = True before the first step / graph capture): split weight gradients go
# through per-range slabs folded in a fixed order instead of red.global.add — with the slab split-K forward this makes a
# UNet train step reproduce bit for bit from run to run (the text encoder's token-embedding gradient still uses atomics)
DETERMINISTIC = os.environ.get("SDB_DETERMINISTIC", "0") == "1"
_SKIP_GEMM = set(x for x in os.environ.get("SDB_SKIP", "").split(",") if x.startswith("gemm_")) \
    if _lib.experiments_enabled() else set()


def _launch(desc: GemmDesc, impl: str) -> None:
    if _SKIP_GEMM and ("gemm_fwd" if desc.mode == SDB_GEMM_FWD else "gemm_wgrad") in _SKIP_GEMM:
        return  # timing experiments only (see _lib.lib)
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
             temb=None, residual=None, impl="tc", d_offset: int = 0, d_strides=None) -> None:
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
    d.N, d.K = N, K
    nkb = len(taps) * -(-K // 64)
    mt, d.block_n = _pick_tile(N, (W, H, NI), bool(b_mn), nkb)
    out2 = out.view(-1, out.shape[-1])
    sk = _pick_splitk(N, (W, H, NI), bool(b_mn), nkb, mt, d.block_n) if (impl == "tc" and d_strides is None) else None
    if sk is not None:
        # split-K: fp32 partial sums in scratch, then one pass adds bias / time embedding / residual and converts
        mt, d.block_n, d.splits = sk
        d.m_halves = mt
        d.bw, d.bh, d.bn = _pixel_box(W, H, NI, 128 * mt)
        rows = W * H * NI
        # the library's effective number of K ranges (include/sdb200.h: sdb_gemm_desc.splits)
        nslabs = -(-nkb // -(-nkb // d.splits))
        slabs = not _SPLITK_RED
        ws = workspace(out.device, rows * N * 4 * (nslabs if slabs else 1))
        d.d = ws.data_ptr()
        d.ldd = N
        d.splitk_slab = rows * N if slabs else 0
        _launch(d, impl)
        r2 = residual.view(-1, residual.shape[-1]) if residual is not None else None
        _lib.check(_lib.lib().sdb_splitk_finish(
            ws.data_ptr(), N, nslabs if slabs else 1, rows * N, _lib.ptr(bias),
            temb.data_ptr() if temb is not None else None,
            temb.stride(0) if temb is not None else 0, W * H, r2.data_ptr() if r2 is not None else None,
            r2.stride(0) if r2 is not None else 0, out2.data_ptr() + d_offset * out2.element_size(), out2.stride(0),
            rows, N, _lib.stream_ptr()), "sdb_splitk_finish")
        return
    d.m_halves = mt
    d.bw, d.bh, d.bn = _pixel_box(W, H, NI, 128 * mt)
    d.splits = 1
    d.d = out2.data_ptr() + d_offset * out2.element_size()
    d.ldd = out2.stride(0)
    if d_strides is not None:
        d.d_strideW, d.d_strideH, d.d_strideN = d_strides
    d.bias = _lib.ptr(bias)
    if temb is not None:
        d.temb = temb.data_ptr()
        d.ldt = temb.stride(0)
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
    # balanced widths: N = 320 -> 2 x 160 rather than 256 + 64 (MN-major boxes round the smem footprint up to 64)
    tiles_n = -(-N // 256)
    d.block_n = min(256, (-(-N // tiles_n) + 15) // 16 * 16)
    # two 128-row halves share each dY stage: measured to pay for the 3x3 convolutions of the large feature maps
    # (many pixel blocks per tile), not for linears or the 16x16 / 8x8 levels (r4d: 256-row tiles everywhere cost
    # +1.3 ms per step)
    d.m_halves = 2 if (M > 128 and len(taps) > 1 and W * H * NI >= 8192 and os.environ.get("SDB_MT", "0") != "1") else 1

    tiles = len(taps) * -(-M // (128 * d.m_halves)) * -(-N // d.block_n)
    total_pb = -(-W // d.bw) * -(-H // d.bh) * -(-NI // d.bn)
    splits = max(1, min(total_pb, (2 * NUM_SMS) // max(1, tiles)))
    d.splits = splits  # > 1: sdb_gemm_tc zeroes the output rows itself (memset node) before the red.add accumulation
    if DETERMINISTIC and splits > 1 and not (impl == "tc" and out.is_contiguous()):
        d.splits = splits = 1   # the CUDA-core cross-check path / strided outputs: one pixel range, no atomics
    if DETERMINISTIC and splits > 1:
        # deterministic mode: every pixel range writes its own dW-shaped slab, folded in a fixed order
        nslabs = -(-total_pb // -(-total_pb // splits))   # the library's effective number of ranges
        n = out.numel()
        ws = workspace(out.device, nslabs * n * 4)
        d.d = ws.data_ptr()
        d.ldd = N
        d.splitk_slab = n
        _launch(d, impl)
        _lib.check(_lib.lib().sdb_splitk_finish_f32(ws.data_ptr(), nslabs, n, out.data_ptr(), n, _lib.stream_ptr()),
                   "sdb_splitk_finish_f32")
        return
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


# ---- conv_in (4 -> C) / conv_out (C -> 4): CUDA-core kernels (csrc/conv_small.cu) -----------------------------------
def conv3x3_c4in(x4, w, *, bias=None, out=None):
    B, H, W, _ = x4.shape
    Cc = w.shape[3]
    out = torch.empty((B, H, W, Cc), device=x4.device, dtype=BF16) if out is None else out
    _lib.check(_lib.lib().sdb_conv3x3_c4in_fwd(x4.data_ptr(), w.data_ptr(), _lib.ptr(bias), out.data_ptr(), B, H, W, Cc,
                                               _lib.stream_ptr()), "sdb_conv3x3_c4in_fwd")
    return out


def conv3x3_c4out(x, w, *, bias=None, out=None):
    B, H, W, Cc = x.shape
    out = torch.empty((B, H, W, 4), device=x.device, dtype=BF16) if out is None else out
    _lib.check(_lib.lib().sdb_conv3x3_c4out_fwd(x.data_ptr(), w.data_ptr(), _lib.ptr(bias), out.data_ptr(), B, H, W, Cc,
                                                _lib.stream_ptr()), "sdb_conv3x3_c4out_fwd")
    return out


def conv3x3_c4out_dgrad(dy4, w, *, out=None):
    B, H, W, _ = dy4.shape
    Cc = w.shape[2]
    out = torch.empty((B, H, W, Cc), device=dy4.device, dtype=BF16) if out is None else out
    _lib.check(_lib.lib().sdb_conv3x3_c4out_dgrad(dy4.data_ptr(), w.data_ptr(), out.data_ptr(), B, H, W, Cc,
                                                  _lib.stream_ptr()), "sdb_conv3x3_c4out_dgrad")
    return out


def conv3x3_c4_wgrad(x, dy, *, out):
    """dw HWIO fp32 for conv_in (x has 4 channels) or conv_out (dy has 4 channels)."""
    B, H, W, cin = x.shape
    c4_is_input = cin == 4
    wide, c4 = (dy, x) if c4_is_input else (x, dy)
    # per-CTA partial sums folded in CTA order (no atomics: the same bits every run; +~5 us on a 0.13 ms kernel)
    ws = workspace(out.device, _lib.lib().sdb_conv3x3_c4_wgrad_workspace_bytes(wide.shape[3]))
    _lib.check(_lib.lib().sdb_conv3x3_c4_wgrad(wide.data_ptr(), c4.data_ptr(), out.data_ptr(), B, H, W, wide.shape[3],
                                               int(c4_is_input), ws.data_ptr(), ws.numel(), _lib.stream_ptr()),
               "sdb_conv3x3_c4_wgrad")
    return out


# ---- 3x3 stride-2 convolution (FlaxDownsample2D): the input is read through four parity sub-grids -----------------
def _parity_views(x: torch.Tensor):
    """x NHWC [B,H,W,C] (H, W even) -> 4 views x[:, ph::2, pw::2, :], index ph*2+pw."""
    B, H, W, Cc = x.shape
    es = x.element_size()
    return [View4(x.data_ptr() + (ph * x.stride(1) + pw * x.stride(2)) * es, Cc, W // 2, H // 2, B, 2 * x.stride(2),
                  2 * x.stride(1), x.stride(0)) for ph in (0, 1) for pw in (0, 1)]


def _s2_taps(cin: int):
    # output (ho, wo) reads input (2ho + kh - 1, 2wo + kw - 1): kh=0 -> parity 1 shifted by -1, kh=1 -> parity 0,
    # kh=2 -> parity 1 (the -1 row of output 0 is the zero padding: TMA out-of-bounds fill)
    par = {0: (1, -1), 1: (0, 0), 2: (1, 0)}
    return [(par[kh][0] * 2 + par[kw][0], par[kh][1], par[kw][1], (kh * 3 + kw) * cin)
            for kh in range(3) for kw in range(3)]


def conv3x3_s2(x, w, *, bias=None, out=None, impl="tc"):
    B, H, W, Cin = x.shape
    Cout = w.shape[3]
    if out is None:
        out = torch.empty((B, H // 2, W // 2, Cout), device=x.device, dtype=BF16)
    gemm_fwd(_parity_views(x), _s2_taps(Cin), w.view(9 * Cin, Cout), True, out, (W // 2, H // 2, B), Cin, Cout,
             bias=bias, impl=impl)
    return out


def conv3x3_s2_wgrad(x, dy, *, out=None, impl="tc"):
    B, H, W, Cin = x.shape
    Cout = dy.shape[3]
    if out is None:
        out = torch.empty((3, 3, Cin, Cout), device=x.device, dtype=torch.float32)
    gemm_wgrad(_parity_views(x), _s2_taps(Cin), _view_nhwc(dy), out.view(9 * Cin, Cout), (W // 2, H // 2, B), Cin,
               Cout, impl=impl)
    return out


def conv3x3_s2_dgrad(dy, w, *, out=None, impl="tc"):
    """dx[2a+ph, 2c+pw] = sum over the taps of matching parity of dy[a+dh, c+dw] @ w[kh,kw]^T — one launch per output
    parity class, each writing its strided sub-grid of dx (exact FLOPs, no zero-insertion)."""
    B, Ho, Wo, Cout = dy.shape
    Cin = w.shape[2]
    H, W = 2 * Ho, 2 * Wo
    if out is None:
        out = torch.empty((B, H, W, Cin), device=dy.device, dtype=BF16)
    sel = {0: [(1, 0)], 1: [(0, 1), (2, 0)]}  # parity -> [(k, shift)]
    for ph in (0, 1):
        for pw in (0, 1):
            taps = [(0, dh, dw, (kh * 3 + kw) * Cin) for kh, dh in sel[ph] for kw, dw in sel[pw]]
            gemm_fwd([_view_nhwc(dy)], taps, w.view(9 * Cin, Cout), False, out, (Wo, Ho, B), Cout, Cin, impl=impl,
                     d_offset=(ph * W + pw) * Cin, d_strides=(2 * Cin, 2 * W * Cin, H * W * Cin))
    return out


# ---------------------------------------------------------------------------------------------------------------
# normalisation / activation / attention / data movement
_WS = {}
_WS_TAG = "main"


class side_workspace:
    """Ops enqueued inside this context use a second scratch buffer: they run on another stream than the main chain
    (unet.Graph forks the parameter-gradient GEMMs / reductions), so they must not share its stream-ordered scratch."""

    def __enter__(self):
        global _WS_TAG
        self._old, _WS_TAG = _WS_TAG, "side"

    def __exit__(self, *exc):
        global _WS_TAG
        _WS_TAG = self._old
        return False


_WS_RETIRED: list = []
_WS_FROZEN = False


def freeze_workspaces() -> None:
    """Called before the first CUDA-graph capture: graphs bake the scratch pointers into their kernel nodes, so from
    here on a workspace that has to grow keeps its predecessor alive (never freed, never handed to another tensor)."""
    global _WS_FROZEN
    _WS_FROZEN = True


def workspace(device, nbytes: int = 64 << 20) -> torch.Tensor:
    """One stream-ordered scratch buffer per device and stream role (partials of reductions; every op finishes with it
    in-stream). Growth after a graph capture retires the old buffer instead of freeing it (captured graphs of other
    batch shapes still write to it)."""
    key = (device.type, device.index, _WS_TAG)
    ws = _WS.get(key)
    if ws is None or ws.numel() < nbytes:
        if ws is not None and _WS_FROZEN:
            _WS_RETIRED.append(ws)
        ws = torch.empty(max(nbytes, 64 << 20), dtype=torch.uint8, device=device)
        _WS[key] = ws
    return ws


def _L():
    return _lib.lib()


def groupnorm_fwd(x, gamma, beta, *, silu: bool, eps: float = 1e-5, out=None):
    """x [B, HW, C] (or NHWC) bf16 -> (y, stats[B,32,2])."""
    B, C = x.shape[0], x.shape[-1]
    HW = x.numel() // (B * C)
    y = torch.empty_like(x) if out is None else out
    stats = torch.empty((B, 32, 2), device=x.device, dtype=torch.float32)
    ws = workspace(x.device)
    _lib.check(_L().sdb_groupnorm_fwd(x.data_ptr(), gamma.data_ptr(), beta.data_ptr(), y.data_ptr(), stats.data_ptr(),
                                      B, HW, C, eps, int(silu), ws.data_ptr(), ws.numel(), _lib.stream_ptr()),
               "sdb_groupnorm_fwd")
    return y, stats


def groupnorm_bwd(x, dy, stats, gamma, beta, dgamma, dbeta, *, silu: bool, dx_add=None, out=None):
    B, C = x.shape[0], x.shape[-1]
    HW = x.numel() // (B * C)
    dx = torch.empty_like(x) if out is None else out
    ws = workspace(x.device)
    _lib.check(_L().sdb_groupnorm_bwd(x.data_ptr(), dy.data_ptr(), stats.data_ptr(), gamma.data_ptr(), beta.data_ptr(),
                                      _lib.ptr(dx_add), dx.data_ptr(), dgamma.data_ptr(), dbeta.data_ptr(), B, HW, C,
                                      int(silu), ws.data_ptr(), ws.numel(), _lib.stream_ptr()), "sdb_groupnorm_bwd")
    return dx


def layernorm_fwd(x, gamma, beta, *, eps: float = 1e-5):
    C = x.shape[-1]
    rows = x.numel() // C
    y = torch.empty_like(x)
    stats = torch.empty((rows, 2), device=x.device, dtype=torch.float32)
    _lib.check(_L().sdb_layernorm_fwd(x.data_ptr(), gamma.data_ptr(), beta.data_ptr(), y.data_ptr(), stats.data_ptr(),
                                      rows, C, eps, _lib.stream_ptr()), "sdb_layernorm_fwd")
    return y, stats


def layernorm_bwd(x, dy, stats, gamma, dgamma, dbeta, *, dx_add=None, out=None):
    C = x.shape[-1]
    rows = x.numel() // C
    dx = torch.empty_like(x) if out is None else out
    ws = workspace(x.device)
    _lib.check(_L().sdb_layernorm_bwd(x.data_ptr(), dy.data_ptr(), stats.data_ptr(), gamma.data_ptr(),
                                      _lib.ptr(dx_add), dx.data_ptr(), dgamma.data_ptr(), dbeta.data_ptr(), rows, C,
                                      ws.data_ptr(), ws.numel(), _lib.stream_ptr()), "sdb_layernorm_bwd")
    return dx


def geglu_fwd(u):
    H = u.shape[-1] // 2
    rows = u.numel() // (2 * H)
    out = torch.empty(u.shape[:-1] + (H,), device=u.device, dtype=BF16)
    _lib.check(_L().sdb_geglu_fwd(u.data_ptr(), out.data_ptr(), rows, H, _lib.stream_ptr()), "sdb_geglu_fwd")
    return out


def geglu_bwd(u, dout):
    H = u.shape[-1] // 2
    rows = u.numel() // (2 * H)
    du = torch.empty_like(u)
    _lib.check(_L().sdb_geglu_bwd(u.data_ptr(), dout.data_ptr(), du.data_ptr(), rows, H, _lib.stream_ptr()),
               "sdb_geglu_bwd")
    return du


def attention_fwd(q, k, v, heads: int, *, kbias2=None, out=None):
    """q [B,Sq,C], k/v [B,Skv,C] bf16 (last dim contiguous, arbitrary row stride) -> (o [B,Sq,C], lse2 [B,H,Sq])."""
    B, Sq, Cc = q.shape
    Skv = k.shape[1]
    d = Cc // heads
    o = torch.empty((B, Sq, Cc), device=q.device, dtype=BF16) if out is None else out
    lse2 = torch.empty((B, heads, Sq), device=q.device, dtype=torch.float32)
    _lib.check(_L().sdb_attn_fwd(q.data_ptr(), q.stride(1), k.data_ptr(), k.stride(1), v.data_ptr(), v.stride(1),
                                 o.data_ptr(), o.stride(1), lse2.data_ptr(), _lib.ptr(kbias2), B, heads, Sq, Skv, d,
                                 1.0 / d ** 0.5, _lib.stream_ptr()), "sdb_attn_fwd")
    return o, lse2


def attention_bwd(q, k, v, o, dO, lse2, heads: int, *, kbias2=None, dq=None, dk=None, dv=None):
    B, Sq, Cc = q.shape
    Skv = k.shape[1]
    d = Cc // heads
    dq = torch.empty((B, Sq, Cc), device=q.device, dtype=BF16) if dq is None else dq
    dk = torch.empty((B, Skv, Cc), device=q.device, dtype=BF16) if dk is None else dk
    dv = torch.empty((B, Skv, Cc), device=q.device, dtype=BF16) if dv is None else dv
    ws = workspace(q.device)
    _lib.check(_L().sdb_attn_bwd(q.data_ptr(), q.stride(1), k.data_ptr(), k.stride(1), v.data_ptr(), v.stride(1),
                                 o.data_ptr(), o.stride(1), dO.data_ptr(), dO.stride(1), lse2.data_ptr(),
                                 _lib.ptr(kbias2), dq.data_ptr(), dq.stride(1), dk.data_ptr(), dk.stride(1),
                                 dv.data_ptr(), dv.stride(1), B, heads, Sq, Skv, d, 1.0 / d ** 0.5, ws.data_ptr(),
                                 ws.numel(), _lib.stream_ptr()), "sdb_attn_bwd")
    return dq, dk, dv


def cast_bf16(src: torch.Tensor, out=None) -> torch.Tensor:
    out = torch.empty(src.shape, device=src.device, dtype=BF16) if out is None else out
    _lib.check(_L().sdb_cast_f32_bf16(src.data_ptr(), out.data_ptr(), src.numel(), _lib.stream_ptr()), "sdb_cast")
    return out


def nchw_to_nhwc_bf16(src: torch.Tensor) -> torch.Tensor:
    B, Cc, H, W = src.shape
    out = torch.empty((B, H, W, Cc), device=src.device, dtype=BF16)
    _lib.check(_L().sdb_nchw_f32_to_nhwc_bf16(src.data_ptr(), out.data_ptr(), B, Cc, H * W, _lib.stream_ptr()), "nhwc")
    return out


def upsample2x_fwd(x):
    B, H, W, Cc = x.shape
    y = torch.empty((B, 2 * H, 2 * W, Cc), device=x.device, dtype=BF16)
    _lib.check(_L().sdb_upsample2x_fwd(x.data_ptr(), y.data_ptr(), B, H, W, Cc, _lib.stream_ptr()), "upsample_fwd")
    return y


def upsample2x_bwd(dy):
    B, H2, W2, Cc = dy.shape
    dx = torch.empty((B, H2 // 2, W2 // 2, Cc), device=dy.device, dtype=BF16)
    _lib.check(_L().sdb_upsample2x_bwd(dy.data_ptr(), dx.data_ptr(), B, H2 // 2, W2 // 2, Cc, _lib.stream_ptr()),
               "upsample_bwd")
    return dx


def copy2d(src2d, dst2d, accumulate: bool = False):
    """dst[r, :cols] (=|+=) src[r, :cols]; both 2-D bf16 views with contiguous last dim."""
    rows, cols = src2d.shape
    _lib.check(_L().sdb_copy2d_bf16(src2d.data_ptr(), src2d.stride(0), dst2d.data_ptr(), dst2d.stride(0), rows, cols,
                                    int(accumulate), _lib.stream_ptr()), "sdb_copy2d_bf16")


def add_(a, b):
    """a += b (bf16, same shape, contiguous)."""
    _lib.check(_L().sdb_add_bf16(a.data_ptr(), b.data_ptr(), a.data_ptr(), a.numel(), _lib.stream_ptr()), "sdb_add")
    return a


def colsum(x2d, out=None, *, nseg: int = 1, out_seg=None):
    """out[c] = sum_r x[r, c] (fp32, bias gradient); out_seg[s, c] (bf16) = the same per segment of rows/nseg rows."""
    rows, cols = x2d.shape
    ws = workspace(x2d.device)
    _lib.check(_L().sdb_colsum_bf16(x2d.data_ptr(), x2d.stride(0), nseg, rows // nseg, cols, _lib.ptr(out),
                                    _lib.ptr(out_seg), ws.data_ptr(), ws.numel(), _lib.stream_ptr()),
               "sdb_colsum_bf16")
    return out


def silu_fwd(x):
    y = torch.empty_like(x)
    _lib.check(_L().sdb_silu_fwd(x.data_ptr(), y.data_ptr(), x.numel(), _lib.stream_ptr()), "sdb_silu_fwd")
    return y


def silu_bwd(x, dy):
    dx = torch.empty_like(x)
    _lib.check(_L().sdb_silu_bwd(x.data_ptr(), dy.data_ptr(), dx.data_ptr(), x.numel(), _lib.stream_ptr()), "silu_bwd")
    return dx


def timestep_embedding(t, dim: int, flip_sin_to_cos: bool = True, freq_shift: float = 0.0):
    out = torch.empty((t.shape[0], dim), device=t.device, dtype=BF16)
    _lib.check(_L().sdb_timestep_embedding(t.data_ptr(), out.data_ptr(), t.shape[0], dim, int(flip_sin_to_cos),
                                           freq_shift, _lib.stream_ptr()), "sdb_timestep_embedding")
    return out


# ---- text encoder (csrc/clip.cu) -----------------------------------------------------------------------------------
def embed_fwd(input_ids, tok_w, pos_w):
    """input_ids int32 [N, S]; tok_w bf16 [vocab, D]; pos_w bf16 [>= S, D] -> bf16 [N * S, D]."""
    N, S = input_ids.shape
    D = tok_w.shape[1]
    out = torch.empty((N * S, D), device=tok_w.device, dtype=BF16)
    _lib.check(_L().sdb_embed_fwd(input_ids.data_ptr(), tok_w.data_ptr(), pos_w.data_ptr(), out.data_ptr(), N * S, S, D,
                                  tok_w.shape[0], _lib.stream_ptr()), "sdb_embed_fwd")
    return out


def embed_bwd(input_ids, dout, dtok, dpos):
    N, S = input_ids.shape
    D = dout.shape[1]
    _lib.check(_L().sdb_embed_bwd(input_ids.data_ptr(), dout.data_ptr(), dtok.data_ptr(), dpos.data_ptr(), N * S, S, D,
                                  dtok.shape[0], _lib.stream_ptr()), "sdb_embed_bwd")


def attention_small_fwd(q, k, v, N: int, heads: int, scale: float, causal: bool):
    """q / k / v bf16 [N * S, C] (column slices allowed) -> (o [N * S, C], lse [N, heads, S])."""
    C_ = q.shape[1]
    S = q.shape[0] // N
    o = torch.empty((N * S, C_), device=q.device, dtype=BF16)
    lse = torch.empty((N, heads, S), device=q.device, dtype=torch.float32)
    _lib.check(_L().sdb_attn_small_fwd(q.data_ptr(), q.stride(0), k.data_ptr(), k.stride(0), v.data_ptr(), v.stride(0),
                                       o.data_ptr(), o.stride(0), lse.data_ptr(), N, heads, S, C_ // heads, scale,
                                       int(causal), _lib.stream_ptr()), "sdb_attn_small_fwd")
    return o, lse


def attention_small_bwd(q, k, v, o, dO, lse, N: int, heads: int, scale: float, causal: bool):
    C_ = q.shape[1]
    S = q.shape[0] // N
    dq, dk, dv = (torch.empty((N * S, C_), device=q.device, dtype=BF16) for _ in range(3))
    _lib.check(_L().sdb_attn_small_bwd(q.data_ptr(), q.stride(0), k.data_ptr(), k.stride(0), v.data_ptr(), v.stride(0),
                                       o.data_ptr(), o.stride(0), dO.data_ptr(), dO.stride(0), lse.data_ptr(),
                                       dq.data_ptr(), dq.stride(0), dk.data_ptr(), dk.stride(0), dv.data_ptr(),
                                       dv.stride(0), N, heads, S, C_ // heads, scale, int(causal), _lib.stream_ptr()),
               "sdb_attn_small_bwd")
    return dq, dk, dv


ACT_KINDS = {"quick_gelu": 0, "gelu": 1}


def act_fwd(x, kind: str):
    y = torch.empty_like(x)
    _lib.check(_L().sdb_act_fwd(x.data_ptr(), y.data_ptr(), x.numel(), ACT_KINDS[kind], _lib.stream_ptr()), "sdb_act_fwd")
    return y


def act_bwd(x, dy, kind: str):
    dx = torch.empty_like(x)
    _lib.check(_L().sdb_act_bwd(x.data_ptr(), dy.data_ptr(), dx.data_ptr(), x.numel(), ACT_KINDS[kind],
                                _lib.stream_ptr()), "sdb_act_bwd")
    return dx


def gather_rows(src2d, row_index, rows_out: int):
    """out[r, :] = src[row_index[r], :] (bf16, contiguous)."""
    out = torch.empty((rows_out, src2d.shape[1]), device=src2d.device, dtype=BF16)
    _lib.check(_L().sdb_gather_rows_bf16(src2d.data_ptr(), row_index.data_ptr(), out.data_ptr(), rows_out,
                                         src2d.shape[1], 0, _lib.stream_ptr()), "sdb_gather_rows_bf16")
    return out


def scatter_rows(dout2d, row_index, rows_src: int):
    """Adjoint of gather_rows: dsrc zero except dsrc[row_index[r], :] = dout[r, :] (the call zeroes dsrc itself)."""
    dsrc = torch.empty((rows_src, dout2d.shape[1]), device=dout2d.device, dtype=BF16)
    _lib.check(_L().sdb_gather_rows_bf16(dout2d.data_ptr(), row_index.data_ptr(), dsrc.data_ptr(), dout2d.shape[0],
                                         dout2d.shape[1], rows_src, _lib.stream_ptr()), "sdb_gather_rows_bf16")
    return dsrc
