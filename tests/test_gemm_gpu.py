"""GPU parity of the tcgen05 GEMM / implicit-GEMM conv (sdb_gemm_tc) against torch fp32 on the same bf16 inputs.

Tolerance: inputs are exact bf16, accumulation fp32, output rounded to bf16 once -> |err| <= 2^-8 * |ref| + small
absolute slack for cancellation (stated per test).
"""
import pytest
import torch

from stable_diffusion_training_b200 import ops

pytestmark = pytest.mark.gpu


def _rand(shape, dev, scale=1.0, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(torch.bfloat16).to(dev)


def _close(out, ref, rtol=1.0 / 128, atol=2e-2):
    out = out.float()
    err = (out - ref).abs()
    tol = atol + rtol * ref.abs()
    bad = (err > tol).sum().item()
    assert bad == 0, f"{bad} mismatches, max err {err.max().item()} (ref max {ref.abs().max().item()})"


@pytest.mark.parametrize("M,K,N", [(128, 64, 64), (256, 128, 128), (512, 320, 320), (1024, 1280, 640), (384, 96, 48),
                                   (8192, 640, 5120), (616, 768, 320)])
@pytest.mark.parametrize("impl", ["tc", "simt"])
def test_linear_fwd(cuda, M, K, N, impl):
    if impl == "simt" and M * N * K > 2 ** 28:
        pytest.skip("simt reference path is slow")
    x = _rand((M, K), cuda, seed=1)
    w = _rand((K, N), cuda, 1.0 / K ** 0.5, seed=2)
    bias = torch.randn(N, device=cuda)
    res = _rand((M, N), cuda, seed=3)
    y = ops.linear(x, w, bias=bias, residual=res, impl=impl)
    torch.cuda.synchronize()
    ref = x.float() @ w.float() + bias + res.float()
    _close(y, ref)


@pytest.mark.parametrize("M,K,N", [(128, 64, 64), (512, 320, 320), (1024, 640, 1280), (384, 48, 96)])
def test_linear_dgrad(cuda, M, K, N):
    # dx[M,K] = dy[M,N] @ w[K,N]^T : weights used as [N' = K rows, K' = N cols], K-major B
    dy = _rand((M, N), cuda, seed=1)
    w = _rand((K, N), cuda, 1.0 / N ** 0.5, seed=2)
    dx = ops.linear_dgrad(dy, w)
    torch.cuda.synchronize()
    _close(dx, dy.float() @ w.float().t())


@pytest.mark.parametrize("M,K,N", [(128, 64, 64), (512, 320, 320), (4096, 640, 1280), (8192, 96, 48), (616, 768, 320)])
def test_linear_wgrad(cuda, M, K, N):
    x = _rand((M, K), cuda, seed=1)
    dy = _rand((M, N), cuda, seed=2)
    dw = ops.linear_wgrad(x, dy)
    torch.cuda.synchronize()
    ref = x.float().t() @ dy.float()
    err = (dw - ref).abs().max().item()
    assert err <= 1e-3 * max(1.0, ref.abs().max().item()), err


@pytest.mark.parametrize("B,H,W,Cin,Cout", [(2, 8, 8, 64, 64), (2, 16, 16, 320, 320), (1, 32, 32, 128, 64),
                                            (2, 16, 8, 96, 32), (8, 8, 8, 1280, 1280)])
def test_conv3x3_fwd_dgrad_wgrad(cuda, B, H, W, Cin, Cout):
    x = _rand((B, H, W, Cin), cuda, seed=1)
    w = _rand((3, 3, Cin, Cout), cuda, 1.0 / (9 * Cin) ** 0.5, seed=2)
    bias = torch.randn(Cout, device=cuda)
    y = ops.conv3x3(x, w, bias=bias)
    torch.cuda.synchronize()
    xt = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    wt = w.float().permute(3, 2, 0, 1).contiguous().requires_grad_(True)
    ref = torch.nn.functional.conv2d(xt, wt, bias, padding=1)
    _close(y, ref.permute(0, 2, 3, 1))
    dy = _rand((B, H, W, Cout), cuda, seed=3)
    ref.backward(dy.float().permute(0, 3, 1, 2))
    dx = ops.conv3x3_dgrad(dy, w)
    dw = ops.conv3x3_wgrad(x, dy)
    torch.cuda.synchronize()
    _close(dx, xt.grad.permute(0, 2, 3, 1), atol=5e-2)
    ref_dw = wt.grad.permute(2, 3, 1, 0)
    err = (dw - ref_dw).abs().max().item()
    assert err <= 2e-3 * max(1.0, ref_dw.abs().max().item()), err


@pytest.mark.parametrize("B,H,W,Cin,Cout", [(2, 16, 16, 64, 64), (2, 32, 32, 320, 320), (1, 8, 8, 128, 128)])
def test_conv3x3_stride2(cuda, B, H, W, Cin, Cout):
    x = _rand((B, H, W, Cin), cuda, seed=1)
    w = _rand((3, 3, Cin, Cout), cuda, 1.0 / (9 * Cin) ** 0.5, seed=2)
    bias = torch.randn(Cout, device=cuda)
    y = ops.conv3x3_s2(x, w, bias=bias)
    torch.cuda.synchronize()
    xt = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    wt = w.float().permute(3, 2, 0, 1).contiguous().requires_grad_(True)
    ref = torch.nn.functional.conv2d(xt, wt, bias, stride=2, padding=1)
    _close(y, ref.permute(0, 2, 3, 1))
    dy = _rand((B, H // 2, W // 2, Cout), cuda, seed=3)
    ref.backward(dy.float().permute(0, 3, 1, 2))
    dx = ops.conv3x3_s2_dgrad(dy, w)
    dw = ops.conv3x3_s2_wgrad(x, dy)
    torch.cuda.synchronize()
    _close(dx, xt.grad.permute(0, 2, 3, 1), atol=5e-2)
    ref_dw = wt.grad.permute(2, 3, 1, 0)
    err = (dw - ref_dw).abs().max().item()
    assert err <= 2e-3 * max(1.0, ref_dw.abs().max().item()), err


@pytest.mark.parametrize("Cin,Cout", [(4, 320), (320, 4), (4, 32)])
def test_conv3x3_simt_small_channels(cuda, Cin, Cout):
    """conv_in / conv_out shapes (4 channels) run on the CUDA-core kernel."""
    B, H, W = 2, 16, 16
    x = _rand((B, H, W, Cin), cuda, seed=1)
    w = _rand((3, 3, Cin, Cout), cuda, 1.0 / (9 * Cin) ** 0.5, seed=2)
    bias = torch.randn(Cout, device=cuda)
    y = ops.conv3x3(x, w, bias=bias, impl="simt")
    xt = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    wt = w.float().permute(3, 2, 0, 1).contiguous().requires_grad_(True)
    ref = torch.nn.functional.conv2d(xt, wt, bias, padding=1)
    _close(y, ref.permute(0, 2, 3, 1))
    dy = _rand((B, H, W, Cout), cuda, seed=3)
    ref.backward(dy.float().permute(0, 3, 1, 2))
    dx = ops.conv3x3_dgrad(dy, w, impl="simt")
    dw = ops.conv3x3_wgrad(x, dy, impl="simt")
    torch.cuda.synchronize()
    _close(dx, xt.grad.permute(0, 2, 3, 1), atol=5e-2)
    ref_dw = wt.grad.permute(2, 3, 1, 0)
    assert (dw - ref_dw).abs().max().item() <= 2e-3 * max(1.0, ref_dw.abs().max().item())


@pytest.fixture
def force_tile(monkeypatch):
    """Force the tile picker (SDB_MT: 128-row halves per tile, SDB_BN: tile width) and drop its cache around a test."""
    def _set(mt=None, bn=None):
        for k, v in (("SDB_MT", mt), ("SDB_BN", bn)):
            if v is None:
                monkeypatch.delenv(k, raising=False)
            else:
                monkeypatch.setenv(k, str(v))
        ops._TILE_CACHE.clear()
    yield _set
    ops._TILE_CACHE.clear()


@pytest.mark.parametrize("mt", [1, 2])
@pytest.mark.parametrize("bn", [None, 80, 160, 256])
@pytest.mark.parametrize("B,H,W,Cin,Cout", [(2, 32, 32, 128, 320), (8, 32, 32, 320, 160), (1, 16, 16, 64, 96),
                                            (3, 12, 20, 64, 64)])
def test_conv3x3_tile_shapes(cuda, force_tile, mt, bn, B, H, W, Cin, Cout):
    """256-row tiles (two issuing warps sharing each B stage) and every tile width give the same result as 128-row
    tiles: forward (MN-major weights), dgrad (K-major weights, fused gradient accumulation) and wgrad."""
    force_tile(mt, bn)
    x = _rand((B, H, W, Cin), cuda, seed=1)
    w = _rand((3, 3, Cin, Cout), cuda, 1.0 / (9 * Cin) ** 0.5, seed=2)
    bias = torch.randn(Cout, device=cuda)
    res = _rand((B, H, W, Cout), cuda, seed=5)
    y = ops.conv3x3(x, w, bias=bias, residual=res)
    xt = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    wt = w.float().permute(3, 2, 0, 1).contiguous().requires_grad_(True)
    ref = torch.nn.functional.conv2d(xt, wt, bias, padding=1)
    torch.cuda.synchronize()
    _close(y, ref.permute(0, 2, 3, 1) + res.float())
    dy = _rand((B, H, W, Cout), cuda, seed=3)
    ref.backward(dy.float().permute(0, 3, 1, 2))
    dx = ops.conv3x3_dgrad(dy, w)
    dw = ops.conv3x3_wgrad(x, dy)
    torch.cuda.synchronize()
    _close(dx, xt.grad.permute(0, 2, 3, 1), atol=4e-2)
    ref_dw = wt.grad.permute(2, 3, 1, 0)
    err = (dw - ref_dw).abs().max().item()
    assert err <= 2e-3 * max(1.0, ref_dw.abs().max().item()), err


@pytest.fixture
def force_splitk(monkeypatch):
    """Force the split-K forward path (SDB_SPLITK_FORCE = "halves,block_n,splits") and drop the picker caches."""
    def _set(spec):
        monkeypatch.setenv("SDB_SPLITK_FORCE", spec)
        ops._SPLITK_CACHE.clear()
    yield _set
    ops._SPLITK_CACHE.clear()


@pytest.mark.parametrize("spec", ["1,128,2", "1,256,3", "2,256,7", "2,160,4"])
@pytest.mark.parametrize("B,H,W,Cin,Cout", [(8, 16, 16, 320, 320), (8, 8, 8, 192, 640), (2, 16, 16, 64, 96),
                                            (3, 12, 20, 64, 200)])
def test_conv3x3_splitk(cuda, force_splitk, spec, B, H, W, Cin, Cout):
    """Split-K forward / dgrad (fp32 partial sums per K range, then sdb_splitk_finish folds them and adds bias, time
    embedding and residual) gives the same result as the fused-epilogue kernel: every K range, ragged last split, 256-row tiles."""
    force_splitk(spec)
    x = _rand((B, H, W, Cin), cuda, seed=1)
    w = _rand((3, 3, Cin, Cout), cuda, 1.0 / (9 * Cin) ** 0.5, seed=2)
    bias = torch.randn(Cout, device=cuda)
    temb = _rand((B, Cout), cuda, seed=7)
    res = _rand((B, H, W, Cout), cuda, seed=5)
    y = ops.conv3x3(x, w, bias=bias, temb=temb, residual=res)
    xt = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    wt = w.float().permute(3, 2, 0, 1).contiguous().requires_grad_(True)
    ref = torch.nn.functional.conv2d(xt, wt, bias, padding=1)
    torch.cuda.synchronize()
    _close(y, ref.permute(0, 2, 3, 1) + res.float() + temb.float()[:, None, None, :])
    dy = _rand((B, H, W, Cout), cuda, seed=3)
    ref.backward(dy.float().permute(0, 3, 1, 2))
    dx = ops.conv3x3_dgrad(dy, w)
    torch.cuda.synchronize()
    _close(dx, xt.grad.permute(0, 2, 3, 1), atol=4e-2)


@pytest.mark.parametrize("red", [False, True])
def test_splitk_slab_mode_is_deterministic(cuda, force_splitk, monkeypatch, red):
    """Default split-K (per-range slabs folded in a fixed order by sdb_splitk_finish) returns the same bits on every
    run — XLA's result is deterministic too; the atomic mode (experiment switch) is checked for parity only."""
    force_splitk("1,128,6")
    monkeypatch.setattr(ops, "_SPLITK_RED", red)
    B, H, W, Cin, Cout = 8, 8, 8, 1280, 1280
    x = _rand((B, H, W, Cin), cuda, seed=1)
    w = _rand((3, 3, Cin, Cout), cuda, 1.0 / (9 * Cin) ** 0.5, seed=2)
    bias = torch.randn(Cout, device=cuda)
    outs = [ops.conv3x3(x, w, bias=bias).clone() for _ in range(6)]
    torch.cuda.synchronize()
    ref = torch.nn.functional.conv2d(x.float().permute(0, 3, 1, 2), w.float().permute(3, 2, 0, 1), bias, padding=1)
    _close(outs[0], ref.permute(0, 2, 3, 1))
    if not red:
        for o in outs[1:]:
            assert torch.equal(o, outs[0])


@pytest.mark.parametrize("spec", ["1,128,2", "2,256,5"])
@pytest.mark.parametrize("M,K,N", [(2048, 1280, 320), (616, 768, 320), (300, 192, 200)])
def test_linear_splitk(cuda, force_splitk, spec, M, K, N):
    force_splitk(spec)
    x = _rand((M, K), cuda, seed=1)
    w = _rand((K, N), cuda, 1.0 / K ** 0.5, seed=2)
    bias = torch.randn(N, device=cuda)
    res = _rand((M, N), cuda, seed=4)
    y = ops.linear(x, w, bias=bias, residual=res)
    dy = _rand((M, N), cuda, seed=3)
    dx = ops.linear_dgrad(dy, w)
    torch.cuda.synchronize()
    _close(y, x.float() @ w.float() + bias + res.float())
    _close(dx, dy.float() @ w.float().t(), atol=4e-2)


@pytest.mark.parametrize("mt", [1, 2])
@pytest.mark.parametrize("M,K,N", [(512, 320, 320), (8192, 640, 1280), (616, 768, 320), (300, 64, 200)])
def test_linear_tile_shapes(cuda, force_tile, mt, M, K, N):
    force_tile(mt, None)
    x = _rand((M, K), cuda, seed=1)
    w = _rand((K, N), cuda, 1.0 / K ** 0.5, seed=2)
    bias = torch.randn(N, device=cuda)
    y = ops.linear(x, w, bias=bias)
    dy = _rand((M, N), cuda, seed=3)
    dx = ops.linear_dgrad(dy, w)
    torch.cuda.synchronize()
    _close(y, x.float() @ w.float() + bias)
    _close(dx, dy.float() @ w.float().t(), atol=4e-2)
