"""GPU parity of the 4-channel-side convolutions (conv_in 4 -> C, conv_out C -> 4; csrc/conv_small.cu) against torch fp32
conv2d / autograd on the same bf16 inputs. Tolerance: fp32 accumulation, one bf16 rounding of the output
(2^-8 relative + small absolute slack); fp32 weight gradients to 1e-3 relative of the largest entry."""
import pytest
import torch
import torch.nn.functional as F

from stable_diffusion_training_b200 import ops

pytestmark = pytest.mark.gpu


def _rand(shape, dev, scale=1.0, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(torch.bfloat16).to(dev)


def _ref(x, w, bias, dy):
    """NHWC/HWIO fp32 reference through torch autograd: returns y, dx, dw (HWIO)."""
    xr = x.float().permute(0, 3, 1, 2).clone().requires_grad_(True)
    wr = w.float().permute(3, 2, 0, 1).clone().requires_grad_(True)
    y = F.conv2d(xr, wr, bias, padding=1)
    y.backward(dy.float().permute(0, 3, 1, 2))
    return y.permute(0, 2, 3, 1), xr.grad.permute(0, 2, 3, 1), wr.grad.permute(2, 3, 1, 0)


def _close(out, ref, rtol=1.0 / 128, atol=2e-2):
    err = (out.float() - ref).abs()
    bad = (err > atol + rtol * ref.abs()).sum().item()
    assert bad == 0, f"{bad} mismatches, max err {err.max().item()}"


@pytest.mark.parametrize("B,H,W,C", [(2, 8, 8, 32), (1, 5, 7, 64), (8, 64, 64, 320), (2, 16, 24, 904)])
def test_conv_in(cuda, B, H, W, C):
    x = _rand((B, H, W, 4), cuda, seed=1)
    w = _rand((3, 3, 4, C), cuda, 1 / 6.0, seed=2)
    bias = torch.randn(C, device=cuda)
    dy = _rand((B, H, W, C), cuda, seed=3)
    y = ops.conv3x3_c4in(x, w, bias=bias)
    dw = ops.conv3x3_c4_wgrad(x, dy, out=torch.empty((3, 3, 4, C), device=cuda))
    torch.cuda.synchronize()
    ry, _, rdw = _ref(x, w, bias, dy)
    _close(y, ry)
    assert (dw - rdw).abs().max().item() <= 1e-3 * rdw.abs().max().item() + 1e-3


@pytest.mark.parametrize("B,H,W,C", [(2, 8, 8, 32), (1, 5, 7, 64), (8, 64, 64, 320), (2, 16, 24, 904)])
def test_conv_out(cuda, B, H, W, C):
    x = _rand((B, H, W, C), cuda, seed=1)
    w = _rand((3, 3, C, 4), cuda, 1.0 / (9 * C) ** 0.5, seed=2)
    bias = torch.randn(4, device=cuda)
    dy = _rand((B, H, W, 4), cuda, seed=3)
    y = ops.conv3x3_c4out(x, w, bias=bias)
    dx = ops.conv3x3_c4out_dgrad(dy, w)
    dw = ops.conv3x3_c4_wgrad(x, dy, out=torch.empty((3, 3, C, 4), device=cuda))
    torch.cuda.synchronize()
    ry, rdx, rdw = _ref(x, w, bias, dy)
    _close(y, ry)
    _close(dx, rdx)
    assert (dw - rdw).abs().max().item() <= 1e-3 * rdw.abs().max().item() + 1e-3


def test_bad_channel_count_is_an_error(cuda):
    from stable_diffusion_training_b200._lib import SdbError
    x = _rand((1, 4, 4, 4), cuda)
    w = _rand((3, 3, 4, 12), cuda)
    with pytest.raises(SdbError, match="multiple of 8"):
        ops.conv3x3_c4in(x, w)
