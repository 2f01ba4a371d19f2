"""GPU parity of GroupNorm(+SiLU) / LayerNorm / GEGLU / attention / data-movement kernels against plain torch fp32
references on the same bf16 inputs (floating-point kernels: tolerance stated per check = bf16 output rounding
2^-8 relative + absolute slack)."""
import math

import pytest
import torch
import torch.nn.functional as F

from stable_diffusion_training_b200 import ops

pytestmark = pytest.mark.gpu


def _rand(shape, dev, scale=1.0, seed=0, dtype=torch.bfloat16):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(dtype).to(dev)


def _close(out, ref, rtol=1.0 / 100, atol=2e-2, what=""):
    err = (out.float() - ref).abs()
    tol = atol + rtol * ref.abs()
    bad = (err > tol).sum().item()
    assert bad == 0, f"{what}: {bad} mismatches, max err {err.max().item()} (ref max {ref.abs().max().item()})"


@pytest.mark.parametrize("B,HW,C", [(2, 1024, 32), (2, 256, 96), (8, 4096, 320), (8, 64, 2560), (4, 1024, 960),
                                    (3, 256, 1920), (2, 64, 1280), (2, 1024, 640), (2, 1024, 128)])
@pytest.mark.parametrize("silu", [True, False])
def test_groupnorm_fwd_bwd(cuda, B, HW, C, silu):
    x = _rand((B, HW, C), cuda, 1.5, 1) + 0.3
    gamma = 1 + 0.1 * torch.randn(C, device=cuda)
    beta = 0.1 * torch.randn(C, device=cuda)
    dy = _rand((B, HW, C), cuda, 1.0, 2)
    add = _rand((B, HW, C), cuda, 1.0, 3)
    y, stats = ops.groupnorm_fwd(x, gamma, beta, silu=silu)
    dgamma, dbeta = torch.empty(C, device=cuda), torch.empty(C, device=cuda)
    dx = ops.groupnorm_bwd(x, dy, stats, gamma, beta, dgamma, dbeta, silu=silu, dx_add=add)
    torch.cuda.synchronize()
    xr = x.float().requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    yr = F.group_norm(xr.transpose(1, 2), 32, gr, br, eps=1e-5).transpose(1, 2)
    if silu:
        yr = F.silu(yr)
    yr.backward(dy.float())
    _close(y, yr.detach(), what="y")
    _close(dx, xr.grad + add.float(), atol=3e-2, what="dx")
    assert torch.allclose(dgamma, gr.grad, rtol=2e-3, atol=2e-3 * gr.grad.abs().max().item())
    assert torch.allclose(dbeta, br.grad, rtol=2e-3, atol=2e-3 * br.grad.abs().max().item())


@pytest.mark.parametrize("rows,C", [(2048, 32), (512, 64), (32768, 320), (8192, 640), (2048, 1280), (77, 1024)])
def test_layernorm_fwd_bwd(cuda, rows, C):
    x = _rand((rows, C), cuda, 2.0, 1) + 0.5
    gamma = 1 + 0.1 * torch.randn(C, device=cuda)
    beta = 0.1 * torch.randn(C, device=cuda)
    dy = _rand((rows, C), cuda, 1.0, 2)
    y, stats = ops.layernorm_fwd(x, gamma, beta)
    dgamma, dbeta = torch.empty(C, device=cuda), torch.empty(C, device=cuda)
    dx = ops.layernorm_bwd(x, dy, stats, gamma, dgamma, dbeta)
    torch.cuda.synchronize()
    xr = x.float().requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    yr = F.layer_norm(xr, (C,), gr, br, eps=1e-5)
    yr.backward(dy.float())
    _close(y, yr.detach(), what="y")
    _close(dx, xr.grad, what="dx")
    assert torch.allclose(dgamma, gr.grad, rtol=2e-3, atol=2e-3 * gr.grad.abs().max().item())
    assert torch.allclose(dbeta, br.grad, rtol=2e-3, atol=2e-3 * br.grad.abs().max().item())


@pytest.mark.parametrize("rows,H", [(2048, 128), (32768, 1280), (512, 5120)])
def test_geglu(cuda, rows, H):
    u = _rand((rows, 2 * H), cuda, 1.5, 1)
    dout = _rand((rows, H), cuda, 1.0, 2)
    out = ops.geglu_fwd(u)
    du = ops.geglu_bwd(u, dout)
    torch.cuda.synchronize()
    ur = u.float().requires_grad_(True)
    ref = ur[:, :H] * F.gelu(ur[:, H:], approximate="tanh")
    ref.backward(dout.float())
    _close(out, ref.detach(), what="out")
    _close(du, ur.grad, what="du")


def _attn_ref(q, k, v, heads, kbias2):
    B, Sq, C = q.shape
    d = C // heads
    qh = q.view(B, Sq, heads, d).transpose(1, 2)
    kh = k.view(B, -1, heads, d).transpose(1, 2)
    vh = v.view(B, -1, heads, d).transpose(1, 2)
    s = (qh @ kh.transpose(-1, -2)) / math.sqrt(d)
    if kbias2 is not None:
        s = s + kbias2.float() * math.log(2.0)
    p = torch.softmax(s, dim=-1)
    return (p @ vh).transpose(1, 2).reshape(B, Sq, C)


@pytest.mark.parametrize("B,heads,Sq,Skv,d,bias", [
    (1, 2, 128, 128, 64, False), (2, 2, 256, 256, 16, False), (2, 2, 1024, 77, 32, False), (2, 8, 1024, 1024, 40, False),
    (1, 8, 4096, 4096, 40, False), (2, 8, 1024, 77, 80, False), (2, 8, 256, 256, 160, False), (2, 8, 256, 77, 160, False),
    (2, 8, 64, 64, 160, False), (2, 8, 64, 77, 160, True), (2, 5, 576, 576, 64, False), (2, 4, 144, 227, 64, True),
    (1, 2, 300, 200, 40, True), (1, 2, 300, 1100, 40, False), (1, 1, 130, 2048, 40, True), (1, 2, 1100, 300, 80, False),
    (1, 3, 200, 1000, 160, False)])
def test_attention_fwd_bwd(cuda, B, heads, Sq, Skv, d, bias):
    C = heads * d
    # q/k/v as column slices of a fused buffer -> exercises row strides
    # (batch stride must be S * row stride: the kernels take [B*S, ld] buffers)
    if Sq == Skv:
        qkv = _rand((B, Sq, 3 * C), cuda, 1.0, 1)
        q, k, v = qkv[:, :, :C], qkv[:, :, C:2 * C], qkv[:, :, 2 * C:]
    else:
        q = _rand((B, Sq, C), cuda, 1.0, 1)
        kv = _rand((B, Skv, 2 * C), cuda, 1.0, 3)
        k, v = kv[:, :, :C], kv[:, :, C:]
    dO = _rand((B, Sq, C), cuda, 1.0, 2)
    kb = None
    if bias:
        kb = torch.zeros(Skv, device=cuda)
        kb[Skv // 4: Skv // 2] = 1.0  # multiplicity 2
    o, lse2 = ops.attention_fwd(q, k, v, heads, kbias2=kb)
    dq, dk, dv = ops.attention_bwd(q, k, v, o, dO, lse2, heads, kbias2=kb)
    torch.cuda.synchronize()
    qr, kr, vr = (t.float().clone().requires_grad_(True) for t in (q, k, v))
    ref = _attn_ref(qr, kr, vr, heads, kb)
    ref.backward(dO.float())
    _close(o, ref.detach(), atol=1e-2, what="o")
    gscale = max(1.0, qr.grad.abs().max().item(), kr.grad.abs().max().item(), vr.grad.abs().max().item())
    _close(dq, qr.grad, rtol=2e-2, atol=2e-2 * gscale, what="dq")
    _close(dk, kr.grad, rtol=2e-2, atol=2e-2 * gscale, what="dk")
    _close(dv, vr.grad, rtol=2e-2, atol=2e-2 * gscale, what="dv")


def test_data_movement(cuda):
    x = _rand((2, 8, 8, 64), cuda, 1.0, 1)
    y = ops.upsample2x_fwd(x)
    assert torch.equal(y, x.repeat_interleave(2, 1).repeat_interleave(2, 2))
    dy = _rand((2, 16, 16, 64), cuda, 1.0, 2)
    dx = ops.upsample2x_bwd(dy)
    ref = dy.float().view(2, 8, 2, 8, 2, 64).sum((2, 4))
    _close(dx, ref, what="upsample bwd")
    a, b = _rand((512, 64), cuda, 1.0, 3), _rand((512, 32), cuda, 1.0, 4)
    cat = torch.zeros((512, 96), device=cuda, dtype=torch.bfloat16)
    ops.copy2d(a, cat[:, :64])
    ops.copy2d(b, cat[:, 64:])
    assert torch.equal(cat, torch.cat([a, b], 1))
    acc = a.clone()
    ops.copy2d(cat[:, :64], acc, accumulate=True)
    _close(acc, 2 * a.float(), what="accumulate")
    big = _rand((5000, 320), cuda, 1.0, 5)
    out = torch.empty(320, device=cuda)
    ops.colsum(big, out)
    assert torch.allclose(out, big.float().sum(0), rtol=1e-4, atol=1e-2)
    wide = _rand((2048, 10240), cuda, 1.0, 7)  # short and wide: the FF gradient shape
    outw = torch.empty(10240, device=cuda)
    ops.colsum(wide, outw)
    assert torch.allclose(outw, wide.float().sum(0), rtol=1e-4, atol=2e-2)
    seg = torch.empty((5, 320), device=cuda, dtype=torch.bfloat16)
    ops.colsum(big, out, nseg=5, out_seg=seg)
    assert torch.allclose(out, big.float().sum(0), rtol=1e-4, atol=1e-2)
    _close(seg, big.float().view(5, 1000, 320).sum(1), what="segmented colsum")
    t = torch.tensor([0, 1, 500, 999], dtype=torch.int32, device=cuda)
    emb = ops.timestep_embedding(t, 320)
    f = torch.exp(-math.log(10000.0) * torch.arange(160, device=cuda, dtype=torch.float32) / 160)
    arg = t.float()[:, None] * f[None]
    _close(emb, torch.cat([arg.cos(), arg.sin()], 1), atol=1e-2, what="timestep")
    s = _rand((8, 1280), cuda, 2.0, 6)
    _close(ops.silu_fwd(s), F.silu(s.float()), what="silu")
    src = torch.randn(2, 4, 16, 16, device=cuda)
    assert torch.equal(ops.nchw_to_nhwc_bf16(src), src.permute(0, 2, 3, 1).to(torch.bfloat16))
    w = torch.randn(1000, 37, device=cuda)
    assert torch.equal(ops.cast_bf16(w), w.to(torch.bfloat16))
