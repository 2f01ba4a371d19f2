"""ORACLE (test infrastructure only — never imported by the product path).

CPU restatement (PyTorch-CPU tensors + autograd, fp64 / fp32 / bf16-emulated) of the arithmetic that
/root/reference/training_utils.py:678-684 runs through `unet.apply(...)`: diffusers==0.21.4 FlaxUNet2DConditionModel
(requirements.txt:1; NOT vendored under /root/reference, restated here from its published source — SURVEY.md
appendix B) including the reference's own patch /root/reference/key_chunk_patch.patch:1-9 (key_chunk_size =
flatten_latent_dim), plus the loss of training_utils.py:688-709.

PARITY UNPINNED for the UNet: the reference ships no tests / golden vectors for this path and jax/flax/diffusers
cannot be imported here, so this file is pinned only by (a) the public parameter census (SD-1.5: 686 leaves,
859 520 964 parameters; SD-2.1: 865 910 724) and (b) self-consistency checks in tests/test_oracle_unet.py.

Layout conventions follow Flax: activations NHWC, Dense kernel (in, out), Conv kernel HWIO, params = {"a/b/c": tensor}.
Rounding model (`mode`):
  "fp64" / "fp32": everything in that dtype (the "truth" / plain fp32 restatement).
  "bf16": Flax dtype=bfloat16 emulation — values are rounded to bf16 at the points where the Flax modules round
          (SURVEY.md B.5): Conv/Dense inputs+kernels+bias cast to bf16, output bf16, bias add in bf16; LayerNorm output
          bf16; GroupNorm output fp32; attention internals bf16. Rounding is straight-through for autograd.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, Tuple

import torch
import torch.nn.functional as F


@dataclass
class UNetConfig:
    in_channels: int = 4
    out_channels: int = 4
    block_out_channels: Tuple[int, ...] = (320, 640, 1280, 1280)
    down_block_types: Tuple[str, ...] = ("CrossAttnDownBlock2D", "CrossAttnDownBlock2D", "CrossAttnDownBlock2D", "DownBlock2D")
    up_block_types: Tuple[str, ...] = ("UpBlock2D", "CrossAttnUpBlock2D", "CrossAttnUpBlock2D", "CrossAttnUpBlock2D")
    layers_per_block: int = 2
    attention_head_dim: object = 8          # Flax quirk: this is the NUMBER OF HEADS (int or per-block tuple)
    cross_attention_dim: int = 768
    use_linear_projection: bool = False
    flip_sin_to_cos: bool = True
    freq_shift: int = 0

    def heads(self, i: int) -> int:
        a = self.attention_head_dim
        return a if isinstance(a, int) else a[i]


SD15 = UNetConfig()
SD21 = UNetConfig(attention_head_dim=(5, 10, 20, 20), cross_attention_dim=1024, use_linear_projection=True)
TINY = UNetConfig(block_out_channels=(32, 64), down_block_types=("CrossAttnDownBlock2D", "DownBlock2D"),
                  up_block_types=("UpBlock2D", "CrossAttnUpBlock2D"), layers_per_block=1, attention_head_dim=2,
                  cross_attention_dim=32)


# ----------------------------------------------------------------------------------------------------------------
# parameter tree
def _resnet_shapes(prefix, cin, cout, temb_dim, out):
    out[f"{prefix}/norm1/scale"] = (cin,)
    out[f"{prefix}/norm1/bias"] = (cin,)
    out[f"{prefix}/conv1/kernel"] = (3, 3, cin, cout)
    out[f"{prefix}/conv1/bias"] = (cout,)
    out[f"{prefix}/time_emb_proj/kernel"] = (temb_dim, cout)
    out[f"{prefix}/time_emb_proj/bias"] = (cout,)
    out[f"{prefix}/norm2/scale"] = (cout,)
    out[f"{prefix}/norm2/bias"] = (cout,)
    out[f"{prefix}/conv2/kernel"] = (3, 3, cout, cout)
    out[f"{prefix}/conv2/bias"] = (cout,)
    if cin != cout:
        out[f"{prefix}/conv_shortcut/kernel"] = (1, 1, cin, cout)
        out[f"{prefix}/conv_shortcut/bias"] = (cout,)


def _transformer_shapes(prefix, c, ctx_dim, linear_proj, out):
    out[f"{prefix}/norm/scale"] = (c,)
    out[f"{prefix}/norm/bias"] = (c,)
    pshape = (c, c) if linear_proj else (1, 1, c, c)
    out[f"{prefix}/proj_in/kernel"] = pshape
    out[f"{prefix}/proj_in/bias"] = (c,)
    t = f"{prefix}/transformer_blocks_0"
    for n in ("norm1", "norm2", "norm3"):
        out[f"{t}/{n}/scale"] = (c,)
        out[f"{t}/{n}/bias"] = (c,)
    for a, kd in (("attn1", c), ("attn2", ctx_dim)):
        out[f"{t}/{a}/to_q/kernel"] = (c, c)
        out[f"{t}/{a}/to_k/kernel"] = (kd, c)
        out[f"{t}/{a}/to_v/kernel"] = (kd, c)
        out[f"{t}/{a}/to_out_0/kernel"] = (c, c)
        out[f"{t}/{a}/to_out_0/bias"] = (c,)
    out[f"{t}/ff/net_0/proj/kernel"] = (c, 8 * c)
    out[f"{t}/ff/net_0/proj/bias"] = (8 * c,)
    out[f"{t}/ff/net_2/kernel"] = (4 * c, c)
    out[f"{t}/ff/net_2/bias"] = (c,)
    out[f"{prefix}/proj_out/kernel"] = pshape
    out[f"{prefix}/proj_out/bias"] = (c,)


def param_shapes(cfg: UNetConfig) -> Dict[str, tuple]:
    out: Dict[str, tuple] = {}
    boc = cfg.block_out_channels
    c0, temb = boc[0], 4 * boc[0]
    out["conv_in/kernel"] = (3, 3, cfg.in_channels, c0)
    out["conv_in/bias"] = (c0,)
    out["time_embedding/linear_1/kernel"] = (c0, temb)
    out["time_embedding/linear_1/bias"] = (temb,)
    out["time_embedding/linear_2/kernel"] = (temb, temb)
    out["time_embedding/linear_2/bias"] = (temb,)
    skips = [c0]
    ch = c0
    for i, btype in enumerate(cfg.down_block_types):
        cout = boc[i]
        for j in range(cfg.layers_per_block):
            _resnet_shapes(f"down_blocks_{i}/resnets_{j}", ch, cout, temb, out)
            ch = cout
            if btype == "CrossAttnDownBlock2D":
                _transformer_shapes(f"down_blocks_{i}/attentions_{j}", ch, cfg.cross_attention_dim,
                                    cfg.use_linear_projection, out)
            skips.append(ch)
        if i != len(boc) - 1:
            out[f"down_blocks_{i}/downsamplers_0/conv/kernel"] = (3, 3, ch, ch)
            out[f"down_blocks_{i}/downsamplers_0/conv/bias"] = (ch,)
            skips.append(ch)
    _resnet_shapes("mid_block/resnets_0", ch, ch, temb, out)
    _transformer_shapes("mid_block/attentions_0", ch, cfg.cross_attention_dim, cfg.use_linear_projection, out)
    _resnet_shapes("mid_block/resnets_1", ch, ch, temb, out)
    rev = list(reversed(boc))
    for i, btype in enumerate(cfg.up_block_types):
        cout = rev[i]
        for j in range(cfg.layers_per_block + 1):
            cin = ch + skips.pop()
            _resnet_shapes(f"up_blocks_{i}/resnets_{j}", cin, cout, temb, out)
            ch = cout
            if btype == "CrossAttnUpBlock2D":
                _transformer_shapes(f"up_blocks_{i}/attentions_{j}", ch, cfg.cross_attention_dim,
                                    cfg.use_linear_projection, out)
        if i != len(boc) - 1:
            out[f"up_blocks_{i}/upsamplers_0/conv/kernel"] = (3, 3, ch, ch)
            out[f"up_blocks_{i}/upsamplers_0/conv/bias"] = (ch,)
    assert not skips
    out["conv_norm_out/scale"] = (c0,)
    out["conv_norm_out/bias"] = (c0,)
    out["conv_out/kernel"] = (3, 3, c0, cfg.out_channels)
    out["conv_out/bias"] = (cfg.out_channels,)
    return dict(sorted(out.items()))


# ----------------------------------------------------------------------------------------------------------------
class _Ctx:
    def __init__(self, params, mode):
        self.mode = mode
        self.dtype = torch.float64 if mode == "fp64" else torch.float32
        self.p = params

    def rnd(self, x):
        """Flax bf16 rounding point (straight-through gradient); identity in fp64/fp32 modes."""
        if self.mode != "bf16":
            return x
        return x + (x.detach().to(torch.bfloat16).to(x.dtype) - x.detach())

    def w(self, name):
        return self.rnd(self.p[name].to(self.dtype))


def _conv(c: _Ctx, x, name, stride=1):
    """flax nn.Conv(dtype): NHWC x HWIO cross-correlation, explicit padding ((1,1),(1,1)) for 3x3, none for 1x1."""
    k = c.w(name + "/kernel")
    pad = 1 if k.shape[0] == 3 else 0
    y = F.conv2d(c.rnd(x).permute(0, 3, 1, 2), k.permute(3, 2, 0, 1), None, stride=stride, padding=pad)
    y = c.rnd(y.permute(0, 2, 3, 1))
    return c.rnd(y + c.w(name + "/bias"))


def _dense(c: _Ctx, x, name, bias=True):
    y = c.rnd(c.rnd(x) @ c.w(name + "/kernel"))
    return c.rnd(y + c.w(name + "/bias")) if bias else y


def _norm_stats(x, dims):
    mean = x.mean(dims, keepdim=True)
    var = ((x * x).mean(dims, keepdim=True) - mean * mean).clamp_min(0)  # use_fast_variance
    return mean, var


def _group_norm(c: _Ctx, x, name, groups=32, eps=1e-5):
    """flax nn.GroupNorm(32, eps 1e-5), no dtype -> fp32 statistics and fp32 output."""
    B, H, W, C = x.shape
    xg = x.reshape(B, H * W, groups, C // groups)
    mean, var = _norm_stats(xg, (1, 3))
    y = ((xg - mean) * torch.rsqrt(var + eps)).reshape(B, H, W, C)
    return y * c.p[name + "/scale"].to(c.dtype) + c.p[name + "/bias"].to(c.dtype)


def _layer_norm(c: _Ctx, x, name, eps=1e-5):
    mean, var = _norm_stats(x, (-1,))
    y = (x - mean) * torch.rsqrt(var + eps)
    return c.rnd(y * c.p[name + "/scale"].to(c.dtype) + c.p[name + "/bias"].to(c.dtype))


def key_chunk_starts(sq: int, skv: int):
    """Patched reference: key_chunk_size = flatten_latent_dim = Sq, then min(Sq, Skv); chunk starts arange(0, Skv, kc),
    lax.dynamic_slice clamps each start to Skv - kc."""
    kc = min(sq, skv)
    return kc, [min(s, skv - kc) for s in range(0, skv, kc)]


def key_multiplicity(sq: int, skv: int):
    kc, starts = key_chunk_starts(sq, skv)
    m = [0] * skv
    for s in starts:
        for k in range(s, s + kc):
            m[k] += 1
    return m


def _chunked_attention(c: _Ctx, q, k, v):
    """jax_memory_efficient_attention + _query_chunk_attention [diffusers 0.21.4 attention_flax.py] with the patch.
    q [BH, Sq, d], k/v [BH, Skv, d]. Query chunking is row-independent and omitted; key chunking is literal."""
    d = q.shape[-1]
    kc, starts = key_chunk_starts(q.shape[1], k.shape[1])
    q = c.rnd(q / math.sqrt(d))
    vals, wts, maxs = [], [], []
    for s in starts:
        kk, vv = k[:, s:s + kc], v[:, s:s + kc]
        w = c.rnd(q @ kk.transpose(1, 2))
        mx = w.max(-1, keepdim=True).values.detach()
        e = c.rnd(torch.exp(w - mx))
        vals.append(c.rnd(e @ vv))
        wts.append(c.rnd(e.sum(-1)))
        maxs.append(mx.squeeze(-1))
    cm = torch.stack(maxs)                      # [n, BH, Sq]
    gm = cm.max(0, keepdim=True).values
    diff = c.rnd(torch.exp(cm - gm))
    allv = c.rnd((torch.stack(vals) * diff.unsqueeze(-1)).sum(0))
    allw = c.rnd((torch.stack(wts) * diff).sum(0)).unsqueeze(-1)
    return c.rnd(allv / allw)


def _attention(c: _Ctx, x, ctx, name, heads):
    B, S, C = x.shape
    d = C // heads
    q = _dense(c, x, name + "/to_q", bias=False)
    k = _dense(c, ctx, name + "/to_k", bias=False)
    v = _dense(c, ctx, name + "/to_v", bias=False)

    def split(t):
        return t.reshape(B, -1, heads, d).permute(0, 2, 1, 3).reshape(B * heads, -1, d)

    o = _chunked_attention(c, split(q), split(k), split(v))
    o = o.reshape(B, heads, S, d).permute(0, 2, 1, 3).reshape(B, S, C)
    return _dense(c, o, name + "/to_out_0")


def _gelu_tanh(z):
    return 0.5 * z * (1.0 + torch.tanh(math.sqrt(2.0 / math.pi) * (z + 0.044715 * z ** 3)))


def _transformer(c: _Ctx, x, ctx, name, heads, linear_proj):
    B, H, W, C = x.shape
    r = x
    h = _group_norm(c, x, name + "/norm")
    if linear_proj:
        h = _dense(c, h.reshape(B, H * W, C), name + "/proj_in")
    else:
        h = _conv(c, h, name + "/proj_in").reshape(B, H * W, C)
    t = name + "/transformer_blocks_0"
    hn = _layer_norm(c, h, t + "/norm1")
    h = c.rnd(h + _attention(c, hn, hn, t + "/attn1", heads))
    h = c.rnd(h + _attention(c, _layer_norm(c, h, t + "/norm2"), ctx, t + "/attn2", heads))
    u = _dense(c, _layer_norm(c, h, t + "/norm3"), t + "/ff/net_0/proj")
    lin, gate = u.chunk(2, dim=-1)
    h = c.rnd(h + _dense(c, c.rnd(lin * _gelu_tanh(gate)), t + "/ff/net_2"))
    if linear_proj:
        h = _dense(c, h, name + "/proj_out").reshape(B, H, W, C)
    else:
        h = _conv(c, h.reshape(B, H, W, C), name + "/proj_out")
    return c.rnd(h + r)


def _resnet(c: _Ctx, x, temb, name):
    h = _conv(c, F.silu(_group_norm(c, x, name + "/norm1")), name + "/conv1")
    t = _dense(c, F.silu(temb), name + "/time_emb_proj")
    h = c.rnd(h + t[:, None, None, :])
    h = _conv(c, F.silu(_group_norm(c, h, name + "/norm2")), name + "/conv2")
    if (name + "/conv_shortcut/kernel") in c.p:
        x = _conv(c, x, name + "/conv_shortcut")
    return c.rnd(h + x)


def timestep_embedding(t, dim, flip_sin_to_cos=True, freq_shift=0, dtype=torch.float32):
    half = dim // 2
    f = torch.exp(-math.log(10000.0) * torch.arange(half, dtype=dtype) / (half - freq_shift))
    arg = t.to(dtype)[:, None] * f[None]
    return torch.cat([arg.cos(), arg.sin()] if flip_sin_to_cos else [arg.sin(), arg.cos()], dim=1)


def unet_forward(cfg: UNetConfig, params: Dict[str, torch.Tensor], sample_nchw, timesteps, context, mode="fp32"):
    """FlaxUNet2DConditionModel.__call__ (train=True, dropout 0): sample [B,C,H,W], timesteps int [B],
    context [B,L,D] -> prediction [B,C,H,W]."""
    c = _Ctx(params, mode)
    x = sample_nchw.to(c.dtype).permute(0, 2, 3, 1)
    ctx = context.to(c.dtype)
    boc = cfg.block_out_channels
    temb = timestep_embedding(timesteps, boc[0], cfg.flip_sin_to_cos, cfg.freq_shift, c.dtype)
    temb = _dense(c, temb, "time_embedding/linear_1")
    temb = _dense(c, F.silu(temb), "time_embedding/linear_2")
    x = _conv(c, x, "conv_in")
    skips = [x]
    for i, btype in enumerate(cfg.down_block_types):
        for j in range(cfg.layers_per_block):
            x = _resnet(c, x, temb, f"down_blocks_{i}/resnets_{j}")
            if btype == "CrossAttnDownBlock2D":
                x = _transformer(c, x, ctx, f"down_blocks_{i}/attentions_{j}", cfg.heads(i), cfg.use_linear_projection)
            skips.append(x)
        if i != len(boc) - 1:
            x = _conv(c, x, f"down_blocks_{i}/downsamplers_0/conv", stride=2)
            skips.append(x)
    x = _resnet(c, x, temb, "mid_block/resnets_0")
    x = _transformer(c, x, ctx, "mid_block/attentions_0", cfg.heads(len(boc) - 1), cfg.use_linear_projection)
    x = _resnet(c, x, temb, "mid_block/resnets_1")
    for i, btype in enumerate(cfg.up_block_types):
        for j in range(cfg.layers_per_block + 1):
            x = torch.cat([x, skips.pop()], dim=-1)
            x = _resnet(c, x, temb, f"up_blocks_{i}/resnets_{j}")
            if btype == "CrossAttnUpBlock2D":
                hi = len(boc) - 1 - i  # reversed_attention_head_dim[i]
                x = _transformer(c, x, ctx, f"up_blocks_{i}/attentions_{j}", cfg.heads(hi), cfg.use_linear_projection)
        if i != len(boc) - 1:
            x = x.repeat_interleave(2, dim=1).repeat_interleave(2, dim=2)  # jax.image.resize(nearest)
            x = _conv(c, x, f"up_blocks_{i}/upsamplers_0/conv")
    x = _conv(c, F.silu(_group_norm(c, x, "conv_norm_out")), "conv_out")
    return x.permute(0, 3, 1, 2)


def init_params(cfg: UNetConfig, seed: int = 0) -> Dict[str, torch.Tensor]:
    """Synthetic weights of SURVEY.md 8(d): NumPy default_rng(seed) in sorted-path order; kernels N(0, 1/fan_in),
    biases N(0, 0.02^2), norm scale 1 + N(0, 0.1^2), norm bias N(0, 0.1^2). fp32."""
    import numpy as np
    rng = np.random.default_rng(seed)
    out = {}
    for name, shape in param_shapes(cfg).items():
        leaf = name.rsplit("/", 1)[1]
        parent = name.rsplit("/", 2)[-2]
        if leaf == "kernel":
            fan_in = int(np.prod(shape[:-1]))
            a = rng.standard_normal(shape, dtype=np.float32) * np.float32(1.0 / math.sqrt(fan_in))
        elif leaf == "scale":
            a = np.float32(1) + rng.standard_normal(shape, dtype=np.float32) * np.float32(0.1)
        elif parent.startswith("norm") or parent == "conv_norm_out":
            a = rng.standard_normal(shape, dtype=np.float32) * np.float32(0.1)
        else:
            a = rng.standard_normal(shape, dtype=np.float32) * np.float32(0.02)
        out[name] = torch.from_numpy(a.astype(np.float32))
    return out


def loss_and_grads(cfg, params, latents, noise, timesteps, context, alphas_cumprod, prediction_type="epsilon",
                   mode="fp32", snr_weight=None):
    """compute_loss + value_and_grad of training_utils.py:570-729 restricted to the UNet (latents given):
    add_noise -> UNet -> target (epsilon | v) -> mean((target - pred)^2 [* w_b]). Returns (loss, grads dict, pred)."""
    from . import sched
    import numpy as np
    dt = torch.float64 if mode == "fp64" else torch.float32
    lat, nz, t = latents.numpy(), noise.numpy(), timesteps.numpy()
    noisy = torch.from_numpy(sched.add_noise(alphas_cumprod, lat, nz, t))
    if prediction_type == "epsilon":
        target = noise
    elif prediction_type == "v_prediction":
        target = torch.from_numpy(sched.get_velocity(alphas_cumprod, lat, nz, t))
    else:
        raise ValueError(f"Unknown prediction type {prediction_type}")
    leaves = {k: v.detach().clone().to(dt).requires_grad_(True) for k, v in params.items()}
    pred = unet_forward(cfg, leaves, noisy, timesteps, context, mode)
    l = (target.to(pred.dtype) - pred) ** 2
    if snr_weight is not None:
        l = l * torch.as_tensor(np.asarray(snr_weight)).to(pred.dtype).view(-1, 1, 1, 1)
    loss = l.mean()
    loss.backward()
    return loss.detach(), {k: v.grad for k, v in leaves.items()}, pred.detach()
