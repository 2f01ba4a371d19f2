"""ORACLE (test infrastructure only — never imported by the product path).

CPU restatement (PyTorch-CPU, fp64 / fp32 / bf16-emulated) of the frozen VAE ENCODE the reference runs at the top of
its step (/root/reference/training_utils.py:574-586: `frozen_vae_state.call.apply(..., sample=batch["pixel_values"],
deterministic=True, method=encode)`, `latent_dist.sample(sample_rng)`, NHWC -> NCHW, `* 0.18215`; the model is
diffusers==0.21.4 FlaxAutoencoderKL loaded at 218-222 with dtype bf16).

PARITY UNPINNED by third-party code: diffusers / flax cannot be imported here, so vae_flax.py is
restated from its published source — FlaxEncoder (conv_in, FlaxDownEncoderBlock2D x 4 with FlaxResnetBlock2D (GroupNorm 32,
eps 1e-6, swish, conv3x3, 1x1 conv_shortcut when the width changes) and FlaxDownsample2D (pad ((0,1),(0,1)) then 3x3
stride-2 VALID conv), FlaxUNetMidBlock2D (resnet, FlaxAttentionBlock: one head, scale C^-1/4 on q and k, softmax over
all H*W tokens, proj_attn, residual; resnet), conv_norm_out + swish + conv_out), quant_conv 1x1,
FlaxDiagonalGaussianDistribution (logvar clipped to [-30, 20]). It is hardened by an independent torch.nn NCHW
restatement (tests/test_oracle_vae.py). Only the encoder half and quant_conv are on the TRAINING path.

The DECODE half (post_quant_conv 1x1, FlaxDecoder: conv_in, mid block, FlaxUpDecoderBlock2D x 4 over the reversed
widths with layers_per_block + 1 resnets each and FlaxUpsample2D (nearest 2x, conv3x3) on all but the last,
conv_norm_out + swish + conv_out) is the last stage of the validation sampler
(/root/reference/models/pipeline_flax_stable_diffusion.py:246-251): `decoder_param_shapes`,
`init_decoder_params`, `decode`.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, Tuple

import torch
import torch.nn.functional as F


@dataclass
class VAEConfig:
    in_channels: int = 3
    latent_channels: int = 4
    block_out_channels: Tuple[int, ...] = (128, 256, 512, 512)
    layers_per_block: int = 2
    norm_num_groups: int = 32
    scaling_factor: float = 0.18215   # hard-coded in the reference (training_utils.py:586)


SD_VAE = VAEConfig()
TINY_VAE = VAEConfig(block_out_channels=(32, 64), layers_per_block=1)


def param_shapes(cfg: VAEConfig) -> Dict[str, tuple]:
    s: Dict[str, tuple] = {}

    def conv(n, cin, cout, k=3):
        s[n + "/kernel"] = (k, k, cin, cout)
        s[n + "/bias"] = (cout,)

    def dense(n, cin, cout):
        s[n + "/kernel"] = (cin, cout)
        s[n + "/bias"] = (cout,)

    def norm(n, c):
        s[n + "/scale"] = (c,)
        s[n + "/bias"] = (c,)

    def resnet(n, cin, cout):
        norm(n + "/norm1", cin)
        conv(n + "/conv1", cin, cout)
        norm(n + "/norm2", cout)
        conv(n + "/conv2", cout, cout)
        if cin != cout:
            conv(n + "/conv_shortcut", cin, cout, 1)

    boc = cfg.block_out_channels
    conv("encoder/conv_in", cfg.in_channels, boc[0])
    ch = boc[0]
    for i, c in enumerate(boc):
        for j in range(cfg.layers_per_block):
            resnet(f"encoder/down_blocks_{i}/resnets_{j}", ch, c)
            ch = c
        if i != len(boc) - 1:
            conv(f"encoder/down_blocks_{i}/downsamplers_0/conv", ch, ch)
    resnet("encoder/mid_block/resnets_0", ch, ch)
    a = "encoder/mid_block/attentions_0"
    norm(a + "/group_norm", ch)
    for d in ("query", "key", "value", "proj_attn"):
        dense(f"{a}/{d}", ch, ch)
    resnet("encoder/mid_block/resnets_1", ch, ch)
    norm("encoder/conv_norm_out", ch)
    conv("encoder/conv_out", ch, 2 * cfg.latent_channels)
    conv("quant_conv", 2 * cfg.latent_channels, 2 * cfg.latent_channels, 1)
    return dict(sorted(s.items()))


def init_params(cfg: VAEConfig, seed: int = 0) -> Dict[str, torch.Tensor]:
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
        elif "norm" in parent:
            a = rng.standard_normal(shape, dtype=np.float32) * np.float32(0.1)
        else:
            a = rng.standard_normal(shape, dtype=np.float32) * np.float32(0.02)
        out[name] = torch.from_numpy(a.astype(np.float32))
    return out


class _Ctx:
    def __init__(self, params, mode):
        self.mode = mode
        self.dtype = torch.float64 if mode == "fp64" else torch.float32
        self.p = params

    def rnd(self, x):
        if self.mode != "bf16":
            return x
        return x.to(torch.bfloat16).to(x.dtype)

    def w(self, name):
        return self.rnd(self.p[name].to(self.dtype))


def _conv(c, x, name, stride=1, pad=1):
    """NHWC x HWIO; pad: symmetric int, or "down" = ((0,1),(0,1)) of FlaxDownsample2D."""
    k = c.w(name + "/kernel")
    xin = c.rnd(x).permute(0, 3, 1, 2)
    if pad == "down":
        xin = F.pad(xin, (0, 1, 0, 1))
        pad = 0
    elif k.shape[0] == 1:
        pad = 0
    y = F.conv2d(xin, k.permute(3, 2, 0, 1), None, stride=stride, padding=pad)
    y = c.rnd(y.permute(0, 2, 3, 1))
    return c.rnd(y + c.w(name + "/bias"))


def _dense(c, x, name):
    return c.rnd(c.rnd(c.rnd(x) @ c.w(name + "/kernel")) + c.w(name + "/bias"))


def _group_norm(c, x, name, groups, eps=1e-6):
    B, H, W, C = x.shape
    xg = x.reshape(B, H * W, groups, C // groups)
    mean = xg.mean((1, 3), keepdim=True)
    var = ((xg * xg).mean((1, 3), keepdim=True) - mean * mean).clamp_min(0)
    y = ((xg - mean) * torch.rsqrt(var + eps)).reshape(B, H, W, C)
    return y * c.p[name + "/scale"].to(c.dtype) + c.p[name + "/bias"].to(c.dtype)


def _resnet(c, x, name, groups):
    h = _conv(c, F.silu(_group_norm(c, x, name + "/norm1", groups)), name + "/conv1")
    h = _conv(c, F.silu(_group_norm(c, h, name + "/norm2", groups)), name + "/conv2")
    if (name + "/conv_shortcut/kernel") in c.p:
        x = _conv(c, x, name + "/conv_shortcut")
    return c.rnd(h + x)


def _attention(c, x, name, groups):
    B, H, W, C = x.shape
    h = _group_norm(c, x, name + "/group_norm", groups).reshape(B, H * W, C)
    q, k, v = (_dense(c, h, f"{name}/{d}") for d in ("query", "key", "value"))
    scale = 1.0 / math.sqrt(math.sqrt(C))
    w = c.rnd(c.rnd(q * scale) @ c.rnd(k * scale).transpose(1, 2))
    w = c.rnd(torch.softmax(w, dim=-1))
    o = c.rnd(w @ v)
    o = _dense(c, o, name + "/proj_attn").reshape(B, H, W, C)
    return c.rnd(o + x)


def encode_moments(cfg: VAEConfig, params: Dict[str, torch.Tensor], pixel_values_nchw, mode="fp32"):
    """FlaxAutoencoderKL.encode -> moments NHWC [B, H/8, W/8, 2 * latent] (mean | logvar before clipping)."""
    c = _Ctx(params, mode)
    g = cfg.norm_num_groups
    boc = cfg.block_out_channels
    x = pixel_values_nchw.to(c.dtype).permute(0, 2, 3, 1)
    x = _conv(c, x, "encoder/conv_in")
    for i in range(len(boc)):
        for j in range(cfg.layers_per_block):
            x = _resnet(c, x, f"encoder/down_blocks_{i}/resnets_{j}", g)
        if i != len(boc) - 1:
            x = _conv(c, x, f"encoder/down_blocks_{i}/downsamplers_0/conv", stride=2, pad="down")
    x = _resnet(c, x, "encoder/mid_block/resnets_0", g)
    x = _attention(c, x, "encoder/mid_block/attentions_0", g)
    x = _resnet(c, x, "encoder/mid_block/resnets_1", g)
    x = _conv(c, F.silu(_group_norm(c, x, "encoder/conv_norm_out", g)), "encoder/conv_out")
    return _conv(c, x, "quant_conv")


def sample_latents(cfg: VAEConfig, moments_nhwc, eps_nhwc):
    """training_utils.py:582-586: latent_dist.sample (eps = the N(0,1) draw in the NHWC mean shape), NCHW, * 0.18215."""
    L = cfg.latent_channels
    mean, logvar = moments_nhwc[..., :L], moments_nhwc[..., L:].clamp(-30.0, 20.0)
    lat = mean + torch.exp(0.5 * logvar) * eps_nhwc.to(mean.dtype)
    return lat.permute(0, 3, 1, 2) * cfg.scaling_factor


# ---- decode half (validation sampling only) ------------------------------------------------------------------
def decoder_param_shapes(cfg: VAEConfig) -> Dict[str, tuple]:
    s: Dict[str, tuple] = {}

    def conv(n, cin, cout, k=3):
        s[n + "/kernel"] = (k, k, cin, cout)
        s[n + "/bias"] = (cout,)

    def dense(n, cin, cout):
        s[n + "/kernel"] = (cin, cout)
        s[n + "/bias"] = (cout,)

    def norm(n, c):
        s[n + "/scale"] = (c,)
        s[n + "/bias"] = (c,)

    def resnet(n, cin, cout):
        norm(n + "/norm1", cin)
        conv(n + "/conv1", cin, cout)
        norm(n + "/norm2", cout)
        conv(n + "/conv2", cout, cout)
        if cin != cout:
            conv(n + "/conv_shortcut", cin, cout, 1)

    rev = tuple(reversed(cfg.block_out_channels))
    conv("post_quant_conv", cfg.latent_channels, cfg.latent_channels, 1)
    conv("decoder/conv_in", cfg.latent_channels, rev[0])
    ch = rev[0]
    resnet("decoder/mid_block/resnets_0", ch, ch)
    a = "decoder/mid_block/attentions_0"
    norm(a + "/group_norm", ch)
    for d in ("query", "key", "value", "proj_attn"):
        dense(f"{a}/{d}", ch, ch)
    resnet("decoder/mid_block/resnets_1", ch, ch)
    for i, c in enumerate(rev):
        for j in range(cfg.layers_per_block + 1):
            resnet(f"decoder/up_blocks_{i}/resnets_{j}", ch, c)
            ch = c
        if i != len(rev) - 1:
            conv(f"decoder/up_blocks_{i}/upsamplers_0/conv", ch, ch)
    norm("decoder/conv_norm_out", ch)
    conv("decoder/conv_out", ch, cfg.in_channels)
    return dict(sorted(s.items()))


def _init(shapes, seed):
    import numpy as np
    rng = np.random.default_rng(seed)
    out = {}
    for name, shape in shapes.items():
        leaf = name.rsplit("/", 1)[1]
        parent = name.rsplit("/", 2)[-2]
        if leaf == "kernel":
            fan_in = int(np.prod(shape[:-1]))
            a = rng.standard_normal(shape, dtype=np.float32) * np.float32(1.0 / math.sqrt(fan_in))
        elif leaf == "scale":
            a = np.float32(1) + rng.standard_normal(shape, dtype=np.float32) * np.float32(0.1)
        elif "norm" in parent:
            a = rng.standard_normal(shape, dtype=np.float32) * np.float32(0.1)
        else:
            a = rng.standard_normal(shape, dtype=np.float32) * np.float32(0.02)
        out[name] = torch.from_numpy(a.astype(np.float32))
    return out


def init_decoder_params(cfg: VAEConfig, seed: int = 0) -> Dict[str, torch.Tensor]:
    """Decoder + post_quant_conv leaves, drawn from their own stream so the encoder's synthetic weights do not move."""
    return _init(decoder_param_shapes(cfg), seed + 7919)


def _upsample(c, x, name):
    """FlaxUpsample2D: jax.image.resize(method="nearest") to (2H, 2W), then conv3x3."""
    x = x.repeat_interleave(2, dim=1).repeat_interleave(2, dim=2)
    return _conv(c, x, name + "/conv")


def decode(cfg: VAEConfig, params: Dict[str, torch.Tensor], latents_nchw, mode="fp32"):
    """FlaxAutoencoderKL.decode(latents NCHW).sample -> NCHW [B, in_channels, 8h, 8w]."""
    c = _Ctx(params, mode)
    g = cfg.norm_num_groups
    rev = tuple(reversed(cfg.block_out_channels))
    x = latents_nchw.to(c.dtype).permute(0, 2, 3, 1)
    x = _conv(c, x, "post_quant_conv")
    x = _conv(c, x, "decoder/conv_in")
    x = _resnet(c, x, "decoder/mid_block/resnets_0", g)
    x = _attention(c, x, "decoder/mid_block/attentions_0", g)
    x = _resnet(c, x, "decoder/mid_block/resnets_1", g)
    for i in range(len(rev)):
        for j in range(cfg.layers_per_block + 1):
            x = _resnet(c, x, f"decoder/up_blocks_{i}/resnets_{j}", g)
        if i != len(rev) - 1:
            x = _upsample(c, x, f"decoder/up_blocks_{i}/upsamplers_0")
    x = _conv(c, F.silu(_group_norm(c, x, "decoder/conv_norm_out", g)), "decoder/conv_out")
    return x.permute(0, 3, 1, 2)
