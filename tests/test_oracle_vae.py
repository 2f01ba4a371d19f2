"""CPU hardening of the VAE-encode oracle (oracle/vae.py): diffusers cannot be imported here, so — as for the UNet — a
second, structurally independent restatement from stock torch.nn modules in NCHW / OIHW (nn.Conv2d incl. the
asymmetric-pad stride-2 downsample, nn.GroupNorm(eps 1e-6), single-head attention via
F.scaled_dot_product_attention, whose 1/sqrt(C) equals the reference's C^-1/4 on q and on k) is loaded through the
Flax -> PyTorch conversion rule and must agree in fp64; plus the parameter census of the public SD VAE. The decode
half (validation sampler) is hardened the same way (nn.Upsample(nearest) for FlaxUpsample2D)."""
import numpy as np
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from oracle import vae as OV


class _Res(nn.Module):
    def __init__(self, cin, cout, g):
        super().__init__()
        self.n1, self.c1 = nn.GroupNorm(g, cin, eps=1e-6), nn.Conv2d(cin, cout, 3, padding=1)
        self.n2, self.c2 = nn.GroupNorm(g, cout, eps=1e-6), nn.Conv2d(cout, cout, 3, padding=1)
        self.sc = nn.Conv2d(cin, cout, 1) if cin != cout else None

    def forward(self, x):
        h = self.c2(F.silu(self.n2(self.c1(F.silu(self.n1(x))))))
        return h + (x if self.sc is None else self.sc(x))


class _Attn(nn.Module):
    def __init__(self, c, g):
        super().__init__()
        self.gn = nn.GroupNorm(g, c, eps=1e-6)
        self.q, self.k, self.v, self.o = (nn.Linear(c, c) for _ in range(4))

    def forward(self, x):
        B, C, H, W = x.shape
        h = self.gn(x).flatten(2).transpose(1, 2)
        a = F.scaled_dot_product_attention(self.q(h)[:, None], self.k(h)[:, None], self.v(h)[:, None])[:, 0]
        return self.o(a).transpose(1, 2).reshape(B, C, H, W) + x


def _loaders(params):
    P = {k: v.double() for k, v in params.items()}

    def conv(m, n):
        m.weight.data, m.bias.data = P[n + "/kernel"].permute(3, 2, 0, 1).contiguous(), P[n + "/bias"]

    def lin(m, n):
        m.weight.data, m.bias.data = P[n + "/kernel"].t().contiguous(), P[n + "/bias"]

    def norm(m, n):
        m.weight.data, m.bias.data = P[n + "/scale"], P[n + "/bias"]

    return conv, lin, norm


def _build_decoder(cfg, params):
    mods, g = [], cfg.norm_num_groups
    conv, lin, norm = _loaders(params)

    def res(n, cin, cout):
        r = _Res(cin, cout, g).double()
        norm(r.n1, n + "/norm1"); conv(r.c1, n + "/conv1"); norm(r.n2, n + "/norm2"); conv(r.c2, n + "/conv2")
        if r.sc is not None:
            conv(r.sc, n + "/conv_shortcut")
        return r

    rev = tuple(reversed(cfg.block_out_channels))
    L = cfg.latent_channels
    m = nn.Conv2d(L, L, 1).double(); conv(m, "post_quant_conv"); mods.append(m)
    m = nn.Conv2d(L, rev[0], 3, padding=1).double(); conv(m, "decoder/conv_in"); mods.append(m)
    ch = rev[0]
    mods.append(res("decoder/mid_block/resnets_0", ch, ch))
    a = _Attn(ch, g).double()
    norm(a.gn, "decoder/mid_block/attentions_0/group_norm")
    for mod, nm in ((a.q, "query"), (a.k, "key"), (a.v, "value"), (a.o, "proj_attn")):
        lin(mod, "decoder/mid_block/attentions_0/" + nm)
    mods.append(a)
    mods.append(res("decoder/mid_block/resnets_1", ch, ch))
    for i, c in enumerate(rev):
        for j in range(cfg.layers_per_block + 1):
            mods.append(res(f"decoder/up_blocks_{i}/resnets_{j}", ch, c)); ch = c
        if i != len(rev) - 1:
            u = nn.Conv2d(ch, ch, 3, padding=1).double(); conv(u, f"decoder/up_blocks_{i}/upsamplers_0/conv")
            mods.append(nn.Sequential(nn.Upsample(scale_factor=2, mode="nearest"), u))
    n = nn.GroupNorm(g, ch, eps=1e-6).double(); norm(n, "decoder/conv_norm_out")
    co = nn.Conv2d(ch, cfg.in_channels, 3, padding=1).double(); conv(co, "decoder/conv_out")
    return nn.Sequential(*mods, n, nn.SiLU(), co)


def _build(cfg, params):
    mods, g = [], cfg.norm_num_groups
    P = {k: v.double() for k, v in params.items()}

    def conv(m, n):
        m.weight.data, m.bias.data = P[n + "/kernel"].permute(3, 2, 0, 1).contiguous(), P[n + "/bias"]

    def lin(m, n):
        m.weight.data, m.bias.data = P[n + "/kernel"].t().contiguous(), P[n + "/bias"]

    def norm(m, n):
        m.weight.data, m.bias.data = P[n + "/scale"], P[n + "/bias"]

    def res(n, cin, cout):
        r = _Res(cin, cout, g).double()
        norm(r.n1, n + "/norm1"); conv(r.c1, n + "/conv1"); norm(r.n2, n + "/norm2"); conv(r.c2, n + "/conv2")
        if r.sc is not None:
            conv(r.sc, n + "/conv_shortcut")
        return r

    boc = cfg.block_out_channels
    m = nn.Conv2d(cfg.in_channels, boc[0], 3, padding=1).double(); conv(m, "encoder/conv_in"); mods.append(m)
    ch = boc[0]
    for i, c in enumerate(boc):
        for j in range(cfg.layers_per_block):
            mods.append(res(f"encoder/down_blocks_{i}/resnets_{j}", ch, c)); ch = c
        if i != len(boc) - 1:
            d = nn.Conv2d(ch, ch, 3, stride=2).double(); conv(d, f"encoder/down_blocks_{i}/downsamplers_0/conv")
            mods.append(nn.Sequential(nn.ZeroPad2d((0, 1, 0, 1)), d))
    mods.append(res("encoder/mid_block/resnets_0", ch, ch))
    a = _Attn(ch, g).double()
    norm(a.gn, "encoder/mid_block/attentions_0/group_norm")
    for mod, nm in ((a.q, "query"), (a.k, "key"), (a.v, "value"), (a.o, "proj_attn")):
        lin(mod, "encoder/mid_block/attentions_0/" + nm)
    mods.append(a)
    mods.append(res("encoder/mid_block/resnets_1", ch, ch))
    n = nn.GroupNorm(g, ch, eps=1e-6).double(); norm(n, "encoder/conv_norm_out")
    co = nn.Conv2d(ch, 2 * cfg.latent_channels, 3, padding=1).double(); conv(co, "encoder/conv_out")
    qc = nn.Conv2d(2 * cfg.latent_channels, 2 * cfg.latent_channels, 1).double(); conv(qc, "quant_conv")
    return nn.Sequential(*mods, n, nn.SiLU(), co, qc)


@pytest.mark.parametrize("cfg,H", [(OV.TINY_VAE, 32), (OV.VAEConfig(block_out_channels=(32, 64, 64), layers_per_block=2), 40)],
                         ids=["tiny", "3-level"])
def test_oracle_encode_equals_independent_torch_nn_restatement(cfg, H):
    params = OV.init_params(cfg, 2)
    g = torch.Generator().manual_seed(1)
    x = torch.randn(2, 3, H, H + 8, generator=g, dtype=torch.float64)
    want = OV.encode_moments(cfg, params, x, mode="fp64")
    got = _build(cfg, params)(x).permute(0, 2, 3, 1)
    assert want.shape == got.shape == (2, H >> (len(cfg.block_out_channels) - 1), (H + 8) >> (len(cfg.block_out_channels) - 1), 8)
    assert (got - want).abs().max().item() <= 1e-10 * want.abs().max().item()
    eps = torch.randn(want.shape[:-1] + (4,), generator=g, dtype=torch.float64)
    lat = OV.sample_latents(cfg, want, eps)
    mean, logvar = got[..., :4], got[..., 4:].clamp(-30, 20)
    ref = ((mean + torch.exp(0.5 * logvar) * eps) * 0.18215).permute(0, 3, 1, 2)
    assert lat.shape == ref.shape and (lat - ref).abs().max().item() < 1e-12


def test_encoder_parameter_census():
    s = OV.param_shapes(OV.SD_VAE)
    # the public SD VAE (AutoencoderKL, 83 653 863 parameters): encoder 34 163 592 + quant_conv 72
    assert sum(int(np.prod(v)) for v in s.values()) == 34_163_664 and len(s) == 108


@pytest.mark.parametrize("cfg,h", [(OV.TINY_VAE, 8), (OV.VAEConfig(block_out_channels=(32, 64, 64), layers_per_block=2), 5)],
                         ids=["tiny", "3-level"])
def test_oracle_decode_equals_independent_torch_nn_restatement(cfg, h):
    params = OV.init_decoder_params(cfg, 2)
    g = torch.Generator().manual_seed(3)
    z = torch.randn(2, 4, h, h + 3, generator=g, dtype=torch.float64)
    want = OV.decode(cfg, params, z, mode="fp64")
    got = _build_decoder(cfg, params)(z)
    up = 1 << (len(cfg.block_out_channels) - 1)
    assert want.shape == got.shape == (2, 3, h * up, (h + 3) * up)
    assert (got - want).abs().max().item() <= 1e-10 * want.abs().max().item()
    # the bf16-rounding emulation stays close to exact arithmetic (it is the yardstick of the GPU parity test)
    emu = OV.decode(cfg, params, z, mode="bf16")
    assert ((emu.double() - want).norm() / want.norm()).item() < 3e-2


def test_decoder_parameter_census():
    s = OV.decoder_param_shapes(OV.SD_VAE)
    # public SD VAE: 83 653 863 parameters = encoder + quant_conv (34 163 664) + decoder + post_quant_conv
    assert sum(int(np.prod(v)) for v in s.values()) == 83_653_863 - 34_163_664 and len(s) == 140
    assert not set(s) & set(OV.param_shapes(OV.SD_VAE))
