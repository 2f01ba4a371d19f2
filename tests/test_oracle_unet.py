"""CPU checks of the UNet oracle (oracle/unet.py). The reference ships no tests or golden vectors for the UNet and
diffusers / flax / jax cannot be imported here (DESIGN.md section 3 records the probe), so the oracle is hardened with
what CAN be pinned:

 1. a second, structurally independent restatement of the forward built from stock torch.nn modules in NCHW / OIHW
    (nn.Conv2d, nn.GroupNorm, nn.LayerNorm, nn.Linear, F.scaled_dot_product_attention, F.gelu(approximate="tanh"),
    F.interpolate(nearest)) organised like the PyTorch UNet2DConditionModel, loaded from the Flax-named parameters with
    the published Flax -> PyTorch conversion rule (conv kernel HWIO -> OIHW, Dense kernel (in,out) -> weight (out,in),
    scale -> weight). It shares no code with oracle/unet.py and catches layout / ordering / head-split slips;
 2. the identity "literal clamped key chunks == plain softmax with a log(multiplicity) key bias"
    (/root/reference/key_chunk_patch.patch:5-6) on which the CUDA kernels' kbias2 input rests;
 3. the public parameter census (SD-1.5: 686 leaves / 859 520 964 parameters; SD-2.1: 865 910 724) and that the
    product's parameter tree is the oracle's, name by name.

What stays unpinnable: diffusers 0.21.4's exact op order inside a module and the bf16 rounding points of the Flax
modules — they are restated from the published source, not executed.
"""
import math

import numpy as np
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from oracle import unet as OU


# ---------------------------------------------------------------------------------------------------------------
# independent restatement: module classes in NCHW, PyTorch parameter layouts
class _Attention(nn.Module):
    def __init__(self, c, ctx_dim, heads):
        super().__init__()
        self.heads = heads
        self.to_q = nn.Linear(c, c, bias=False)
        self.to_k = nn.Linear(ctx_dim, c, bias=False)
        self.to_v = nn.Linear(ctx_dim, c, bias=False)
        self.to_out = nn.Linear(c, c)

    def forward(self, x, ctx):
        B, S, C = x.shape
        h = self.heads
        q = self.to_q(x).view(B, S, h, C // h).transpose(1, 2)
        k = self.to_k(ctx).view(B, -1, h, C // h).transpose(1, 2)
        v = self.to_v(ctx).view(B, -1, h, C // h).transpose(1, 2)
        o = F.scaled_dot_product_attention(q, k, v)  # softmax(q k^T / sqrt(d)) v
        return self.to_out(o.transpose(1, 2).reshape(B, S, C))


class _Block(nn.Module):
    def __init__(self, c, ctx_dim, heads):
        super().__init__()
        self.norm1, self.norm2, self.norm3 = nn.LayerNorm(c), nn.LayerNorm(c), nn.LayerNorm(c)
        self.attn1 = _Attention(c, c, heads)
        self.attn2 = _Attention(c, ctx_dim, heads)
        self.ff_in = nn.Linear(c, 8 * c)
        self.ff_out = nn.Linear(4 * c, c)

    def forward(self, x, ctx):
        n = self.norm1(x)
        x = x + self.attn1(n, n)
        x = x + self.attn2(self.norm2(x), ctx)
        a, gate = self.ff_in(self.norm3(x)).chunk(2, dim=-1)
        return x + self.ff_out(a * F.gelu(gate, approximate="tanh"))


class _Transformer2D(nn.Module):
    def __init__(self, c, ctx_dim, heads, linear_proj):
        super().__init__()
        self.linear_proj = linear_proj
        self.norm = nn.GroupNorm(32, c, eps=1e-5)
        self.proj_in = nn.Linear(c, c) if linear_proj else nn.Conv2d(c, c, 1)
        self.block = _Block(c, ctx_dim, heads)
        self.proj_out = nn.Linear(c, c) if linear_proj else nn.Conv2d(c, c, 1)

    def forward(self, x, ctx):
        B, C, H, W = x.shape
        h = self.norm(x)
        if self.linear_proj:
            h = self.proj_in(h.flatten(2).transpose(1, 2))
        else:
            h = self.proj_in(h).flatten(2).transpose(1, 2)
        h = self.block(h, ctx)
        if self.linear_proj:
            h = self.proj_out(h).transpose(1, 2).reshape(B, C, H, W)
        else:
            h = self.proj_out(h.transpose(1, 2).reshape(B, C, H, W))
        return h + x


class _ResNet(nn.Module):
    def __init__(self, cin, cout, temb):
        super().__init__()
        self.norm1 = nn.GroupNorm(32, cin, eps=1e-5)
        self.conv1 = nn.Conv2d(cin, cout, 3, padding=1)
        self.time_emb_proj = nn.Linear(temb, cout)
        self.norm2 = nn.GroupNorm(32, cout, eps=1e-5)
        self.conv2 = nn.Conv2d(cout, cout, 3, padding=1)
        self.shortcut = nn.Conv2d(cin, cout, 1) if cin != cout else None

    def forward(self, x, temb):
        h = self.conv1(F.silu(self.norm1(x)))
        h = h + self.time_emb_proj(F.silu(temb))[:, :, None, None]
        h = self.conv2(F.silu(self.norm2(h)))
        return h + (x if self.shortcut is None else self.shortcut(x))


class _RefUNet(nn.Module):
    """PyTorch-style UNet2DConditionModel skeleton; `flax_names` maps every torch parameter to its Flax leaf."""

    def __init__(self, cfg: OU.UNetConfig):
        super().__init__()
        self.cfg = cfg
        boc = cfg.block_out_channels
        n = len(boc)
        temb = 4 * boc[0]
        self.names = {}

        def reg(mod, flax_prefix):
            self.names[flax_prefix] = mod
            return mod

        self.conv_in = reg(nn.Conv2d(cfg.in_channels, boc[0], 3, padding=1), "conv_in")
        self.time1 = reg(nn.Linear(boc[0], temb), "time_embedding/linear_1")
        self.time2 = reg(nn.Linear(temb, temb), "time_embedding/linear_2")
        self.down = nn.ModuleList()
        ch = boc[0]
        skip = [ch]
        for i, kind in enumerate(cfg.down_block_types):
            blk = nn.ModuleDict({"res": nn.ModuleList(), "att": nn.ModuleList()})
            for j in range(cfg.layers_per_block):
                blk["res"].append(reg(_ResNet(ch, boc[i], temb), f"down_blocks_{i}/resnets_{j}"))
                ch = boc[i]
                if kind == "CrossAttnDownBlock2D":
                    blk["att"].append(reg(_Transformer2D(ch, cfg.cross_attention_dim, cfg.heads(i),
                                                         cfg.use_linear_projection), f"down_blocks_{i}/attentions_{j}"))
                skip.append(ch)
            if i != n - 1:
                blk["down"] = reg(nn.Conv2d(ch, ch, 3, stride=2, padding=1), f"down_blocks_{i}/downsamplers_0/conv")
                skip.append(ch)
            self.down.append(blk)
        self.mid_res0 = reg(_ResNet(ch, ch, temb), "mid_block/resnets_0")
        self.mid_att = reg(_Transformer2D(ch, cfg.cross_attention_dim, cfg.heads(n - 1), cfg.use_linear_projection),
                           "mid_block/attentions_0")
        self.mid_res1 = reg(_ResNet(ch, ch, temb), "mid_block/resnets_1")
        self.up = nn.ModuleList()
        for i, kind in enumerate(cfg.up_block_types):
            cout = boc[n - 1 - i]
            blk = nn.ModuleDict({"res": nn.ModuleList(), "att": nn.ModuleList()})
            for j in range(cfg.layers_per_block + 1):
                blk["res"].append(reg(_ResNet(ch + skip.pop(), cout, temb), f"up_blocks_{i}/resnets_{j}"))
                ch = cout
                if kind == "CrossAttnUpBlock2D":
                    blk["att"].append(reg(_Transformer2D(ch, cfg.cross_attention_dim, cfg.heads(n - 1 - i),
                                                         cfg.use_linear_projection), f"up_blocks_{i}/attentions_{j}"))
            if i != n - 1:
                blk["up"] = reg(nn.Conv2d(ch, ch, 3, padding=1), f"up_blocks_{i}/upsamplers_0/conv")
            self.up.append(blk)
        self.norm_out = reg(nn.GroupNorm(32, ch, eps=1e-5), "conv_norm_out")
        self.conv_out = reg(nn.Conv2d(ch, cfg.out_channels, 3, padding=1), "conv_out")

    def forward(self, x, t, ctx):
        half = self.cfg.block_out_channels[0] // 2
        freqs = torch.exp(-math.log(10000.0) * torch.arange(half, dtype=x.dtype) / half)
        arg = t.to(x.dtype)[:, None] * freqs[None]
        temb = self.time2(F.silu(self.time1(torch.cat([arg.cos(), arg.sin()], dim=1))))
        x = self.conv_in(x)
        stack = [x]
        for blk in self.down:
            for j, res in enumerate(blk["res"]):
                x = res(x, temb)
                if len(blk["att"]):
                    x = blk["att"][j](x, ctx)
                stack.append(x)
            if "down" in blk:
                x = blk["down"](x)
                stack.append(x)
        x = self.mid_res1(self.mid_att(self.mid_res0(x, temb), ctx), temb)
        for blk in self.up:
            for j, res in enumerate(blk["res"]):
                x = res(torch.cat([x, stack.pop()], dim=1), temb)
                if len(blk["att"]):
                    x = blk["att"][j](x, ctx)
            if "up" in blk:
                x = blk["up"](F.interpolate(x, scale_factor=2.0, mode="nearest"))
        return self.conv_out(F.silu(self.norm_out(x)))


def _load_flax(model: _RefUNet, flax: dict) -> int:
    """Published Flax -> PyTorch conversion: conv kernel HWIO -> weight OIHW; Dense kernel (in,out) -> weight (out,in);
    norm scale -> weight. Returns the number of Flax leaves consumed."""
    used = set()

    def take(name):
        used.add(name)
        return flax[name].double()

    def conv(m, p):
        m.weight.data = take(p + "/kernel").permute(3, 2, 0, 1).contiguous()
        m.bias.data = take(p + "/bias")

    def lin(m, p, bias=True):
        k = take(p + "/kernel")
        m.weight.data = k.reshape(-1, k.shape[-1]).t().contiguous()
        if bias:
            m.bias.data = take(p + "/bias")

    def norm(m, p):
        m.weight.data = take(p + "/scale")
        m.bias.data = take(p + "/bias")

    for prefix, m in model.names.items():
        if isinstance(m, nn.Conv2d):
            conv(m, prefix)
        elif isinstance(m, nn.Linear):
            lin(m, prefix)
        elif isinstance(m, nn.GroupNorm):
            norm(m, prefix)
        elif isinstance(m, _ResNet):
            norm(m.norm1, prefix + "/norm1")
            conv(m.conv1, prefix + "/conv1")
            lin(m.time_emb_proj, prefix + "/time_emb_proj")
            norm(m.norm2, prefix + "/norm2")
            conv(m.conv2, prefix + "/conv2")
            if m.shortcut is not None:
                conv(m.shortcut, prefix + "/conv_shortcut")
        elif isinstance(m, _Transformer2D):
            norm(m.norm, prefix + "/norm")
            (lin if m.linear_proj else conv)(m.proj_in, prefix + "/proj_in")
            (lin if m.linear_proj else conv)(m.proj_out, prefix + "/proj_out")
            t = prefix + "/transformer_blocks_0"
            b = m.block
            for i, ln in enumerate((b.norm1, b.norm2, b.norm3), 1):
                norm(ln, f"{t}/norm{i}")
            for a, nm in ((b.attn1, "attn1"), (b.attn2, "attn2")):
                lin(a.to_q, f"{t}/{nm}/to_q", bias=False)
                lin(a.to_k, f"{t}/{nm}/to_k", bias=False)
                lin(a.to_v, f"{t}/{nm}/to_v", bias=False)
                lin(a.to_out, f"{t}/{nm}/to_out_0")
            lin(b.ff_in, t + "/ff/net_0/proj")
            lin(b.ff_out, t + "/ff/net_2")
        else:
            raise AssertionError(type(m))
    assert used == set(flax), sorted(set(flax) - used)[:5]
    return len(used)


_MINI_SD15 = dict(block_out_channels=(32, 64, 64), layers_per_block=2,
                  down_block_types=("CrossAttnDownBlock2D", "CrossAttnDownBlock2D", "DownBlock2D"),
                  up_block_types=("UpBlock2D", "CrossAttnUpBlock2D", "CrossAttnUpBlock2D"), attention_head_dim=4,
                  cross_attention_dim=24)
_MINI_SD21 = dict(block_out_channels=(64, 128), down_block_types=("CrossAttnDownBlock2D", "DownBlock2D"),
                  up_block_types=("UpBlock2D", "CrossAttnUpBlock2D"), layers_per_block=2, attention_head_dim=(1, 2),
                  cross_attention_dim=48, use_linear_projection=True)


@pytest.mark.parametrize("kw,H", [(None, 32), (_MINI_SD15, 40), (_MINI_SD21, 20)], ids=["tiny", "mini-sd15", "mini-sd21"])
def test_oracle_forward_equals_independent_torch_nn_restatement(kw, H):
    """fp64 forward AND parameter gradients of oracle/unet.py == the stock-torch.nn NCHW model loaded through the
    Flax -> PyTorch conversion rule. Sizes are chosen so that every attention has Sq >= Skv or kc | Skv (no key-chunk
    double counting): with the quirk out of the way the two must agree to fp64 rounding."""
    cfg = OU.TINY if kw is None else OU.UNetConfig(**kw)
    torch.manual_seed(0)
    params = OU.init_params(cfg, 3)
    B, L = 2, 77
    g = torch.Generator().manual_seed(5)
    x = torch.randn(B, 4, H, H, generator=g, dtype=torch.float64)
    t = torch.tensor([17, 803])
    ctx = torch.randn(B, L, cfg.cross_attention_dim, generator=g, dtype=torch.float64)
    n_levels = len(cfg.block_out_channels)
    for lvl in range(n_levels):  # no level may trigger the quirk
        sq = (H >> lvl) ** 2
        assert all(m == 1 for m in OU.key_multiplicity(sq, L)), (lvl, sq)

    leaves = {k: v.double().requires_grad_(True) for k, v in params.items()}
    want = OU.unet_forward(cfg, leaves, x, t, ctx, mode="fp64")
    w = torch.randn(want.shape, generator=g, dtype=torch.float64)
    (want * w).sum().backward()

    ref = _RefUNet(cfg).double()
    assert _load_flax(ref, params) == len(params)
    got = ref(x, t, ctx)
    (got * w).sum().backward()
    scale = want.abs().max().item()
    assert (got - want).abs().max().item() <= 1e-10 * scale
    # gradients of a few structurally different leaves, converted back to the Flax layout
    checks = [("conv_in/kernel", ref.conv_in.weight.grad.permute(2, 3, 1, 0)),
              ("time_embedding/linear_1/kernel", ref.time1.weight.grad.t()),
              ("conv_norm_out/scale", ref.norm_out.weight.grad)]
    att = ref.names["mid_block/attentions_0"]
    tb = "mid_block/attentions_0/transformer_blocks_0"
    checks += [(f"{tb}/attn2/to_k/kernel", att.block.attn2.to_k.weight.grad.t()),
               (f"{tb}/attn1/to_q/kernel", att.block.attn1.to_q.weight.grad.t()),
               (f"{tb}/ff/net_0/proj/kernel", att.block.ff_in.weight.grad.t()),
               (f"{tb}/norm3/scale", att.block.norm3.weight.grad)]
    up_last = ref.names[f"up_blocks_{n_levels - 1}/resnets_0"]
    checks += [(f"up_blocks_{n_levels - 1}/resnets_0/conv1/kernel", up_last.conv1.weight.grad.permute(2, 3, 1, 0)),
               (f"up_blocks_{n_levels - 1}/resnets_0/conv_shortcut/kernel", up_last.shortcut.weight.grad.permute(2, 3, 1, 0)),
               (f"up_blocks_{n_levels - 1}/resnets_0/time_emb_proj/kernel", up_last.time_emb_proj.weight.grad.t())]
    # (some leaves have an exactly-zero true gradient — time_emb_proj in front of a GroupNorm with one channel per
    #  group — where only rounding noise remains: tolerances are relative to the largest checked gradient)
    gscale = max(leaves[name].grad.abs().max().item() for name, _ in checks)
    for name, gref in checks:
        gor = leaves[name].grad
        gref = gref.reshape(gor.shape)
        assert (gor - gref).abs().max().item() <= 1e-9 * max(gor.abs().max().item(), 1e-3 * gscale), name


@pytest.mark.parametrize("sq,skv", [(64, 77), (64, 227), (144, 227), (144, 77), (256, 77), (16, 77)])
def test_literal_key_chunks_equal_multiplicity_bias(sq, skv):
    """The patched reference's clamped key chunks (oracle: literal) == softmax with an additive log(m_k) key bias."""
    g = torch.Generator().manual_seed(sq * 1000 + skv)
    BH, d = 3, 16
    q = torch.randn(BH, sq, d, generator=g, dtype=torch.float64)
    k = torch.randn(BH, skv, d, generator=g, dtype=torch.float64)
    v = torch.randn(BH, skv, d, generator=g, dtype=torch.float64)
    lit = OU._chunked_attention(OU._Ctx({}, "fp64"), q, k, v)
    m = torch.tensor(OU.key_multiplicity(sq, skv), dtype=torch.float64)
    bias = torch.log(m)[None, None, :].expand(BH, sq, skv)
    ref = F.scaled_dot_product_attention(q[None], k[None], v[None], attn_mask=bias[None])[0]
    assert (lit - ref).abs().max().item() < 1e-12
    # SURVEY.md A.3 table
    table = {(64, 77): (13, 63), (64, 227): (163, 191), (144, 227): (83, 143)}
    if (sq, skv) in table:
        lo, hi = table[(sq, skv)]
        assert [i for i, x in enumerate(m.tolist()) if x == 2] == list(range(lo, hi + 1))
    else:
        assert ((m == 1).all().item()) == (skv % min(sq, skv) == 0 or sq >= skv)


def test_parameter_census_and_product_tree_match():
    from stable_diffusion_training_b200 import unet as PU
    for ocfg, pcfg, leaves, total in ((OU.SD15, PU.SD15, 686, 859_520_964), (OU.SD21, PU.SD21, 686, 865_910_724)):
        so = OU.param_shapes(ocfg)
        sp = PU.param_shapes(pcfg)
        assert so == sp
        assert len(so) == leaves
        assert sum(int(np.prod(s)) for s in so.values()) == total
    assert OU.param_shapes(OU.TINY) == PU.param_shapes(PU.TINY)
    assert PU.key_multiplicity(64, 77) == OU.key_multiplicity(64, 77)
