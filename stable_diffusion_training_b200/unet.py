"""SD UNet forward + hand-scheduled backward on the sm_100a kernels (host side).

Mirrors diffusers==0.21.4 FlaxUNet2DConditionModel as the reference calls it
(/root/reference/training_utils.py:209-214 load, 678-684 apply; un-vendored upstream, wiring restated in SURVEY.md
appendix B) including the reference's attention patch (/root/reference/key_chunk_patch.patch:1-9). Parameter names and
layouts are Flax's (Dense kernel (in,out), Conv kernel HWIO), so `state.params` trees interchange with the reference.

There is no autograd and no eager fallback: every op below enqueues kernels of libsdb200.so and registers an explicit
backward closure on a tape; `backward()` replays the tape in reverse. torch is used for device buffers only.
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from . import ops

BF16 = torch.bfloat16


@dataclass(frozen=True)
class UNetConfig:
    in_channels: int = 4
    out_channels: int = 4
    block_out_channels: Tuple[int, ...] = (320, 640, 1280, 1280)
    down_block_types: Tuple[str, ...] = ("CrossAttnDownBlock2D",) * 3 + ("DownBlock2D",)
    up_block_types: Tuple[str, ...] = ("UpBlock2D",) + ("CrossAttnUpBlock2D",) * 3
    layers_per_block: int = 2
    attention_head_dim: object = 8   # Flax UNet quirk: used as the NUMBER of heads (int or per-block tuple)
    cross_attention_dim: int = 768
    use_linear_projection: bool = False
    flip_sin_to_cos: bool = True
    freq_shift: int = 0

    def heads(self, level: int) -> int:
        a = self.attention_head_dim
        return int(a) if isinstance(a, int) else int(a[level])

    @property
    def temb_dim(self) -> int:
        return 4 * self.block_out_channels[0]


SD15 = UNetConfig()
SD21 = UNetConfig(attention_head_dim=(5, 10, 20, 20), cross_attention_dim=1024, use_linear_projection=True)
TINY = UNetConfig(block_out_channels=(32, 64), down_block_types=("CrossAttnDownBlock2D", "DownBlock2D"),
                  up_block_types=("UpBlock2D", "CrossAttnUpBlock2D"), layers_per_block=1, attention_head_dim=2,
                  cross_attention_dim=32)


# ---------------------------------------------------------------------------------------------------------------
# Layer program: one list drives both the parameter census and the executor.
def layer_program(cfg: UNetConfig) -> List[tuple]:
    prog: List[tuple] = [("conv_in", "conv_in", cfg.in_channels, cfg.block_out_channels[0]), ("push",)]
    boc = cfg.block_out_channels
    n = len(boc)
    ch = boc[0]
    skip_ch = [ch]
    for i, kind in enumerate(cfg.down_block_types):
        for j in range(cfg.layers_per_block):
            prog.append(("resnet", f"down_blocks_{i}/resnets_{j}", ch, boc[i]))
            ch = boc[i]
            if kind == "CrossAttnDownBlock2D":
                prog.append(("xformer", f"down_blocks_{i}/attentions_{j}", ch, cfg.heads(i)))
            prog.append(("push",))
            skip_ch.append(ch)
        if i != n - 1:
            prog.append(("down", f"down_blocks_{i}/downsamplers_0/conv", ch))
            prog.append(("push",))
            skip_ch.append(ch)
    prog.append(("resnet", "mid_block/resnets_0", ch, ch))
    prog.append(("xformer", "mid_block/attentions_0", ch, cfg.heads(n - 1)))
    prog.append(("resnet", "mid_block/resnets_1", ch, ch))
    for i, kind in enumerate(cfg.up_block_types):
        cout = boc[n - 1 - i]
        for j in range(cfg.layers_per_block + 1):
            prog.append(("popcat",))
            prog.append(("resnet", f"up_blocks_{i}/resnets_{j}", ch + skip_ch.pop(), cout))
            ch = cout
            if kind == "CrossAttnUpBlock2D":
                prog.append(("xformer", f"up_blocks_{i}/attentions_{j}", ch, cfg.heads(n - 1 - i)))
        if i != n - 1:
            prog.append(("up", f"up_blocks_{i}/upsamplers_0/conv", ch))
    assert not skip_ch
    prog.append(("out", "conv_norm_out", "conv_out", ch, cfg.out_channels))
    return prog


def param_shapes(cfg: UNetConfig) -> Dict[str, tuple]:
    """Flax parameter tree as {"a/b/c": shape}, sorted by path (jax pytree leaf order)."""
    s: Dict[str, tuple] = {}
    td = cfg.temb_dim

    def conv(name, cin, cout, k=3):
        s[name + "/kernel"] = (k, k, cin, cout)
        s[name + "/bias"] = (cout,)

    def dense(name, cin, cout, bias=True):
        s[name + "/kernel"] = (cin, cout)
        if bias:
            s[name + "/bias"] = (cout,)

    def norm(name, c):
        s[name + "/scale"] = (c,)
        s[name + "/bias"] = (c,)

    dense("time_embedding/linear_1", cfg.block_out_channels[0], td)
    dense("time_embedding/linear_2", td, td)
    for op in layer_program(cfg):
        kind = op[0]
        if kind == "conv_in":
            conv(op[1], op[2], op[3])
        elif kind in ("down", "up"):
            conv(op[1], op[2], op[2])
        elif kind == "resnet":
            _, name, cin, cout = op
            norm(name + "/norm1", cin)
            conv(name + "/conv1", cin, cout)
            dense(name + "/time_emb_proj", td, cout)
            norm(name + "/norm2", cout)
            conv(name + "/conv2", cout, cout)
            if cin != cout:
                conv(name + "/conv_shortcut", cin, cout, k=1)
        elif kind == "xformer":
            _, name, c, _h = op
            norm(name + "/norm", c)
            for pj in ("proj_in", "proj_out"):
                if cfg.use_linear_projection:
                    dense(f"{name}/{pj}", c, c)
                else:
                    conv(f"{name}/{pj}", c, c, k=1)
            t = name + "/transformer_blocks_0"
            for nn_ in ("norm1", "norm2", "norm3"):
                norm(f"{t}/{nn_}", c)
            for a, kd in (("attn1", c), ("attn2", cfg.cross_attention_dim)):
                dense(f"{t}/{a}/to_q", c, c, bias=False)
                dense(f"{t}/{a}/to_k", kd, c, bias=False)
                dense(f"{t}/{a}/to_v", kd, c, bias=False)
                dense(f"{t}/{a}/to_out_0", c, c)
            dense(f"{t}/ff/net_0/proj", c, 8 * c)
            dense(f"{t}/ff/net_2", 4 * c, c)
        elif kind == "out":
            norm(op[1], op[3])
            conv(op[2], op[3], op[4])
    return dict(sorted(s.items()))


def init_params_numpy(cfg: UNetConfig, seed: int = 0) -> Dict[str, np.ndarray]:
    """Synthetic random-init weights: NumPy default_rng(seed), sorted-path order; kernels N(0, 1/fan_in),
    biases N(0, 0.02^2), norm scale 1 + N(0, 0.1^2), norm bias N(0, 0.1^2)."""
    rng = np.random.default_rng(seed)
    out = {}
    for name, shape in param_shapes(cfg).items():
        parts = name.split("/")
        leaf, parent = parts[-1], parts[-2]
        if leaf == "kernel":
            a = rng.standard_normal(shape, dtype=np.float32) * np.float32(1.0 / math.sqrt(int(np.prod(shape[:-1]))))
        elif leaf == "scale":
            a = np.float32(1) + rng.standard_normal(shape, dtype=np.float32) * np.float32(0.1)
        elif parent.startswith("norm") or parent == "conv_norm_out":
            a = rng.standard_normal(shape, dtype=np.float32) * np.float32(0.1)
        else:
            a = rng.standard_normal(shape, dtype=np.float32) * np.float32(0.02)
        out[name] = a.astype(np.float32)
    return out


def key_multiplicity(sq: int, skv: int) -> List[int]:
    """How often each key is counted by the patched reference (key_chunk_patch.patch:5-6: key_chunk_size =
    flatten_latent_dim): chunk size kc = min(Sq, Skv), chunk starts arange(0, Skv, kc) clamped to Skv - kc by
    lax.dynamic_slice, so the last chunk re-reads keys when kc does not divide Skv."""
    kc = min(sq, skv)
    m = [0] * skv
    for s in range(0, skv, kc):
        s = min(s, skv - kc)
        for k in range(s, s + kc):
            m[k] += 1
    return m


# ---------------------------------------------------------------------------------------------------------------
class T:
    """A bf16 activation [rows, C] (+ its gradient slot)."""
    __slots__ = ("v", "g", "rg", "geom")

    def __init__(self, v: torch.Tensor, geom=None, rg: bool = True):
        self.v = v
        self.g: Optional[torch.Tensor] = None
        self.rg = rg
        self.geom = geom  # (B, H, W) for image-shaped activations


class _NoTape(list):
    """Tape of a forward-only pass: the backward closures (and the activations they hold) are dropped at once."""

    def append(self, item) -> None:
        pass


class Graph:
    """One forward/backward pass over device buffers. P: fp32 masters, W: bf16 shadows, G: fp32 gradients
    (dicts path -> tensor views into the arenas). forward_only: nothing is kept for a backward pass (the sampler),
    so every activation is freed as soon as its consumers have been enqueued."""

    def __init__(self, P, W, G, reference_attention_quirk: bool = True, forward_only: bool = False):
        self.P, self.W, self.G = P, W, G
        self.tape = _NoTape() if forward_only else []
        self.quirk = reference_attention_quirk
        self._kb_cache = {}
        # Parameter gradients (wgrad GEMMs, bias column sums) feed nothing but the optimizer, so backward enqueues them
        # on a second stream: their CTAs fill the SMs the main chain leaves idle (partial last waves of the persistent
        # GEMMs, launch gaps). SDB_SIDE_STREAM=0 keeps everything on one stream.
        self.side: Optional[torch.cuda.Stream] = None
        self._side_enabled = os.environ.get("SDB_SIDE_STREAM", "1") != "0"
        self._forked = False
        self._keep: list = []

    # -------- second stream for parameter gradients --------
    def _fork(self, fn, *keep) -> None:
        """Run `fn` (launches that only READ activations / output gradients and write parameter gradients) on the
        side stream, after everything the main stream has enqueued so far. `keep` are tensors the main chain drops
        before the side stream is done with them: they stay referenced until the join so the allocator cannot hand
        their memory to a later main-stream tensor."""
        if not self._side_enabled or not torch.cuda.is_available():
            fn()
            return
        main = torch.cuda.current_stream()
        if self.side is None or self.side.device != main.device:
            self.side = torch.cuda.Stream(device=main.device)
        self.side.wait_stream(main)
        with torch.cuda.stream(self.side), ops.side_workspace():
            fn()
        self._keep.extend(keep)
        self._forked = True

    def join(self) -> None:
        """Main stream waits for the side stream (before gradients are consumed: bucket all-reduce, optimizer, any
        in-place accumulation into a tensor the side stream may still read)."""
        if self._forked:
            torch.cuda.current_stream().wait_stream(self.side)
            self._forked = False

    # -------- gradient plumbing --------
    @staticmethod
    def _acc(t: T, g: torch.Tensor) -> None:
        if not t.rg:
            return
        if t.g is None:
            t.g = g
        else:
            ops.add_(t.g, g)

    def _acc_safe(self, t: T, g: torch.Tensor) -> None:
        if t.rg and t.g is not None:
            self.join()  # in-place add into a tensor that may alias an output gradient the side stream still reads
        self._acc(t, g)

    def backward(self, on_grads=None) -> None:
        """Replay the tape in reverse. `on_grads(names)` is called right after the closure that finishes the gradients
        of those parameter leaves has been enqueued (every leaf is used once per forward, so 'finished' is exact):
        the data-parallel reducer starts a bucket's all-reduce from it while the rest of backward still runs."""
        for fn, names in reversed(self.tape):
            fn()
            if on_grads is not None and names:
                on_grads(names)
        self.join()
        self._keep = []
        self.tape = []

    def _kbias(self, sq: int, skv: int, device):
        if not self.quirk:
            return None
        key = (sq, skv)
        if key not in self._kb_cache:
            m = key_multiplicity(sq, skv)
            self._kb_cache[key] = None if all(x == 1 for x in m) else torch.tensor(
                [math.log2(x) for x in m], dtype=torch.float32, device=device)
        return self._kb_cache[key]

    # -------- primitive layers --------
    def linear(self, x: T, name: str, bias: bool = True, res: Optional[T] = None, geom=None, side: bool = False) -> T:
        """side=True: the WHOLE backward of this layer (input gradient included) runs on the parameter-gradient stream.
        Used for the 22 time_emb_proj projections: their output gradients are per-image column sums that feed nothing
        on the main chain until the time-embedding MLP at the very end of backward, so ~5 tiny dependent launches per
        ResNet leave the critical path (the consumer joins the streams: Graph.silu(join_before_bwd=True))."""
        w = self.W[name + "/kernel"]
        w2 = w.view(-1, w.shape[-1])
        y = T(ops.linear(x.v, w2, bias=self.P[name + "/bias"] if bias else None,
                         residual=None if res is None else res.v), geom if geom is not None else x.geom)

        def bwd():
            dy = y.g
            gw = self.G[name + "/kernel"]

            def param_grads():
                if bias:
                    ops.colsum(dy, self.G[name + "/bias"])
                ops.linear_wgrad(x.v, dy, out=gw.view(-1, gw.shape[-1]))
                if side and x.rg:
                    old = x.g
                    x.g = ops.linear_dgrad(dy, w2, residual=old)
                    if old is not None:
                        self._keep.append(old)
            self._fork(param_grads, dy)
            if x.rg and not side:
                x.g = ops.linear_dgrad(dy, w2, residual=x.g)
            if res is not None:
                self._acc_safe(res, dy)
            y.g = None
        self.tape.append((bwd, [name + "/kernel"] + ([name + "/bias"] if bias else [])))
        return y

    def conv3x3(self, x: T, name: str, temb: Optional[T] = None, res: Optional[T] = None, stride: int = 1) -> T:
        B, H, W = x.geom
        w = self.W[name + "/kernel"]
        cin, cout = w.shape[2], w.shape[3]
        x4 = x.v.view(B, H, W, cin)
        small = cin < 16 or cout < 16  # conv_in / conv_out: CUDA-core kernels
        impl = "simt" if small else "tc"
        c4 = stride == 1 and temb is None and res is None and ((cin == 4 and cout % 8 == 0 and 8 <= cout <= 904) or
                                                                (cout == 4 and cin % 8 == 0 and 8 <= cin <= 904))
        if c4:
            fn = ops.conv3x3_c4in if cin == 4 else ops.conv3x3_c4out
            y4 = fn(x4, w, bias=self.P[name + "/bias"])
            geom = (B, H, W)
        elif stride == 1:
            y4 = ops.conv3x3(x4, w, bias=self.P[name + "/bias"], temb=None if temb is None else temb.v,
                             residual=None if res is None else res.v.view(B, H, W, cout), impl=impl)
            geom = (B, H, W)
        else:
            y4 = ops.conv3x3_s2(x4, w, bias=self.P[name + "/bias"], impl=impl)
            geom = (B, H // 2, W // 2)
        y = T(y4.view(-1, cout), geom)

        def bwd():
            dy = y.g
            dy4 = dy.view(geom[0], geom[1], geom[2], cout)
            if temb is not None:
                # the per-image column sums are the gradient of the broadcast time-embedding add; `temb` (a
                # time_emb_proj output) has this one consumer and its own backward runs on the side stream too
                dt = torch.empty((B, cout), device=dy.device, dtype=BF16)
                self._fork(lambda: ops.colsum(dy, self.G[name + "/bias"], nseg=B, out_seg=dt), dy, dt)
                assert temb.g is None
                temb.g = dt

            def param_grads():
                if temb is None:
                    ops.colsum(dy, self.G[name + "/bias"])
                if c4:
                    ops.conv3x3_c4_wgrad(x4, dy4, out=self.G[name + "/kernel"])
                elif stride == 1:
                    ops.conv3x3_wgrad(x4, dy4, out=self.G[name + "/kernel"], impl=impl)
                else:
                    ops.conv3x3_s2_wgrad(x4, dy4, out=self.G[name + "/kernel"], impl=impl)
            self._fork(param_grads, dy)
            if c4:
                if x.rg:
                    self._acc_safe(x, ops.conv3x3_c4out_dgrad(dy4, w).view(-1, cin))
            elif stride == 1:
                if x.rg:
                    if x.g is None or small:
                        self._acc_safe(x, ops.conv3x3_dgrad(dy4, w, impl=impl).view(-1, cin))
                    else:  # accumulate into the existing gradient inside the GEMM epilogue
                        x.g = _dgrad_acc(dy4, w, x.g.view(B, H, W, cin)).view(-1, cin)
            else:
                if x.rg:
                    self._acc_safe(x, ops.conv3x3_s2_dgrad(dy4, w, impl=impl).view(-1, cin))
            if res is not None:
                self._acc_safe(res, dy)
            y.g = None
        self.tape.append((bwd, [name + "/kernel", name + "/bias"]))
        return y

    def groupnorm(self, x: T, name: str, silu: bool) -> T:
        B = x.geom[0]
        C = x.v.shape[1]
        gamma, beta = self.P[name + "/scale"], self.P[name + "/bias"]
        yv, stats = ops.groupnorm_fwd(x.v.view(B, -1, C), gamma, beta, silu=silu)
        y = T(yv.view(-1, C), x.geom)

        def bwd():
            x.g = ops.groupnorm_bwd(x.v.view(B, -1, C), y.g.view(B, -1, C), stats, gamma, beta,
                                    self.G[name + "/scale"], self.G[name + "/bias"], silu=silu,
                                    dx_add=None if x.g is None else x.g.view(B, -1, C)).view(-1, C)
            y.g = None
        self.tape.append((bwd, [name + "/scale", name + "/bias"]))
        return y

    def layernorm(self, x: T, name: str) -> T:
        gamma, beta = self.P[name + "/scale"], self.P[name + "/bias"]
        yv, stats = ops.layernorm_fwd(x.v, gamma, beta)
        y = T(yv, x.geom)

        def bwd():
            x.g = ops.layernorm_bwd(x.v, y.g, stats, gamma, self.G[name + "/scale"], self.G[name + "/bias"],
                                    dx_add=x.g)
            y.g = None
        self.tape.append((bwd, [name + "/scale", name + "/bias"]))
        return y

    def geglu(self, u: T) -> T:
        y = T(ops.geglu_fwd(u.v), u.geom)

        def bwd():
            self._acc_safe(u, ops.geglu_bwd(u.v, y.g))
            y.g = None
        self.tape.append((bwd, None))
        return y

    def silu(self, x: T, join_before_bwd: bool = False) -> T:
        y = T(ops.silu_fwd(x.v), x.geom)

        def bwd():
            if join_before_bwd:
                self.join()  # y.g was accumulated on the side stream (Graph.linear(side=True))
            self._acc_safe(x, ops.silu_bwd(x.v, y.g))
            y.g = None
        self.tape.append((bwd, None))
        return y

    def attention_core(self, q: T, k: T, v: T, B: int, heads: int) -> T:
        C = q.v.shape[1]
        Sq, Skv = q.v.shape[0] // B, k.v.shape[0] // B
        kb = self._kbias(Sq, Skv, q.v.device)
        q3, k3, v3 = q.v.view(B, Sq, C), k.v.view(B, Skv, C), v.v.view(B, Skv, C)
        o3, lse2 = ops.attention_fwd(q3, k3, v3, heads, kbias2=kb)
        o = T(o3.view(-1, C), q.geom)

        def bwd():
            dq, dk, dv = ops.attention_bwd(q3, k3, v3, o3, o.g.view(B, Sq, C), lse2, heads, kbias2=kb)
            self._acc_safe(q, dq.view(-1, C))
            self._acc_safe(k, dk.view(-1, C))
            self._acc_safe(v, dv.view(-1, C))
            o.g = None
        self.tape.append((bwd, None))
        return o

    def attention(self, x: T, ctx: T, name: str, B: int, heads: int, res: T) -> T:
        q = self.linear(x, name + "/to_q", bias=False)
        k = self.linear(ctx, name + "/to_k", bias=False)
        v = self.linear(ctx, name + "/to_v", bias=False)
        o = self.attention_core(q, k, v, B, heads)
        return self.linear(o, name + "/to_out_0", bias=True, res=res)

    def concat(self, a: T, b: T) -> T:
        ca, cb = a.v.shape[1], b.v.shape[1]
        out = torch.empty((a.v.shape[0], ca + cb), device=a.v.device, dtype=BF16)
        ops.copy2d(a.v, out[:, :ca])
        ops.copy2d(b.v, out[:, ca:])
        y = T(out, a.geom)

        def bwd():
            for t, sl in ((a, y.g[:, :ca]), (b, y.g[:, ca:])):
                if t.g is None:
                    t.g = torch.empty_like(t.v)
                    ops.copy2d(sl, t.g)
                else:
                    self.join()
                    ops.copy2d(sl, t.g, accumulate=True)
            y.g = None
        self.tape.append((bwd, None))
        return y

    def upsample(self, x: T) -> T:
        B, H, W = x.geom
        C = x.v.shape[1]
        y = T(ops.upsample2x_fwd(x.v.view(B, H, W, C)).view(-1, C), (B, 2 * H, 2 * W))

        def bwd():
            self._acc_safe(x, ops.upsample2x_bwd(y.g.view(B, 2 * H, 2 * W, C)).view(-1, C))
            y.g = None
        self.tape.append((bwd, None))
        return y

    # -------- composite blocks (diffusers 0.21.4 FlaxResnetBlock2D / FlaxTransformer2DModel) --------
    def resnet(self, x: T, temb_act: T, name: str) -> T:
        a = self.groupnorm(x, name + "/norm1", silu=True)
        t = self.linear(temb_act, name + "/time_emb_proj", side=True)
        h = self.conv3x3(a, name + "/conv1", temb=t)
        b = self.groupnorm(h, name + "/norm2", silu=True)
        sc = self.linear(x, name + "/conv_shortcut") if (name + "/conv_shortcut/kernel") in self.W else x
        return self.conv3x3(b, name + "/conv2", res=sc)

    def xformer(self, x: T, ctx: T, name: str, heads: int) -> T:
        B = x.geom[0]
        h = self.groupnorm(x, name + "/norm", silu=False)
        h = self.linear(h, name + "/proj_in")
        t = name + "/transformer_blocks_0"
        n1 = self.layernorm(h, t + "/norm1")
        h = self.attention(n1, n1, t + "/attn1", B, heads, res=h)
        h = self.attention(self.layernorm(h, t + "/norm2"), ctx, t + "/attn2", B, heads, res=h)
        u = self.linear(self.layernorm(h, t + "/norm3"), t + "/ff/net_0/proj")
        h = self.linear(self.geglu(u), t + "/ff/net_2", res=h)
        return self.linear(h, name + "/proj_out", res=x)


def _dgrad_acc(dy4, w, add4):
    """conv3x3 dgrad with the existing gradient added in the epilogue."""
    B, H, Wd, Cout = dy4.shape
    Cin = w.shape[2]
    out = torch.empty((B, H, Wd, Cin), device=dy4.device, dtype=BF16)
    ops.gemm_fwd([ops._view_nhwc(dy4)], ops._conv_taps(Cin, True), w.view(9 * Cin, Cout), False, out, (Wd, H, B), Cout,
                 Cin, residual=add4)
    return out


class UNet:
    """forward(sample NHWC bf16, timesteps, context) -> prediction [B*H*W, out_channels] bf16; backward(dpred)."""

    def __init__(self, cfg: UNetConfig, P, W, G, reference_attention_quirk: bool = True):
        self.cfg = cfg
        self.P, self.W, self.G = P, W, G
        self.program = layer_program(cfg)
        self.quirk = reference_attention_quirk
        self.graph: Optional[Graph] = None
        self._kb_cache = {}

    def forward(self, sample_nhwc: torch.Tensor, timesteps: torch.Tensor, context: torch.Tensor,
                forward_only: bool = False) -> T:
        """sample_nhwc [B,H,W,Cin] bf16; timesteps int32 [B]; context [B,L,D] bf16."""
        cfg = self.cfg
        g = Graph(self.P, self.W, self.G, self.quirk, forward_only)
        g._kb_cache = self._kb_cache
        self.graph = g
        B, H, W, Cin = sample_nhwc.shape
        x = T(sample_nhwc.view(-1, Cin), (B, H, W), rg=False)
        if isinstance(context, tuple):   # (T [B * L, D] produced by a trained text encoder, B): its gradient is wanted
            ctx = context[0]
            assert ctx.v.dim() == 2 and ctx.v.shape[0] % context[1] == 0
        else:
            ctx = T(context.view(-1, context.shape[-1]), None, rg=False)
        emb = T(ops.timestep_embedding(timesteps, cfg.block_out_channels[0], cfg.flip_sin_to_cos,
                                       float(cfg.freq_shift)), None, rg=False)
        temb = g.linear(g.silu(g.linear(emb, "time_embedding/linear_1")), "time_embedding/linear_2")
        temb_act = g.silu(temb, join_before_bwd=True)
        skips: List[T] = []
        for op in self.program:
            kind = op[0]
            if kind == "conv_in":
                x = g.conv3x3(x, op[1])
            elif kind == "push":
                skips.append(x)
            elif kind == "popcat":
                x = g.concat(x, skips.pop())
            elif kind == "resnet":
                x = g.resnet(x, temb_act, op[1])
            elif kind == "xformer":
                x = g.xformer(x, ctx, op[1], op[3])
            elif kind == "down":
                x = g.conv3x3(x, op[1], stride=2)
            elif kind == "up":
                x = g.conv3x3(g.upsample(x), op[1])
            elif kind == "out":
                x = g.conv3x3(g.groupnorm(x, op[1], silu=True), op[2])
        return x

    def backward(self, pred: T, dpred: torch.Tensor, on_grads=None) -> None:
        pred.g = dpred
        self.graph.backward(on_grads)
        self.graph = None
