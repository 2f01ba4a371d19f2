"""Frozen VAE encode and decode (8(f)-4, validation sampler) on the sm_100a kernels — host side,
forward only.

Mirrors diffusers==0.21.4 FlaxAutoencoderKL.encode as the reference calls it at the top of its step
(/root/reference/training_utils.py:574-586: `frozen_vae_state.call.apply({"params": ...}, sample=batch["pixel_values"],
deterministic=True, method=encode)`, `latent_dist.sample(sample_rng)`, transpose to NCHW, `* 0.18215`; loaded at 218-222
with dtype bf16). The VAE is frozen (`FrozenModel`), so there is no backward and no optimizer state: the bf16 operand
copies of the encoder's parameters are made once.

Convolutions / GroupNorm(+SiLU) / Dense are the UNet's kernels (implicit-GEMM tcgen05 conv, fused GroupNorm, GEMM);
the stride-2 downsample uses the parity-view trick with FlaxDownsample2D's ((0,1),(0,1)) padding; the mid block's
single-head attention (head dim 512 > the flash kernels' 192) runs per image as GEMM -> row softmax -> GEMM; conv_in
(3 channels) runs on pixels padded to 8 channels. Parameter names and layouts are Flax's (`encoder/conv_in/kernel`
HWIO, ...): trees interchange with the reference.

Decode mirrors FlaxAutoencoderKL.decode as the reference's sampler calls it
(/root/reference/models/pipeline_flax_stable_diffusion.py:246-251: `1 / scaling_factor * latents`, `vae.apply(...,
method=vae.decode).sample`, `(image / 2 + 0.5).clip(0, 1).transpose(0, 2, 3, 1)`): post_quant_conv, conv_in (4 ->
512, the UNet's 4-channel CUDA-core kernel), mid block, four up blocks (nearest 2x + conv3x3), conv_out (128 -> 3,
weights padded to 4 output channels) and the image post-processing fused into the read-out kernel.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from . import _lib, ops
from . import jax_random as JR
from .lion_quant import flatten_dict, unflatten_dict

BF16 = torch.bfloat16


@dataclass(frozen=True)
class VAEConfig:
    in_channels: int = 3
    latent_channels: int = 4
    block_out_channels: Tuple[int, ...] = (128, 256, 512, 512)
    layers_per_block: int = 2
    norm_num_groups: int = 32
    scaling_factor: float = 0.18215   # training_utils.py:586


SD_VAE = VAEConfig()
TINY_VAE = VAEConfig(block_out_channels=(32, 64), layers_per_block=1)


def param_shapes(cfg: VAEConfig) -> Dict[str, tuple]:
    """Encoder + quant_conv leaves of the FlaxAutoencoderKL tree (the decoder is never run in training)."""
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


def init_params_numpy(cfg: VAEConfig, seed: int = 0) -> Dict[str, np.ndarray]:
    """Synthetic random-init weights (same generator and order as oracle/vae.py:init_params)."""
    return _init_numpy(param_shapes(cfg), seed)


def _init_numpy(shapes: Dict[str, tuple], seed: int) -> Dict[str, np.ndarray]:
    rng = np.random.default_rng(seed)
    out = {}
    for name, shape in shapes.items():
        leaf = name.rsplit("/", 1)[1]
        parent = name.rsplit("/", 2)[-2]
        if leaf == "kernel":
            a = rng.standard_normal(shape, dtype=np.float32) * np.float32(1.0 / math.sqrt(int(np.prod(shape[:-1]))))
        elif leaf == "scale":
            a = np.float32(1) + rng.standard_normal(shape, dtype=np.float32) * np.float32(0.1)
        elif "norm" in parent:
            a = rng.standard_normal(shape, dtype=np.float32) * np.float32(0.1)
        else:
            a = rng.standard_normal(shape, dtype=np.float32) * np.float32(0.02)
        out[name] = a.astype(np.float32)
    return out


def _down_taps(cin: int):
    # FlaxDownsample2D: pad ((0,1),(0,1)) then a VALID 3x3 stride-2 conv: output (ho, wo) reads input (2ho + kh, 2wo + kw);
    # kh = 0 -> parity 0, kh = 1 -> parity 1, kh = 2 -> parity 0 shifted by +1 (past the edge: TMA zero fill = the padding)
    par = {0: (0, 0), 1: (1, 0), 2: (0, 1)}
    return [(par[kh][0] * 2 + par[kw][0], par[kh][1], par[kw][1], (kh * 3 + kw) * cin) for kh in range(3) for kw in range(3)]


def decoder_param_shapes(cfg: VAEConfig) -> Dict[str, tuple]:
    """Decoder + post_quant_conv leaves of the FlaxAutoencoderKL tree (validation sampling only)."""
    s: Dict[str, tuple] = {}

    def conv(n, cin, cout, k=3):
        s[n + "/kernel"] = (k, k, cin, cout)
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
        s[f"{a}/{d}/kernel"] = (ch, ch)
        s[f"{a}/{d}/bias"] = (ch,)
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


def init_decoder_params_numpy(cfg: VAEConfig, seed: int = 0) -> Dict[str, np.ndarray]:
    """Synthetic decoder weights (same generator and order as oracle/vae.py:init_decoder_params)."""
    return _init_numpy(decoder_param_shapes(cfg), seed + 7919)


class _VAEExec:
    """Loads the leaves named by `shapes` (fp32 copies P, bf16 GEMM operands W) and holds the layers both halves use."""

    def __init__(self, cfg: VAEConfig, params: dict, device, shapes: Dict[str, tuple], what: str):
        self.cfg = cfg
        self.device = torch.device(device)
        flat = flatten_dict(params)
        missing = sorted(set(shapes) - set(flat))
        if missing:
            raise ValueError(f"VAE params lack {what} leaves (e.g. {missing[:3]})")
        self.P: Dict[str, torch.Tensor] = {}   # fp32 (biases, norm parameters)
        self.W: Dict[str, torch.Tensor] = {}   # bf16 GEMM operands
        for n, shp in shapes.items():
            v = flat[n]
            t = torch.as_tensor(np.asarray(v) if not isinstance(v, torch.Tensor) else v).to(torch.float32)
            if tuple(t.shape) != tuple(shp):
                raise ValueError(f"VAE leaf {n}: shape {tuple(t.shape)} != {shp}")
            t = t.to(self.device).contiguous()
            self.P[n] = t
            if n.endswith("/kernel"):
                self.W[n] = ops.cast_bf16(t)

    # ---- layers (forward only) ----
    def _gn(self, x, B, name, silu):
        C = x.shape[-1]
        y, _ = ops.groupnorm_fwd(x.view(B, -1, C), self.P[name + "/scale"], self.P[name + "/bias"], silu=silu, eps=1e-6)
        return y

    def _conv(self, x4, name, residual=None):
        w = self.W[name + "/kernel"]
        impl = "simt" if (w.shape[2] < 16 or w.shape[3] < 16) else "tc"
        return ops.conv3x3(x4, w, bias=self.P[name + "/bias"], residual=residual, impl=impl)

    def _linear(self, x2, name, residual=None, impl=None):
        w = self.W[name + "/kernel"]
        w2 = w.view(-1, w.shape[-1])
        if impl is None:
            impl = "simt" if (w2.shape[0] < 16 or w2.shape[1] < 16) else "tc"
        return ops.linear(x2, w2, bias=self.P[name + "/bias"], residual=residual, impl=impl)

    def _resnet(self, x4, name):
        B, H, W, C = x4.shape
        h = self._conv(self._gn(x4, B, name + "/norm1", True).view(B, H, W, C), name + "/conv1")
        Co = h.shape[3]
        sc = x4 if (name + "/conv_shortcut/kernel") not in self.W else \
            self._linear(x4.view(-1, C), name + "/conv_shortcut").view(B, H, W, Co)
        return self._conv(self._gn(h, B, name + "/norm2", True).view(B, H, W, Co), name + "/conv2", residual=sc)

    def _downsample(self, x4, name):
        B, H, W, C = x4.shape
        w = self.W[name + "/kernel"]
        out = torch.empty((B, H // 2, W // 2, C), device=x4.device, dtype=BF16)
        ops.gemm_fwd(ops._parity_views(x4), _down_taps(C), w.view(9 * C, C), True, out, (W // 2, H // 2, B), C, C,
                     bias=self.P[name + "/bias"])
        return out

    def _attention(self, x4, name):
        B, H, W, C = x4.shape
        S = H * W
        h = self._gn(x4, B, name + "/group_norm", False).view(B * S, C)
        q, k, v = (self._linear(h, f"{name}/{d}") for d in ("query", "key", "value"))
        o = torch.empty((B * S, C), device=x4.device, dtype=BF16)
        scores = torch.empty((S, S), device=x4.device, dtype=BF16)
        for b in range(B):   # one head of width C: q k^T -> softmax -> P v, per image
            qb, kb, vb = (t[b * S:(b + 1) * S] for t in (q, k, v))
            ops.linear_dgrad(qb, kb, out=scores)                      # [S, C] x [S, C]^T
            _lib.check(_lib.lib().sdb_softmax_rows_bf16(scores.data_ptr(), scores.data_ptr(), S, S, scores.stride(0),
                                                        1.0 / math.sqrt(C), _lib.stream_ptr(x4.device)),
                       "sdb_softmax_rows_bf16")
            ops.linear(scores, vb, out=o[b * S:(b + 1) * S])
        return self._linear(o, name + "/proj_attn", residual=x4.view(B * S, C)).view(B, H, W, C)


class VAEEncoder(_VAEExec):
    """encode(pixel_values NCHW fp32 [B, 3, H, W]) -> moments NHWC bf16 [B, H/8 * W/8, 2 * latent]."""

    def __init__(self, cfg: VAEConfig, params: dict, device):
        super().__init__(cfg, params, device, param_shapes(cfg), "encoder")
        # conv_in on the tensor cores: pixels padded to 8 channels (TMA needs 16-byte pixel rows), the padded input
        # channels have zero weights; a K block of 64 is mostly TMA zero fill, which is still 7x faster than the CUDA-core
        # 4-channel kernel at 512x512 (r4m: 2.27 ms for B = 8)
        k = self.P["encoder/conv_in/kernel"]
        if cfg.in_channels > 8:
            raise NotImplementedError("conv_in with more than 8 input channels")
        kp = torch.zeros((3, 3, 8, k.shape[3]), dtype=torch.float32, device=self.device)
        kp[:, :, :cfg.in_channels] = k
        self._conv_in_w = ops.cast_bf16(kp)

    def encode(self, pixel_values: torch.Tensor) -> torch.Tensor:
        cfg = self.cfg
        if pixel_values.device != self.device:
            raise ValueError(f"VAEEncoder.encode: pixel_values are on {pixel_values.device}, the VAE on {self.device}")
        B, Cc, H, W = pixel_values.shape
        down = 2 ** (len(cfg.block_out_channels) - 1)
        if Cc != cfg.in_channels or H % (2 * down) or W % (2 * down):
            raise ValueError(f"VAEEncoder.encode: expected [B, {cfg.in_channels}, H, W] with H, W multiples of {2 * down}")
        px = pixel_values.to(torch.float32).contiguous()
        x = torch.empty((B, H, W, 8), device=self.device, dtype=BF16)
        _lib.check(_lib.lib().sdb_nchw_f32_to_nhwc_bf16_pad(px.data_ptr(), x.data_ptr(), B, Cc, 8, H * W,
                                                            _lib.stream_ptr(self.device)), "sdb_nchw_f32_to_nhwc_bf16_pad")
        x = ops.conv3x3(x, self._conv_in_w, bias=self.P["encoder/conv_in/bias"])
        n = len(cfg.block_out_channels)
        for i in range(n):
            for j in range(cfg.layers_per_block):
                x = self._resnet(x, f"encoder/down_blocks_{i}/resnets_{j}")
            if i != n - 1:
                x = self._downsample(x, f"encoder/down_blocks_{i}/downsamplers_0/conv")
        x = self._resnet(x, "encoder/mid_block/resnets_0")
        x = self._attention(x, "encoder/mid_block/attentions_0")
        x = self._resnet(x, "encoder/mid_block/resnets_1")
        Bq, Hq, Wq, C = x.shape
        x = self._conv(self._gn(x, B, "encoder/conv_norm_out", True).view(Bq, Hq, Wq, C), "encoder/conv_out")
        mom = self._linear(x.view(-1, x.shape[3]), "quant_conv")
        return mom.view(B, Hq * Wq, 2 * cfg.latent_channels)

    def sample(self, moments: torch.Tensor, hw: Tuple[int, int], key=None, *, batch_global: Optional[int] = None,
               batch_begin: int = 0) -> torch.Tensor:
        """latent_dist.sample(key) -> NCHW fp32 latents * scaling_factor (training_utils.py:582-586). key None: the mode.
        The N(0,1) draw has the NHWC shape of the mean, as FlaxDiagonalGaussianDistribution.sample draws it; under data
        parallelism a rank draws its rows of the global draw."""
        B, HW, twoL = moments.shape
        L = twoL // 2
        eps = None
        if key is not None:
            eps = JR.sample_noise(key, (B, 1, HW * L), self.device, batch_global=batch_global, batch_begin=batch_begin)
        lat = torch.empty((B, L, hw[0], hw[1]), device=self.device, dtype=torch.float32)
        _lib.check(_lib.lib().sdb_vae_sample(moments.data_ptr(), _lib.ptr(eps), lat.data_ptr(), B, L, HW,
                                             float(self.cfg.scaling_factor), _lib.stream_ptr(self.device)), "sdb_vae_sample")
        return lat


class VAEDecoder(_VAEExec):
    """decode_nhwc(latents NCHW fp32 [B, L, h, w], scale) -> NHWC bf16 [B, 8h, 8w, 4] (channel 3 is padding);
    decode() / decode_image() read it out as the reference's `.sample` (NCHW) or as post-processed pixels."""

    def __init__(self, cfg: VAEConfig, params: dict, device):
        super().__init__(cfg, params, device, decoder_param_shapes(cfg), "decoder")
        if cfg.latent_channels != 4 or cfg.in_channels > 4:
            raise NotImplementedError("VAEDecoder: 4 latent channels and at most 4 image channels")
        # conv_out (C -> 3) on the 4-output-channel CUDA-core kernel of the UNet's conv_out: zero fourth output channel
        k = self.P["decoder/conv_out/kernel"]
        kp = torch.zeros((3, 3, k.shape[2], 4), dtype=torch.float32, device=self.device)
        kp[..., :cfg.in_channels] = k
        self._conv_out_w = ops.cast_bf16(kp)
        bp = torch.zeros(4, dtype=torch.float32, device=self.device)
        bp[:cfg.in_channels] = self.P["decoder/conv_out/bias"]
        self._conv_out_b = bp

    def _upsample(self, x4, name):
        return self._conv(ops.upsample2x_fwd(x4), name + "/conv")

    def decode_nhwc(self, latents: torch.Tensor, scale: float = 1.0) -> torch.Tensor:
        cfg = self.cfg
        if latents.device != self.device:
            raise ValueError(f"VAEDecoder.decode: latents are on {latents.device}, the VAE on {self.device}")
        if latents.dim() != 4 or latents.shape[1] != cfg.latent_channels:
            raise ValueError(f"VAEDecoder.decode: expected NCHW latents [B, {cfg.latent_channels}, h, w]")
        B, L, h, w = latents.shape
        lat = latents.to(torch.float32).contiguous()
        x = torch.empty((B, h, w, L), device=self.device, dtype=BF16)
        _lib.check(_lib.lib().sdb_latents_to_nhwc_bf16(lat.data_ptr(), float(np.float32(scale)), 1, x.data_ptr(), B, L,
                                                       h * w, _lib.stream_ptr(self.device)), "sdb_latents_to_nhwc_bf16")
        x = self._linear(x.view(-1, L), "post_quant_conv").view(B, h, w, L)
        x = ops.conv3x3_c4in(x, self.W["decoder/conv_in/kernel"], bias=self.P["decoder/conv_in/bias"])
        x = self._resnet(x, "decoder/mid_block/resnets_0")
        x = self._attention(x, "decoder/mid_block/attentions_0")
        x = self._resnet(x, "decoder/mid_block/resnets_1")
        n = len(cfg.block_out_channels)
        for i in range(n):
            for j in range(cfg.layers_per_block + 1):
                x = self._resnet(x, f"decoder/up_blocks_{i}/resnets_{j}")
            if i != n - 1:
                x = self._upsample(x, f"decoder/up_blocks_{i}/upsamplers_0")
        Bq, Hq, Wq, C = x.shape
        x = self._gn(x, B, "decoder/conv_norm_out", True).view(Bq, Hq, Wq, C)
        return ops.conv3x3_c4out(x, self._conv_out_w, bias=self._conv_out_b)

    def decode(self, latents: torch.Tensor, scale: float = 1.0) -> torch.Tensor:
        """FlaxAutoencoderKL.decode(latents * scale).sample: NCHW fp32 [B, in_channels, 8h, 8w] (bf16-valued)."""
        y = self.decode_nhwc(latents, scale)
        B, H, W, _ = y.shape
        out = torch.empty((B, 4, H, W), dtype=torch.float32, device=self.device)
        _lib.check(_lib.lib().sdb_nhwc_bf16_to_nchw_f32(y.data_ptr(), out.data_ptr(), B, 4, H * W,
                                                        _lib.stream_ptr(self.device)), "sdb_nhwc_bf16_to_nchw_f32")
        return out[:, :self.cfg.in_channels]

    def decode_image(self, latents: torch.Tensor, scale: float = 1.0) -> torch.Tensor:
        """(decode(latents * scale) / 2 + 0.5).clip(0, 1).transpose(0, 2, 3, 1): NHWC fp32 [B, 8h, 8w, in_channels]."""
        y = self.decode_nhwc(latents, scale)
        B, H, W, _ = y.shape
        C = self.cfg.in_channels
        out = torch.empty((B, H, W, C), dtype=torch.float32, device=self.device)
        _lib.check(_lib.lib().sdb_image_postprocess(y.data_ptr(), out.data_ptr(), B * H * W, 4, C,
                                                    _lib.stream_ptr(self.device)), "sdb_image_postprocess")
        return out


class AutoencoderKL:
    """Stand-in for the FlaxAutoencoderKL module object (`FrozenModel(call=vae, params=vae_params)` in the reference):
    `.config`, `.encoder(params, device)` / `.decoder(params, device)` build / cache the executors for a parameter tree."""

    def __init__(self, config: VAEConfig = SD_VAE):
        self.config = config
        self._enc = None
        self._dec = None

    def init_weights(self, seed: int = 0, decoder: bool = False) -> dict:
        """Synthetic tree: the encoder half (what training needs); decoder=True adds the decode half."""
        flat = init_params_numpy(self.config, seed)
        if decoder:
            flat.update(init_decoder_params_numpy(self.config, seed))
        return unflatten_dict(flat)

    def decoder(self, params: dict, device) -> VAEDecoder:
        if self._dec is None or self._dec[0] is not params or self._dec[1].device != torch.device(device):
            self._dec = (params, VAEDecoder(self.config, params, device))
        return self._dec[1]

    def encoder(self, params: dict, device) -> VAEEncoder:
        if self._enc is None or self._enc[0] is not params or self._enc[1].device != torch.device(device):
            self._enc = (params, VAEEncoder(self.config, params, device))
        return self._enc[1]
