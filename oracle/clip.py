"""ORACLE (test infrastructure only — never imported by the product path).

CPU restatement (PyTorch-CPU tensors + autograd, fp64 / fp32 / bf16-emulated) of the text encoder the reference trains
inside its step: transformers' FlaxCLIPTextModel, loaded at /root/reference/training_utils.py:215-217
(`FlaxCLIPTextModel.from_pretrained(..., dtype=jnp.bfloat16)`), called at 635-640 (`text_encoder_state.apply_fn(params=...,
input_ids=batch["input_ids"], dropout_rng=..., train=True)[0]`), followed by the reference's own context assembly
(641-674: reshape to [B, chunks, 77, D], strip BOS / EOS tokens of the inner chunks, concatenate).

transformers' Flax model is third-party and flax is absent here, so the module is restated from its published source
(modeling_flax_clip.py: FlaxCLIPTextEmbeddings, FlaxCLIPAttention with `query * scale` and the causal mask,
FlaxCLIPMLP with ACT2FN[hidden_act], pre-LayerNorm encoder layers, final_layer_norm). PIN: the same architecture ships
in the SAME package as a PyTorch implementation (transformers.CLIPTextModel), which IS importable here —
tests/test_oracle_clip.py loads the Flax-named parameters into it through the published Flax <-> PyTorch conversion rule
and requires equality of the forward and of parameter gradients in fp64. Attention dropout is 0 in the CLIP configs, so
train=True changes nothing.

Parameter tree (Flax names, fp32 masters):
  text_model/embeddings/{token_embedding,position_embedding}/embedding
  text_model/encoder/layers/<i>/{layer_norm1,layer_norm2}/{scale,bias}
  text_model/encoder/layers/<i>/self_attn/{q_proj,k_proj,v_proj,out_proj}/{kernel (in,out), bias}
  text_model/encoder/layers/<i>/mlp/{fc1,fc2}/{kernel, bias}
  text_model/final_layer_norm/{scale,bias}
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict

import torch


@dataclass
class CLIPTextConfig:
    vocab_size: int = 49408
    hidden_size: int = 768
    intermediate_size: int = 3072
    num_hidden_layers: int = 12
    num_attention_heads: int = 12
    max_position_embeddings: int = 77
    hidden_act: str = "quick_gelu"
    layer_norm_eps: float = 1e-5


SD15_CLIP = CLIPTextConfig()
SD21_CLIP = CLIPTextConfig(hidden_size=1024, intermediate_size=4096, num_hidden_layers=23, num_attention_heads=16,
                           hidden_act="gelu")
TINY_CLIP = CLIPTextConfig(vocab_size=1000, hidden_size=32, intermediate_size=64, num_hidden_layers=2,
                           num_attention_heads=2)


def param_shapes(cfg: CLIPTextConfig) -> Dict[str, tuple]:
    D, F = cfg.hidden_size, cfg.intermediate_size
    s = {"text_model/embeddings/token_embedding/embedding": (cfg.vocab_size, D),
         "text_model/embeddings/position_embedding/embedding": (cfg.max_position_embeddings, D)}
    for i in range(cfg.num_hidden_layers):
        p = f"text_model/encoder/layers/{i}"
        for ln in ("layer_norm1", "layer_norm2"):
            s[f"{p}/{ln}/scale"] = (D,)
            s[f"{p}/{ln}/bias"] = (D,)
        for a in ("q_proj", "k_proj", "v_proj", "out_proj"):
            s[f"{p}/self_attn/{a}/kernel"] = (D, D)
            s[f"{p}/self_attn/{a}/bias"] = (D,)
        s[f"{p}/mlp/fc1/kernel"] = (D, F)
        s[f"{p}/mlp/fc1/bias"] = (F,)
        s[f"{p}/mlp/fc2/kernel"] = (F, D)
        s[f"{p}/mlp/fc2/bias"] = (D,)
    s["text_model/final_layer_norm/scale"] = (D,)
    s["text_model/final_layer_norm/bias"] = (D,)
    return dict(sorted(s.items()))


def init_params(cfg: CLIPTextConfig, seed: int = 0) -> Dict[str, torch.Tensor]:
    """Synthetic weights: NumPy default_rng(seed), sorted-path order; kernels N(0, 1/fan_in), embeddings N(0, 0.02^2),
    biases N(0, 0.02^2), norm scale 1 + N(0, 0.1^2), norm bias N(0, 0.1^2)."""
    import numpy as np
    rng = np.random.default_rng(seed)
    out = {}
    for name, shape in param_shapes(cfg).items():
        leaf = name.rsplit("/", 1)[1]
        parent = name.rsplit("/", 2)[-2]
        if leaf == "kernel":
            a = rng.standard_normal(shape, dtype=np.float32) * np.float32(1.0 / math.sqrt(shape[0]))
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
        return x + (x.detach().to(torch.bfloat16).to(x.dtype) - x.detach())

    def w(self, name):
        return self.rnd(self.p[name].to(self.dtype))


def _dense(c: _Ctx, x, name):
    return c.rnd(c.rnd(c.rnd(x) @ c.w(name + "/kernel")) + c.w(name + "/bias"))


def _layer_norm(c: _Ctx, x, name, eps):
    mean = x.mean(-1, keepdim=True)
    var = ((x * x).mean(-1, keepdim=True) - mean * mean).clamp_min(0)  # flax use_fast_variance
    y = (x - mean) * torch.rsqrt(var + eps)
    return c.rnd(y * c.p[name + "/scale"].to(c.dtype) + c.p[name + "/bias"].to(c.dtype))


def _act(cfg, z):
    if cfg.hidden_act == "quick_gelu":
        return z * torch.sigmoid(1.702 * z)
    if cfg.hidden_act == "gelu":  # transformers ACT2FN["gelu"] = nn.gelu(approximate=False)
        return 0.5 * z * (1.0 + torch.erf(z / math.sqrt(2.0)))
    raise ValueError(f"hidden_act {cfg.hidden_act}")


def text_encoder_forward(cfg: CLIPTextConfig, params: Dict[str, torch.Tensor], input_ids, mode="fp32"):
    """FlaxCLIPTextModel(...)[0] = last_hidden_state [N, S, D] (after final_layer_norm)."""
    c = _Ctx(params, mode)
    N, S = input_ids.shape
    H = cfg.num_attention_heads
    D = cfg.hidden_size
    d = D // H
    tok = c.w("text_model/embeddings/token_embedding/embedding")[input_ids.long()]
    pos = c.w("text_model/embeddings/position_embedding/embedding")[:S]
    x = c.rnd(tok + pos[None])
    causal = torch.full((S, S), float("-inf"), dtype=c.dtype).triu(1)
    for i in range(cfg.num_hidden_layers):
        p = f"text_model/encoder/layers/{i}"
        h = _layer_norm(c, x, p + "/layer_norm1", cfg.layer_norm_eps)
        q = c.rnd(_dense(c, h, p + "/self_attn/q_proj") * (d ** -0.5))
        k = _dense(c, h, p + "/self_attn/k_proj")
        v = _dense(c, h, p + "/self_attn/v_proj")

        def heads(t):
            return t.reshape(N, S, H, d).permute(0, 2, 1, 3)

        w = c.rnd(heads(q) @ heads(k).transpose(-1, -2)) + causal
        w = c.rnd(torch.softmax(w, dim=-1))
        o = c.rnd(w @ heads(v)).permute(0, 2, 1, 3).reshape(N, S, D)
        x = c.rnd(x + _dense(c, o, p + "/self_attn/out_proj"))
        h = _layer_norm(c, x, p + "/layer_norm2", cfg.layer_norm_eps)
        h = c.rnd(_act(cfg, _dense(c, h, p + "/mlp/fc1")))
        x = c.rnd(x + _dense(c, h, p + "/mlp/fc2"))
    return _layer_norm(c, x, "text_model/final_layer_norm", cfg.layer_norm_eps)


def assemble_context(hidden, batch: int, strip_bos_eos_token: bool = True):
    """training_utils.py:641-674: [B * chunks, 77, D] -> [B, L, D]. With stripping: the first chunk loses its EOS
    position, inner chunks lose BOS and EOS, the last chunk loses its BOS (chunks == 1: the slices [:, 0, :-1] and
    [:, -1, 1:] are both taken from the single chunk, as the reference's code does)."""
    D = hidden.shape[-1]
    e = hidden.reshape(batch, -1, 77, D)
    if not strip_bos_eos_token:
        return e.reshape(batch, -1, D)
    return torch.cat([e[:, 0, :-1, :], e[:, 1:-1, 1:-1, :].reshape(batch, -1, D), e[:, -1, 1:, :]], dim=1)


def context_row_index(batch: int, chunks: int, strip_bos_eos_token: bool = True):
    """Row indices into the [B * chunks * 77] token rows that assemble_context gathers (same order), as a list."""
    import numpy as np
    idx = np.arange(batch * chunks * 77).reshape(batch, chunks, 77)
    if not strip_bos_eos_token:
        return idx.reshape(batch, -1)
    return np.concatenate([idx[:, 0, :-1], idx[:, 1:-1, 1:-1].reshape(batch, -1), idx[:, -1, 1:]], axis=1)
