"""CLIP text encoder forward + hand-scheduled backward on the sm_100a kernels (host side) — SURVEY.md 8(f)-2.

Mirrors transformers' FlaxCLIPTextModel as the reference loads and trains it
(/root/reference/training_utils.py:215-217 `FlaxCLIPTextModel.from_pretrained(..., dtype=jnp.bfloat16)`; 389-425 its own
`clip_by_global_norm(1) -> lion_8bit` TrainState with `apply_fn=text_encoder.__call__`; 635-640 the call inside the
step; 641-674 the context assembly; 733 `text_encoder_state.apply_gradients`). Parameter names and layouts are Flax's
(`text_model/encoder/layers/<i>/self_attn/q_proj/kernel` (in, out), ...), so trees interchange with the reference.

Built from the UNet's executor (unet.Graph: tape of explicit backward closures, parameter gradients on a second
stream): Dense / LayerNorm / bias gradients are the same kernels; embeddings, the 77-token causal attention, quick-GELU
and the row gather of the context assembly are csrc/clip.cu.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Optional

import numpy as np
import torch

from . import ops
from .unet import Graph, T


@dataclass(frozen=True)
class CLIPTextConfig:
    vocab_size: int = 49408
    hidden_size: int = 768
    intermediate_size: int = 3072
    num_hidden_layers: int = 12
    num_attention_heads: int = 12
    max_position_embeddings: int = 77
    hidden_act: str = "quick_gelu"
    layer_norm_eps: float = 1e-5


SD15_CLIP = CLIPTextConfig()  # openai/clip-vit-large-patch14 text tower
SD21_CLIP = CLIPTextConfig(hidden_size=1024, intermediate_size=4096, num_hidden_layers=23, num_attention_heads=16,
                           hidden_act="gelu")
TINY_CLIP = CLIPTextConfig(vocab_size=1000, hidden_size=32, intermediate_size=64, num_hidden_layers=2,
                           num_attention_heads=2)

_TOK = "text_model/embeddings/token_embedding/embedding"
_POS = "text_model/embeddings/position_embedding/embedding"


def param_shapes(cfg: CLIPTextConfig) -> Dict[str, tuple]:
    D, F = cfg.hidden_size, cfg.intermediate_size
    s = {_TOK: (cfg.vocab_size, D), _POS: (cfg.max_position_embeddings, D)}
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


def init_params_numpy(cfg: CLIPTextConfig, seed: int = 0) -> Dict[str, np.ndarray]:
    """Synthetic random-init weights (same generator and order as oracle/clip.py:init_params)."""
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
        out[name] = a.astype(np.float32)
    return out


def backward_param_order(cfg: CLIPTextConfig) -> List[str]:
    """Leaves in the order backward finishes their gradients: final norm, layers last to first, embeddings."""
    shapes = param_shapes(cfg)
    out = [n for n in shapes if n.startswith("text_model/final_layer_norm/")]
    for i in reversed(range(cfg.num_hidden_layers)):
        p = f"text_model/encoder/layers/{i}/"
        for sub in ("mlp/fc2", "mlp/fc1", "layer_norm2", "self_attn/out_proj", "self_attn/v_proj", "self_attn/k_proj",
                    "self_attn/q_proj", "layer_norm1"):
            out += [n for n in shapes if n.startswith(p + sub + "/")]
    out += [_POS, _TOK]
    assert len(out) == len(shapes) and len(set(out)) == len(out)
    return out


def context_row_index(batch: int, chunks: int, strip_bos_eos_token: bool = True) -> np.ndarray:
    """Row indices into the [batch * chunks * 77] token rows that the reference's context assembly keeps, in its
    order (training_utils.py:641-674): reshape to [B, chunks, 77, D]; with stripping the first chunk loses its last
    position, inner chunks their first and last, the last chunk its first; else plain concatenation."""
    idx = np.arange(batch * chunks * 77, dtype=np.int32).reshape(batch, chunks, 77)
    if not strip_bos_eos_token:
        return idx.reshape(batch, -1)
    return np.concatenate([idx[:, 0, :-1], idx[:, 1:-1, 1:-1].reshape(batch, -1), idx[:, -1, 1:]], axis=1)


class TextGraph(Graph):
    """unet.Graph + the text-encoder specific layers."""

    def embed(self, input_ids: torch.Tensor) -> T:
        tok, pos = self.W[_TOK], self.W[_POS]
        y = T(ops.embed_fwd(input_ids, tok, pos), None)

        def bwd():
            dy = y.g
            self._fork(lambda: ops.embed_bwd(input_ids, dy, self.G[_TOK], self.G[_POS]), dy)
            y.g = None
        self.tape.append((bwd, [_TOK, _POS]))
        return y

    def attention_small(self, q: T, k: T, v: T, N: int, heads: int, scale: float, causal: bool) -> T:
        ov, lse = ops.attention_small_fwd(q.v, k.v, v.v, N, heads, scale, causal)
        o = T(ov, None)

        def bwd():
            dq, dk, dv = ops.attention_small_bwd(q.v, k.v, v.v, ov, o.g, lse, N, heads, scale, causal)
            self._acc_safe(q, dq)
            self._acc_safe(k, dk)
            self._acc_safe(v, dv)
            o.g = None
        self.tape.append((bwd, None))
        return o

    def act(self, x: T, kind: str) -> T:
        y = T(ops.act_fwd(x.v, kind), None)

        def bwd():
            self._acc_safe(x, ops.act_bwd(x.v, y.g, kind))
            y.g = None
        self.tape.append((bwd, None))
        return y

    def gather_rows(self, x: T, row_index: torch.Tensor) -> T:
        """The context assembly: y[r] = x[row_index[r]] (each source row used at most once)."""
        rows_src = x.v.shape[0]
        y = T(ops.gather_rows(x.v, row_index, row_index.numel()), None)

        def bwd():
            if y.g is not None:
                self._acc_safe(x, ops.scatter_rows(y.g, row_index, rows_src))
            y.g = None
        self.tape.append((bwd, None))
        return y


class ClipText:
    """forward(input_ids int32 [N, S]) -> last_hidden_state T [N * S, D]; backward() replays the tape."""

    def __init__(self, cfg: CLIPTextConfig, P, W, G):
        self.cfg = cfg
        self.P, self.W, self.G = P, W, G
        self.graph: Optional[TextGraph] = None

    def forward(self, input_ids: torch.Tensor, forward_only: bool = False) -> T:
        cfg = self.cfg
        g = TextGraph(self.P, self.W, self.G, False, forward_only)
        self.graph = g
        N, S = input_ids.shape
        if S > cfg.max_position_embeddings:
            raise ValueError(f"sequence length {S} exceeds max_position_embeddings {cfg.max_position_embeddings}")
        H = cfg.num_attention_heads
        d = cfg.hidden_size // H
        x = g.embed(input_ids)
        for i in range(cfg.num_hidden_layers):
            p = f"text_model/encoder/layers/{i}"
            h = g.layernorm(x, p + "/layer_norm1")
            q = g.linear(h, p + "/self_attn/q_proj")
            k = g.linear(h, p + "/self_attn/k_proj")
            v = g.linear(h, p + "/self_attn/v_proj")
            o = g.attention_small(q, k, v, N, H, d ** -0.5, True)
            x = g.linear(o, p + "/self_attn/out_proj", res=x)
            h = g.layernorm(x, p + "/layer_norm2")
            h = g.act(g.linear(h, p + "/mlp/fc1"), cfg.hidden_act)
            x = g.linear(h, p + "/mlp/fc2", res=x)
        return g.layernorm(x, "text_model/final_layer_norm")

    def backward(self, on_grads=None) -> None:
        self.graph.backward(on_grads)
        self.graph = None


class CLIPTextModel:
    """Stand-in for the FlaxCLIPTextModel module object: `.config` and `__call__(params=, input_ids=)` (the reference
    passes `text_encoder.__call__` as the TrainState's apply_fn, training_utils.py:419-424)."""

    def __init__(self, config: CLIPTextConfig = SD15_CLIP):
        self.config = config
        self._state = None

    def init_weights(self, seed: int = 0) -> dict:
        from .lion_quant import unflatten_dict
        return unflatten_dict(init_params_numpy(self.config, seed))

    def __call__(self, input_ids=None, params=None, dropout_rng=None, train: bool = False):
        """-> (last_hidden_state [N, S, D] fp32,) — forward only (dropout rates are 0 in the CLIP configs)."""
        st = self._state
        if st is None or (params is not None and params is not st.params):
            raise ValueError("CLIPTextModel.__call__: pass params=state.params of the TrainState created with this model")
        ids = input_ids.to(torch.int32).contiguous()
        out = st.net.forward(ids, forward_only=True)
        st.net.graph = None
        N, S = ids.shape
        return (out.v.view(N, S, -1).float(),)
