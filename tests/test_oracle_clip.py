"""CPU pin of the text-encoder oracle (oracle/clip.py) against a real third-party implementation: transformers'
PyTorch `CLIPTextModel` — the same architecture, shipped by the same package as the FlaxCLIPTextModel the reference loads
(/root/reference/training_utils.py:215-217) — loaded with the oracle's Flax-named parameters through the published
Flax -> PyTorch conversion (Dense kernel (in,out) -> weight (out,in); embedding -> weight; scale -> weight). Forward and
parameter gradients must agree in fp64. Also: the context assembly of training_utils.py:641-674."""
import numpy as np
import pytest
import torch

from oracle import clip as OC


def _hf_model(cfg: OC.CLIPTextConfig, params):
    tf = pytest.importorskip("transformers")
    hf_cfg = tf.CLIPTextConfig(vocab_size=cfg.vocab_size, hidden_size=cfg.hidden_size,
                               intermediate_size=cfg.intermediate_size, num_hidden_layers=cfg.num_hidden_layers,
                               num_attention_heads=cfg.num_attention_heads,
                               max_position_embeddings=cfg.max_position_embeddings, hidden_act=cfg.hidden_act,
                               layer_norm_eps=cfg.layer_norm_eps, attention_dropout=0.0)
    m = tf.CLIPTextModel(hf_cfg).double().eval()
    sd = {}
    for name, v in params.items():
        parts = name.split("/")
        leaf = parts[-1]
        key = ".".join(parts[:-1])
        v = v.double()
        if leaf == "kernel":
            sd[key + ".weight"] = v.t().contiguous()
        elif leaf in ("embedding", "scale"):
            sd[key + ".weight"] = v
        else:
            sd[key + ".bias"] = v
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected
    assert all("position_ids" in k for k in missing), missing
    return m


@pytest.mark.parametrize("cfg", [OC.TINY_CLIP, OC.CLIPTextConfig(vocab_size=500, hidden_size=64, intermediate_size=96,
                                                                 num_hidden_layers=3, num_attention_heads=4,
                                                                 hidden_act="gelu")], ids=["quick_gelu", "gelu"])
def test_oracle_equals_transformers_clip_text_model(cfg):
    params = OC.init_params(cfg, 1)
    m = _hf_model(cfg, params)
    g = torch.Generator().manual_seed(3)
    ids = torch.randint(0, cfg.vocab_size, (3, 77), generator=g)
    leaves = {k: v.double().requires_grad_(True) for k, v in params.items()}
    want = OC.text_encoder_forward(cfg, leaves, ids, mode="fp64")
    got = m(input_ids=ids).last_hidden_state
    assert (got - want).abs().max().item() <= 1e-10 * want.abs().max().item()
    w = torch.randn(want.shape, generator=g, dtype=torch.float64)
    (want * w).sum().backward()
    (got * w).sum().backward()
    hf = dict(m.named_parameters())
    checks = [("text_model/embeddings/token_embedding/embedding", hf["text_model.embeddings.token_embedding.weight"].grad),
              ("text_model/embeddings/position_embedding/embedding", hf["text_model.embeddings.position_embedding.weight"].grad),
              ("text_model/encoder/layers/0/self_attn/q_proj/kernel", hf["text_model.encoder.layers.0.self_attn.q_proj.weight"].grad.t()),
              ("text_model/encoder/layers/1/self_attn/out_proj/bias", hf["text_model.encoder.layers.1.self_attn.out_proj.bias"].grad),
              ("text_model/encoder/layers/1/mlp/fc1/kernel", hf["text_model.encoder.layers.1.mlp.fc1.weight"].grad.t()),
              ("text_model/encoder/layers/0/layer_norm2/scale", hf["text_model.encoder.layers.0.layer_norm2.weight"].grad),
              ("text_model/final_layer_norm/bias", hf["text_model.final_layer_norm.bias"].grad)]
    for name, gref in checks:
        gor = leaves[name].grad
        assert (gor - gref).abs().max().item() <= 1e-9 * max(gor.abs().max().item(), 1e-30), name


def test_parameter_census():
    # openai/clip-vit-large-patch14 text tower (SD-1.5): 123 060 480 parameters; 196 leaves
    s = OC.param_shapes(OC.SD15_CLIP)
    assert len(s) == 2 + 12 * 16 + 2 and sum(int(np.prod(v)) for v in s.values()) == 123_060_480
    # OpenCLIP ViT-H text tower as SD-2.x ships it (23 layers): 340 387 840
    assert sum(int(np.prod(v)) for v in OC.param_shapes(OC.SD21_CLIP).values()) == 340_387_840


@pytest.mark.parametrize("chunks", [1, 2, 3])
@pytest.mark.parametrize("strip", [True, False])
def test_context_assembly(chunks, strip):
    B, D = 2, 4
    h = torch.arange(B * chunks * 77 * D, dtype=torch.float32).reshape(B * chunks, 77, D)
    ctx = OC.assemble_context(h, B, strip)
    L = (76 + 75 * max(chunks - 2, 0) + 76) if strip else 77 * chunks
    assert ctx.shape == (B, L, D)
    idx = OC.context_row_index(B, chunks, strip)
    assert torch.equal(ctx, h.reshape(-1, D)[torch.from_numpy(idx).reshape(-1)].reshape(B, L, D))
    if strip and chunks == 3:  # the repo's default 3 x 75 + 2 context: 227 tokens
        assert L == 227
