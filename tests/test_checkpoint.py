"""Checkpoint container (stable_diffusion_training_b200/checkpoint.py): flax.serialization's msgpack format restated
(flax is an absent dependency) — structure checks with plain msgpack, round trips incl. int8 / 0-d / chunked leaves,
and on the GPU: save_model -> load_unet_params / load_optimizer_state reproduces the TrainState bit for bit
(/root/reference/training_utils.py:986-1025 save, 209-214 load)."""
import json
import os

import msgpack
import numpy as np
import pytest
import torch

from stable_diffusion_training_b200 import checkpoint as CK


def test_msgpack_container_structure():
    tree = {"conv_in": {"kernel": np.arange(24, dtype=np.float32).reshape(1, 2, 3, 4), "bias": np.zeros(4, np.float32)},
            "codes": np.array([[3, -127], [127, 0]], dtype=np.int8), "count": np.int32(7)}
    blob = CK.msgpack_serialize(tree)
    raw = msgpack.unpackb(blob, raw=False)  # no ext hook: ndarrays stay ExtType
    k = raw["conv_in"]["kernel"]
    assert isinstance(k, msgpack.ExtType) and k.code == 1
    shape, dtype_name, buf = msgpack.unpackb(k.data, raw=False)
    assert list(shape) == [1, 2, 3, 4] and dtype_name == "float32"
    assert buf == np.arange(24, dtype=np.float32).tobytes()
    assert raw["count"].code == 3  # numpy scalar
    back = CK.msgpack_restore(blob)
    assert back["codes"].dtype == np.int8 and np.array_equal(back["codes"], tree["codes"])
    assert np.array_equal(back["conv_in"]["kernel"], tree["conv_in"]["kernel"]) and back["count"] == 7


def test_msgpack_round_trip_torch_tuples_and_chunks(monkeypatch):
    monkeypatch.setattr(CK, "_MAX_CHUNK_SIZE", 64)
    rng = np.random.default_rng(0)
    big = rng.standard_normal((5, 7)).astype(np.float32)  # 140 bytes > 64: chunked
    tree = {"w": torch.from_numpy(big), "q": (torch.full((2, 16), 3, dtype=torch.int8), torch.ones(2, 1)),
            "step": torch.zeros([], dtype=torch.int32)}
    blob = CK.msgpack_serialize(tree)
    raw = msgpack.unpackb(blob, ext_hook=CK._ext_unpack, raw=False)
    assert raw["w"]["__msgpack_chunked_array__"] is True and raw["w"]["shape"] == [5, 7] and len(raw["w"]["chunks"]) == 3
    back = CK.msgpack_restore(blob)
    assert np.array_equal(back["w"], big)
    assert np.array_equal(back["q"]["0"], np.full((2, 16), 3, np.int8)) and back["q"]["1"].shape == (2, 1)
    assert back["step"].shape == () and back["step"].dtype == np.int32


@pytest.mark.gpu
def test_save_and_load_reproduce_the_train_state(cuda, tmp_path):
    from stable_diffusion_training_b200 import lion_quant as LQ
    from stable_diffusion_training_b200 import training_utils as TU
    from stable_diffusion_training_b200 import unet as PU
    cfg = PU.TINY
    params = LQ.unflatten_dict(PU.init_params_numpy(cfg, 0))
    state = TU.TrainState.create(apply_fn=TU.UNetModel(cfg).apply, params=params, tx=TU.create_lion_optimizer(params),
                                 device=cuda)
    ema = state.create_ema()
    sched = TU.default_scheduler(cuda, "scaled_linear", "epsilon")
    rng = np.random.default_rng(3)

    def batch(seed):
        r = np.random.default_rng(seed)
        return {k: torch.from_numpy(v).to(cuda) for k, v in {
            "latents": r.standard_normal((2, 4, 16, 16), dtype=np.float32),
            "noise": r.standard_normal((2, 4, 16, 16), dtype=np.float32),
            "timesteps": r.integers(0, 1000, 2).astype(np.int32),
            "encoder_hidden_states": r.standard_normal((2, 77, cfg.cross_attention_dim), dtype=np.float32)}.items()}
    for s in (1, 2):
        state, _ = TU.train_step(state, batch(s), None, sched, unet_ema_params=ema, ema_rate=0.99)
    out = str(tmp_path / "ckpt")
    CK.save_model(state, out, unet_ema_params=ema)
    for rel in ("unet/config.json", "unet/diffusion_flax_model.msgpack", "scheduler/scheduler_config.json",
                "model_index.json", "unet_ema/diffusion_flax_model.msgpack", "optimizer/lion8_state.msgpack"):
        assert os.path.exists(os.path.join(out, rel)), rel
    conf = json.load(open(os.path.join(out, "unet", "config.json")))
    assert conf["block_out_channels"] == [32, 64] and conf["attention_head_dim"] == 2
    assert json.load(open(os.path.join(out, "scheduler", "scheduler_config.json")))["prediction_type"] == "v_prediction"
    loaded = CK.load_unet_params(out)
    flat_l, flat_s = LQ.flatten_dict(loaded), LQ.flatten_dict(state.params)
    assert set(flat_l) == set(flat_s)
    for n in flat_s:
        assert flat_l[n].dtype == np.float32 and np.array_equal(flat_l[n], flat_s[n].cpu().numpy()), n
    # a new state from the checkpoint + the exported optimizer state continues identically
    st2 = TU.TrainState.create(apply_fn=TU.UNetModel(cfg).apply, params=loaded, tx=TU.create_lion_optimizer(loaded),
                               device=cuda)
    CK.load_optimizer_state(st2, out)
    assert int(st2.step) == 2 and torch.equal(st2.arena.codes, state.arena.codes)
    assert torch.equal(st2.arena.inv_scales, state.arena.inv_scales) and torch.equal(st2.arena.mu, state.arena.mu)
    b3 = batch(3)
    state, m1 = TU.train_step(state, b3, None, sched)
    st2, m2 = TU.train_step(st2, b3, None, sched)
    torch.cuda.synchronize()
    assert float(m1["loss"]) == float(m2["loss"])
    assert float((st2.arena.master != state.arena.master).float().mean()) < 1e-3  # wgrad red.add order only


def test_save_pipeline_writes_the_three_models_in_the_diffusers_layout(tmp_path):
    """The reference's save_model argument list (training_utils.py:986-1025); CPU only (NumPy / CPU-tensor trees)."""
    from stable_diffusion_training_b200 import clip as PC
    from stable_diffusion_training_b200 import lion_quant as LQ
    from stable_diffusion_training_b200 import training_utils as TU
    from stable_diffusion_training_b200 import unet as PU
    from stable_diffusion_training_b200 import vae as PV
    up = PU.init_params_numpy(PU.TINY, 0)
    tp = PC.init_params_numpy(PC.TINY_CLIP, 1)
    vp = dict(PV.init_params_numpy(PV.TINY_VAE, 2), **PV.init_decoder_params_numpy(PV.TINY_VAE, 2))
    models = {"unet": TU.UNetModel(PU.TINY), "text_encoder": PC.CLIPTextModel(PC.TINY_CLIP),
              "vae": PV.AutoencoderKL(PV.TINY_VAE)}

    class Tok:
        def save_pretrained(self, d):
            os.makedirs(d, exist_ok=True)
            open(os.path.join(d, "vocab.json"), "w").write("{}")

    ttree = LQ.unflatten_dict({k: torch.from_numpy(v) for k, v in tp.items()})   # tensors and arrays both serialise
    CK.save_pipeline(models, Tok(), LQ.unflatten_dict(up), ttree, LQ.unflatten_dict(vp), str(tmp_path))
    for rel in ("unet/diffusion_flax_model.msgpack", "unet/config.json", "text_encoder/flax_model.msgpack",
                "text_encoder/config.json", "vae/diffusion_flax_model.msgpack", "vae/config.json",
                "scheduler/scheduler_config.json", "model_index.json", "tokenizer/vocab.json"):
        assert os.path.exists(os.path.join(tmp_path, rel)), rel
    idx = json.load(open(os.path.join(tmp_path, "model_index.json")))
    assert idx["_class_name"] == "FlaxStableDiffusionPipeline" and idx["text_encoder"] == ["transformers", "FlaxCLIPTextModel"]
    assert idx["vae"] == ["diffusers", "FlaxAutoencoderKL"] and idx["scheduler"] == ["diffusers", "FlaxDDIMScheduler"]
    sc = json.load(open(os.path.join(tmp_path, "scheduler", "scheduler_config.json")))
    assert sc["prediction_type"] == "v_prediction" and sc["beta_schedule"] == "scaled_linear"   # training_utils.py:997-1004
    cc = json.load(open(os.path.join(tmp_path, "text_encoder", "config.json")))
    assert cc["hidden_size"] == PC.TINY_CLIP.hidden_size and cc["hidden_act"] == "quick_gelu"
    vc = json.load(open(os.path.join(tmp_path, "vae", "config.json")))
    assert vc["block_out_channels"] == list(PV.TINY_VAE.block_out_channels) and vc["scaling_factor"] == 0.18215
    for loaded, want in ((CK.load_unet_params(str(tmp_path)), up), (CK.load_text_encoder_params(str(tmp_path)), tp),
                         (CK.load_vae_params(str(tmp_path)), vp)):
        flat = LQ.flatten_dict(loaded)
        assert set(flat) == set(want)
        assert all(flat[k].dtype == np.float32 and np.array_equal(flat[k], want[k]) for k in want)
    # partial saves: what is None is left out of the folder and of the index
    CK.save_pipeline(models, None, LQ.unflatten_dict(up), None, None, str(tmp_path / "unet_only"))
    idx = json.load(open(os.path.join(tmp_path, "unet_only", "model_index.json")))
    assert "text_encoder" not in idx and "vae" not in idx and "tokenizer" not in idx
