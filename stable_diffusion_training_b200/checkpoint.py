"""Checkpoint I/O in the layout the reference ecosystem reads and writes.

/root/reference/training_utils.py:986-1025 (`save_model`) wraps the trained parameter trees into
`FlaxStableDiffusionPipeline.save_pretrained`, i.e. the diffusers folder layout

    <dir>/unet/config.json                          the FlaxUNet2DConditionModel config
    <dir>/unet/diffusion_flax_model.msgpack         flax.serialization.to_bytes(params)
    <dir>/text_encoder/config.json                  the CLIPTextConfig          (transformers: flax_model.msgpack)
    <dir>/text_encoder/flax_model.msgpack
    <dir>/vae/config.json                           the FlaxAutoencoderKL config
    <dir>/vae/diffusion_flax_model.msgpack
    <dir>/scheduler/scheduler_config.json           the DDIM placeholder of training_utils.py:997-1004
    <dir>/model_index.json

and `from_pretrained` reads the same files back (training_utils.py:209-222). This module writes / reads the UNet, the
text encoder and the VAE parameter trees bit-exactly (`save_model` for a TrainState, `save_pipeline` with the
reference's own `save_model` argument list), plus what the reference never saves but SURVEY asks to keep exportable:
the 8-bit Lion state (`optimizer/lion8_state.msgpack`: int8 codes + fp32 reciprocal-absmax scales per quantised leaf in
the reference's (n/bs, bs) / (n/bs, 1) layout, fp32 momentum otherwise, counters) and the EMA tree.

flax / diffusers are absent third-party dependencies here; the msgpack container is flax.serialization's published
format, restated: a nested dict packed with msgpack (`strict_types=True`), every ndarray as ExtType(1, packb((shape,
dtype.name, raw C-order bytes))), numpy scalars as ExtType(3, ...), arrays above 2**30 bytes split into chunks
({"__msgpack_chunked_array__": True, "shape": ..., "chunks": {"0": ..., ...}}). Parameters are stored in Flax layouts
(Dense (in, out), Conv HWIO, fp32) — exactly `state.params`.
"""
from __future__ import annotations

import json
import os
from typing import Any, Dict, Optional

import msgpack
import numpy as np
import torch

_EXT_NDARRAY, _EXT_NPSCALAR = 1, 3
_MAX_CHUNK_SIZE = 2 ** 30


def _dtype_from_name(name: str):
    if name == "bfloat16":
        raise NotImplementedError("bfloat16 leaves need ml_dtypes; the reference stores fp32 masters")
    return np.dtype(name)


def _ndarray_to_bytes(arr: np.ndarray) -> bytes:
    arr = np.asarray(arr)
    if arr.dtype.hasobject or arr.dtype.isalignedstruct:
        raise ValueError("object and structured dtypes are not serialisable")
    return msgpack.packb((arr.shape, arr.dtype.name, arr.tobytes("C")), use_bin_type=True)


def _ndarray_from_bytes(data: bytes) -> np.ndarray:
    shape, dtype_name, buf = msgpack.unpackb(data, raw=True)
    return np.frombuffer(buf, dtype=_dtype_from_name(dtype_name.decode() if isinstance(dtype_name, bytes) else dtype_name),
                         count=-1, offset=0).reshape(shape, order="C")


def _ext_pack(x):
    if isinstance(x, np.ndarray):
        return msgpack.ExtType(_EXT_NDARRAY, _ndarray_to_bytes(x))
    if isinstance(x, np.generic):
        return msgpack.ExtType(_EXT_NPSCALAR, _ndarray_to_bytes(np.asarray(x)))
    return x


def _ext_unpack(code, data):
    if code == _EXT_NDARRAY:
        return _ndarray_from_bytes(data)
    if code == _EXT_NPSCALAR:
        return _ndarray_from_bytes(data)[()]
    return msgpack.ExtType(code, data)


def _chunk(arr: np.ndarray) -> Dict[str, Any]:
    """flax.serialization._chunk: leaves above 2**30 bytes are split along the flattened axis."""
    chunksize = max(1, int(_MAX_CHUNK_SIZE / arr.dtype.itemsize))
    data = {"__msgpack_chunked_array__": True, "shape": list(arr.shape)}
    flat = arr.reshape(-1)
    chunks = [flat[i:i + chunksize] for i in range(0, flat.size, chunksize)]
    data["chunks"] = {str(i): c for i, c in enumerate(chunks)}
    return data


def _chunk_leaves(tree):
    if isinstance(tree, dict):
        return {k: _chunk_leaves(v) for k, v in tree.items()}
    if isinstance(tree, np.ndarray) and tree.size * tree.dtype.itemsize > _MAX_CHUNK_SIZE:
        return _chunk(tree)
    return tree


def _unchunk_leaves(tree):
    if isinstance(tree, dict):
        if "__msgpack_chunked_array__" in tree:
            n = len(tree["chunks"])
            flat = np.concatenate([tree["chunks"][str(i)] for i in range(n)])
            return flat.reshape(tree["shape"])
        return {k: _unchunk_leaves(v) for k, v in tree.items()}
    return tree


def _to_numpy_tree(tree):
    if isinstance(tree, dict):
        return {str(k): _to_numpy_tree(v) for k, v in tree.items()}
    if isinstance(tree, (tuple, list)):  # flax serialises tuples / lists as dicts with string indices
        return {str(i): _to_numpy_tree(v) for i, v in enumerate(tree)}
    if isinstance(tree, torch.Tensor):
        return tree.detach().cpu().numpy()
    return tree


def msgpack_serialize(tree) -> bytes:
    """flax.serialization.msgpack_serialize of a nested dict of arrays (torch tensors are copied to the host)."""
    return msgpack.packb(_chunk_leaves(_to_numpy_tree(tree)), default=_ext_pack, strict_types=True)


def msgpack_restore(data: bytes):
    """flax.serialization.msgpack_restore -> nested dict of numpy arrays."""
    return _unchunk_leaves(msgpack.unpackb(data, ext_hook=_ext_unpack, raw=False, strict_map_key=False))


# ---------------------------------------------------------------------------------------------------------------
def unet_config_dict(cfg) -> dict:
    """config.json of FlaxUNet2DConditionModel (diffusers 0.21.4 field names) for a unet.UNetConfig."""
    heads = cfg.attention_head_dim
    return {
        "_class_name": "FlaxUNet2DConditionModel", "_diffusers_version": "0.21.4",
        "sample_size": 64, "in_channels": cfg.in_channels, "out_channels": cfg.out_channels,
        "down_block_types": list(cfg.down_block_types), "up_block_types": list(cfg.up_block_types),
        "only_cross_attention": False, "block_out_channels": list(cfg.block_out_channels),
        "layers_per_block": cfg.layers_per_block,
        "attention_head_dim": heads if isinstance(heads, int) else list(heads), "num_attention_heads": None,
        "cross_attention_dim": cfg.cross_attention_dim, "dropout": 0.0,
        "use_linear_projection": bool(cfg.use_linear_projection), "flip_sin_to_cos": bool(cfg.flip_sin_to_cos),
        "freq_shift": cfg.freq_shift, "use_memory_efficient_attention": True, "transformer_layers_per_block": 1,
    }


DDIM_PLACEHOLDER = {  # training_utils.py:997-1004
    "_class_name": "FlaxDDIMScheduler", "_diffusers_version": "0.21.4", "beta_start": 0.00085, "beta_end": 0.012,
    "beta_schedule": "scaled_linear", "num_train_timesteps": 1000, "prediction_type": "v_prediction",
    "set_alpha_to_one": True, "steps_offset": 0, "trained_betas": None,
}


def save_model(state, output_dir: str, *, unet_ema_params: Optional[dict] = None, save_optimizer_state: bool = True
               ) -> None:
    """save_model (training_utils.py:986-1025) for the UNet this framework trains: diffusers Flax folder layout."""
    os.makedirs(os.path.join(output_dir, "unet"), exist_ok=True)
    os.makedirs(os.path.join(output_dir, "scheduler"), exist_ok=True)
    torch.cuda.synchronize(state.arena.device)
    with open(os.path.join(output_dir, "unet", "diffusion_flax_model.msgpack"), "wb") as f:
        f.write(msgpack_serialize(state.params))
    with open(os.path.join(output_dir, "unet", "config.json"), "w") as f:
        json.dump(unet_config_dict(state.net.cfg), f, indent=2)
    with open(os.path.join(output_dir, "scheduler", "scheduler_config.json"), "w") as f:
        json.dump(DDIM_PLACEHOLDER, f, indent=2)
    with open(os.path.join(output_dir, "model_index.json"), "w") as f:
        json.dump({"_class_name": "FlaxStableDiffusionPipeline", "_diffusers_version": "0.21.4",
                   "unet": ["diffusers", "FlaxUNet2DConditionModel"],
                   "scheduler": ["diffusers", "FlaxDDIMScheduler"]}, f, indent=2)
    if unet_ema_params is not None:
        os.makedirs(os.path.join(output_dir, "unet_ema"), exist_ok=True)
        with open(os.path.join(output_dir, "unet_ema", "diffusion_flax_model.msgpack"), "wb") as f:
            f.write(msgpack_serialize(unet_ema_params))
    if save_optimizer_state:
        os.makedirs(os.path.join(output_dir, "optimizer"), exist_ok=True)
        a = state.arena
        lion = state.opt_state[1] if isinstance(state.opt_state, tuple) and len(state.opt_state) == 2 else state.opt_state
        tree = {"step": state.step, "count": lion.lion.count, "schedule_count": lion.schedule.count,
                "block_size": np.int32(a.block_size), "mu_quant": lion.lion.mu_quant}
        with open(os.path.join(output_dir, "optimizer", "lion8_state.msgpack"), "wb") as f:
            f.write(msgpack_serialize(tree))


def clip_config_dict(cfg) -> dict:
    """config.json of transformers' CLIPTextModel for a clip.CLIPTextConfig (transformers field names)."""
    return {"architectures": ["CLIPTextModel"], "model_type": "clip_text_model", "vocab_size": cfg.vocab_size,
            "hidden_size": cfg.hidden_size, "intermediate_size": cfg.intermediate_size,
            "num_hidden_layers": cfg.num_hidden_layers, "num_attention_heads": cfg.num_attention_heads,
            "max_position_embeddings": cfg.max_position_embeddings, "hidden_act": cfg.hidden_act,
            "layer_norm_eps": cfg.layer_norm_eps, "attention_dropout": 0.0, "dropout": 0.0}


def vae_config_dict(cfg) -> dict:
    """config.json of FlaxAutoencoderKL (diffusers 0.21.4 field names) for a vae.VAEConfig."""
    n = len(cfg.block_out_channels)
    return {"_class_name": "FlaxAutoencoderKL", "_diffusers_version": "0.21.4", "in_channels": cfg.in_channels,
            "out_channels": cfg.in_channels, "down_block_types": ["DownEncoderBlock2D"] * n,
            "up_block_types": ["UpDecoderBlock2D"] * n, "block_out_channels": list(cfg.block_out_channels),
            "layers_per_block": cfg.layers_per_block, "act_fn": "silu", "latent_channels": cfg.latent_channels,
            "norm_num_groups": cfg.norm_num_groups, "sample_size": 512, "scaling_factor": cfg.scaling_factor}


def _write_tree(directory: str, filename: str, tree, config: Optional[dict]) -> None:
    os.makedirs(directory, exist_ok=True)
    with open(os.path.join(directory, filename), "wb") as f:
        f.write(msgpack_serialize(tree))
    if config is not None:
        with open(os.path.join(directory, "config.json"), "w") as f:
            json.dump(config, f, indent=2)


def save_pipeline(model_object_dict: dict, tokenizer_object: Any, unet_params: dict, text_encoder_params: dict,
                  vae_params: dict, output_dir: str) -> None:
    """The reference's save_model (training_utils.py:986-1025), same argument list: `model_object_dict` holds the module
    stand-ins {"unet": UNetModel, "text_encoder": clip.CLIPTextModel, "vae": vae.AutoencoderKL}; the three parameter
    trees (Flax layouts; device tensors or NumPy) go to the diffusers folder layout with the DDIM placeholder scheduler.
    `tokenizer_object`: saved through its own `save_pretrained(<dir>/tokenizer)` when it has one (CLIPTokenizer)."""
    devs = {v.device for t in (unet_params, text_encoder_params, vae_params) if t is not None
            for v in _flat_leaves(t) if isinstance(v, torch.Tensor) and v.is_cuda}
    for d in devs:
        torch.cuda.synchronize(d)
    index = {"_class_name": "FlaxStableDiffusionPipeline", "_diffusers_version": "0.21.4",
             "scheduler": ["diffusers", "FlaxDDIMScheduler"]}
    if unet_params is not None:
        _write_tree(os.path.join(output_dir, "unet"), "diffusion_flax_model.msgpack", unet_params,
                    unet_config_dict(model_object_dict["unet"].config))
        index["unet"] = ["diffusers", "FlaxUNet2DConditionModel"]
    if text_encoder_params is not None:
        _write_tree(os.path.join(output_dir, "text_encoder"), "flax_model.msgpack", text_encoder_params,
                    clip_config_dict(model_object_dict["text_encoder"].config))
        index["text_encoder"] = ["transformers", "FlaxCLIPTextModel"]
    if vae_params is not None:
        _write_tree(os.path.join(output_dir, "vae"), "diffusion_flax_model.msgpack", vae_params,
                    vae_config_dict(model_object_dict["vae"].config))
        index["vae"] = ["diffusers", "FlaxAutoencoderKL"]
    if tokenizer_object is not None and hasattr(tokenizer_object, "save_pretrained"):
        tokenizer_object.save_pretrained(os.path.join(output_dir, "tokenizer"))
        index["tokenizer"] = ["transformers", "CLIPTokenizer"]
    os.makedirs(os.path.join(output_dir, "scheduler"), exist_ok=True)
    with open(os.path.join(output_dir, "scheduler", "scheduler_config.json"), "w") as f:
        json.dump(DDIM_PLACEHOLDER, f, indent=2)
    with open(os.path.join(output_dir, "model_index.json"), "w") as f:
        json.dump(index, f, indent=2)


def _flat_leaves(tree):
    for v in tree.values():
        if isinstance(v, dict):
            yield from _flat_leaves(v)
        else:
            yield v


def _load_tree(model_dir: str, subfolder: str, filename: str) -> dict:
    sub = os.path.join(model_dir, subfolder)
    path = os.path.join(sub if os.path.isdir(sub) else model_dir, filename)
    with open(path, "rb") as f:
        return msgpack_restore(f.read())


def load_text_encoder_params(model_dir: str) -> dict:
    """FlaxCLIPTextModel.from_pretrained(..., subfolder="text_encoder") parameter tree (training_utils.py:215-217)."""
    return _load_tree(model_dir, "text_encoder", "flax_model.msgpack")


def load_vae_params(model_dir: str) -> dict:
    """FlaxAutoencoderKL.from_pretrained(..., subfolder="vae") parameter tree (training_utils.py:218-222)."""
    return _load_tree(model_dir, "vae", "diffusion_flax_model.msgpack")


def load_unet_params(model_dir: str) -> dict:
    """FlaxUNet2DConditionModel.from_pretrained(..., subfolder="unet") parameter tree (training_utils.py:209-214): nested
    dict of fp32 numpy arrays in Flax layouts, ready for TrainState.create(params=...)."""
    sub = os.path.join(model_dir, "unet")
    path = os.path.join(sub if os.path.isdir(sub) else model_dir, "diffusion_flax_model.msgpack")
    with open(path, "rb") as f:
        return msgpack_restore(f.read())


def load_optimizer_state(state, model_dir: str) -> None:
    """Restore the exported 8-bit Lion state into a TrainState built over the same parameter tree (in place)."""
    with open(os.path.join(model_dir, "optimizer", "lion8_state.msgpack"), "rb") as f:
        tree = msgpack_restore(f.read())
    a = state.arena
    if int(tree["block_size"]) != a.block_size:
        raise ValueError(f"optimizer state was saved with block_size {int(tree['block_size'])}, the state uses {a.block_size}")
    lion = state.opt_state[1] if isinstance(state.opt_state, tuple) and len(state.opt_state) == 2 else state.opt_state

    def flat_state(t, prefix=""):
        out = {}
        for k, v in t.items():
            name = f"{prefix}/{k}" if prefix else k
            if isinstance(v, dict) and set(v) != {"0", "1"}:
                out.update(flat_state(v, name))
            else:
                out[name] = v
        return out

    have = flat_state(tree["mu_quant"])
    from .lion_quant import _flatten_state
    mine = _flatten_state(lion.lion.mu_quant)
    if set(have) != set(mine):
        raise ValueError("optimizer state does not match the parameter tree")
    for n, m in mine.items():
        if isinstance(m, tuple):
            m[0].copy_(torch.from_numpy(np.ascontiguousarray(have[n]["0"])))
            m[1].copy_(torch.from_numpy(np.ascontiguousarray(have[n]["1"])))
        else:
            m.copy_(torch.from_numpy(np.ascontiguousarray(have[n])))
    state.step.copy_(torch.as_tensor(np.asarray(tree["step"])))
    lion.lion.count.copy_(torch.as_tensor(np.asarray(tree["count"])))
    lion.schedule.count.copy_(torch.as_tensor(np.asarray(tree["schedule_count"])))
    state._host_step = int(np.asarray(tree["schedule_count"]))
