"""ORACLE (test infrastructure only — never imported by the product path).

CPU restatement of the reference's validation sampler, FlaxStableDiffusionPipeline._generate
(/root/reference/models/pipeline_flax_stable_diffusion.py:160-254, vendored in the reference tree),
on top of the other oracles: text encoder (oracle/clip.py), UNet (oracle/unet.py), DDIM step (oracle/sched.py, pinned
by the reference's own scheduler file through tests/golden/ddim_ref.npz), Threefry normal for the initial latents
(oracle/threefry.py, pinned by Random123 / jax known answers) and the VAE decode (oracle/vae.py).

Line by line:
  172-186  prompt_embeds / negative_prompt_embeds = text_encoder(ids)[0]; context = concat([negative, prompt])
  189      guidance_scale as fp32
  191-201  latents = jax.random.normal(prng_seed, (B, C, H/8, W/8), fp32) unless given
  236-241  scheduler.set_timesteps; latents *= init_noise_sigma (1.0 for DDIM)
  203-234  loop body: latents_input = concat([latents] * 2); t = timesteps[step]; scale_model_input (identity for DDIM);
           noise_pred = unet(latents_input, t, context).sample; uncond, text = split(noise_pred, 2);
           noise_pred = uncond + guidance_scale * (text - uncond); latents = scheduler.step(noise_pred, t, latents)
  246-251  latents = 1 / scaling_factor * latents; image = vae.decode(latents).sample;
           image = (image / 2 + 0.5).clip(0, 1).transpose(0, 2, 3, 1)

mode "bf16" puts the roundings where the reference's dtypes put them (UNet / text encoder / VAE compute in bf16 and
return bf16; `text - uncond` is a bf16 subtraction; guidance and the scheduler are fp32; the image arithmetic is bf16);
"fp32" / "fp64" run the same arithmetic without the bf16 roundings (the yardstick's exact side).

PARITY: inherits the pins of its parts; the UNet / CLIP / VAE parts stay "unpinned by third-party code" as their
headers say.
"""
from __future__ import annotations

import numpy as np
import torch

from . import clip as OC
from . import sched as OS
from . import threefry as OT
from . import unet as OU
from . import vae as OV


def ddim_timesteps(num_inference_steps: int, num_train_timesteps: int = 1000, steps_offset: int = 0) -> np.ndarray:
    """scheduling_ddim_flax.py:163-184 (set_timesteps)."""
    ratio = num_train_timesteps // num_inference_steps
    return (np.arange(0, num_inference_steps) * ratio).round()[::-1].astype(np.int64) + steps_offset


def _rnd(x, mode):
    return x.to(torch.bfloat16).to(x.dtype) if mode == "bf16" else x


def generate(unet_cfg, clip_cfg, vae_cfg, params, prompt_ids, neg_prompt_ids, prng_key, num_inference_steps: int,
             height: int, width: int, guidance_scale: float, latents=None, mode: str = "fp32", sched_state=None,
             prediction_type: str = "epsilon", num_train_timesteps: int = 1000, return_latents: bool = False):
    """params: {"unet", "text_encoder", "vae"} flat dicts of torch tensors. Returns images NHWC [B, height, width, 3]
    (and the final latents NCHW fp32 when asked)."""
    if height % 8 != 0 or width % 8 != 0:
        raise ValueError(f"`height` and `width` have to be divisible by 8 but are {height} and {width}.")
    dt = torch.float64 if mode == "fp64" else torch.float32
    ids = torch.as_tensor(np.asarray(prompt_ids)).long()
    neg = torch.as_tensor(np.asarray(neg_prompt_ids)).long()
    B = ids.shape[0]
    pe = OC.text_encoder_forward(clip_cfg, params["text_encoder"], ids, mode=mode)
    ne = OC.text_encoder_forward(clip_cfg, params["text_encoder"], neg, mode=mode)
    context = _rnd(torch.cat([ne, pe]), mode)
    scale = 2 ** (len(vae_cfg.block_out_channels) - 1)
    shape = (B, unet_cfg.in_channels, height // scale, width // scale)
    if latents is None:
        latents = OT.normal(prng_key, shape)
    lat = np.asarray(latents, dtype=np.float32)
    if lat.shape != shape:
        raise ValueError(f"Unexpected latents shape, got {lat.shape}, expected {shape}")
    if sched_state is None:
        sched_state = OS.common_state("scaled_linear", 0.00085, 0.012, num_train_timesteps)
    gs = np.float32(guidance_scale)
    for t in ddim_timesteps(num_inference_steps, num_train_timesteps):
        x2 = torch.from_numpy(np.concatenate([lat, lat]))
        tt = torch.full((2 * B,), int(t), dtype=torch.int64)
        pred = _rnd(OU.unet_forward(unet_cfg, params["unet"], x2.to(dt), tt, context, mode=mode), mode)
        uncond, text = pred[:B], pred[B:]
        diff = _rnd(text - uncond, mode)
        if mode == "fp64":
            noise_pred = (uncond + float(gs) * diff).numpy()
            # the scheduler itself is fp32 in the reference; keep the UNet side exact and round once here
            noise_pred = noise_pred.astype(np.float32)
        else:
            noise_pred = (uncond.float() + torch.tensor(gs) * diff.float()).numpy()
        lat = OS.ddim_step(sched_state, noise_pred, int(t), lat, num_inference_steps, prediction_type=prediction_type,
                           num_train_timesteps=num_train_timesteps)
    z = (np.float32(1.0 / vae_cfg.scaling_factor) * lat).astype(np.float32)
    image = _rnd(OV.decode(vae_cfg, params["vae"], torch.from_numpy(z), mode=mode), mode)
    image = _rnd(_rnd(image / 2, mode) + 0.5, mode).clamp(0, 1).permute(0, 2, 3, 1)
    if return_latents:
        return image, lat
    return image
