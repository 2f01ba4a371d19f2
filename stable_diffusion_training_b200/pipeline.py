"""Validation sampler on the sm_100a kernels (host side) — SURVEY.md 8(f)-4.

Mirrors the reference's vendored FlaxStableDiffusionPipeline (/root/reference/models/pipeline_flax_stable_diffusion.py:
79-152 constructor / prepare_inputs, 160-254 `_generate`, 256-352 `__call__`; built by save_model at
training_utils.py:1006-1014): same constructor arguments, same `params` dict ({"text_encoder", "unet", "vae"[,
"scheduler"]}), same call arguments and outputs ({"images", "nsfw_content_detected"}).

What runs where:
  text encoder      clip.ClipText forward (uncond and prompt ids in one pass; rows are independent)
  initial latents   jax.random.normal(prng_seed, (B, 4, H/8, W/8)) — csrc/rng.cu, the draw jax makes
  loop body         sdb_latents_to_nhwc_bf16 (concatenate([latents] * 2) + entry transpose) -> unet.UNet forward at 2B
                    -> sdb_cfg_combine -> scheduler.step (sdb_ddim_step). `jit=True` (the reference's pmap/jit switch)
                    replays the first three as ONE CUDA graph per step: the forward is ~700 launches of a few
                    microseconds at sampling batch sizes, i.e. launch-bound when enqueued from Python.
  decode            vae.VAEDecoder.decode_image (1 / scaling_factor, decode, (x / 2 + 0.5).clip(0, 1), NHWC)

Executors are built once per parameter tree and cached; a tree that is the `.params` of a live TrainState reuses that
state's executor (the bf16 operand copies the optimizer keeps fresh), so validation between train steps copies nothing.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Union

import numpy as np
import torch

from . import _lib, ops
from . import clip as CL
from . import jax_random as JR
from . import unet as UN
from .lion_quant import flatten_dict

BF16 = torch.bfloat16


class _InferenceWeights:
    """fp32 copies (P) and bf16 GEMM operands (W) of one parameter tree, as two flat device buffers."""

    def __init__(self, shapes: Dict[str, tuple], params: dict, device, what: str):
        flat = flatten_dict(params)
        if set(flat) != set(shapes):
            missing = sorted(set(shapes) - set(flat))[:3]
            extra = sorted(set(flat) - set(shapes))[:3]
            raise ValueError(f"{what} params do not match the model config (missing {missing}, unexpected {extra})")
        off, offs = 0, {}
        for n, shp in shapes.items():
            offs[n] = off
            off += -(-int(np.prod(shp)) // 16) * 16   # 32-byte aligned bf16 leaves (TMA needs 16)
        self.master = torch.zeros(off, dtype=torch.float32, device=device)
        self.P, self.W = {}, {}
        for n, shp in shapes.items():
            v = flat[n]
            t = torch.as_tensor(np.asarray(v) if not isinstance(v, torch.Tensor) else v)
            if tuple(t.shape) != tuple(shp):
                raise ValueError(f"{what} leaf {n}: shape {tuple(t.shape)} != {tuple(shp)}")
            k = int(np.prod(shp))
            self.P[n] = self.master[offs[n]:offs[n] + k].view(shp)
            self.P[n].copy_(t.to(torch.float32))
        self.shadow = ops.cast_bf16(self.master)
        for n, shp in shapes.items():
            k = int(np.prod(shp))
            self.W[n] = self.shadow[offs[n]:offs[n] + k].view(shp)


class FlaxStableDiffusionPipeline:
    """pipeline_flax_stable_diffusion.py:79-152. `vae` / `text_encoder` / `unet` are the module stand-ins of this package
    (vae.AutoencoderKL, clip.CLIPTextModel, training_utils.UNetModel); `tokenizer` is any callable with the
    CLIPTokenizer call signature (only used by prepare_inputs and for the empty negative prompt); `scheduler` a
    schedulers.FlaxDDIMScheduler (the one save_model writes, training_utils.py:997-1004)."""

    def __init__(self, vae, text_encoder, tokenizer, unet, scheduler, dtype=torch.float32, device="cuda",
                 reference_attention_quirk: bool = True):
        self.vae, self.text_encoder, self.tokenizer, self.unet, self.scheduler = vae, text_encoder, tokenizer, unet, scheduler
        self.dtype = dtype
        self.device = torch.device(device)
        self.quirk = reference_attention_quirk
        self.vae_scale_factor = 2 ** (len(self.vae.config.block_out_channels) - 1)
        self._exec: dict = {}     # (kind, id(params)) -> (params, executor)
        self._graphs: dict = {}   # captured loop bodies
        self.launches_last_call = 0
        self._captured_launches = 0
        self.last_latents: Optional[torch.Tensor] = None   # final latents of the last call (diagnostics / tests)

    # ---- executors ------------------------------------------------------------------------------------------------
    def _executor(self, kind: str, params):
        key = (kind, id(params))
        hit = self._exec.get(key)
        if hit is not None and hit[0] is params:
            return hit[1]
        model = self.unet if kind == "unet" else self.text_encoder
        st = getattr(model, "_state", None)
        if st is not None and params is st.params and st.arena.device == self.device:
            ex = st.net
        elif kind == "unet":
            w = _InferenceWeights(UN.param_shapes(model.config), params, self.device, "unet")
            ex = UN.UNet(model.config, w.P, w.W, None, self.quirk)
            ex._weights = w
        else:
            w = _InferenceWeights(CL.param_shapes(model.config), params, self.device, "text_encoder")
            ex = CL.ClipText(model.config, w.P, w.W, None)
            ex._weights = w
        self._exec[key] = (params, ex)
        return ex

    # ---- reference API ----------------------------------------------------------------------------------------------
    def prepare_inputs(self, prompt: Union[str, List[str]]):
        if not isinstance(prompt, (str, list)):
            raise ValueError(f"`prompt` has to be of type `str` or `list` but is {type(prompt)}")
        text_input = self.tokenizer(prompt, padding="max_length", max_length=self.tokenizer.model_max_length,
                                    truncation=True, return_tensors="np")
        return text_input.input_ids

    def _ids(self, ids) -> torch.Tensor:
        t = torch.as_tensor(np.asarray(ids) if not isinstance(ids, torch.Tensor) else ids)
        if t.dim() != 2:
            raise ValueError(f"prompt ids must be [batch, tokens], got {tuple(t.shape)}")
        return t.to(device=self.device, dtype=torch.int32).contiguous()

    @staticmethod
    def _scalar_guidance(guidance_scale) -> float:
        g = np.asarray(guidance_scale.detach().cpu() if isinstance(guidance_scale, torch.Tensor) else guidance_scale,
                       dtype=np.float32).ravel()
        if g.size == 0 or not np.all(g == g[0]):
            # the reference multiplies a [1, B]-shaped array into [B, C, H, W] (pipeline :189, :227), which broadcasts
            # only for one sample per device: a single value per call is what its pmap path computes
            raise NotImplementedError("guidance_scale must be one value for the whole call")
        return float(g[0])

    def _loop_body(self, net, lat, tsteps, ctx, gs: float, out):
        """noise_pred (NCHW fp32, into `out`) for the current latents: pipeline :203-227."""
        B, Cc, h, w = lat.shape
        l = _lib.lib()
        s = _lib.stream_ptr(self.device)
        x2 = torch.empty((2 * B, h, w, Cc), device=self.device, dtype=BF16)
        _lib.check(l.sdb_latents_to_nhwc_bf16(lat.data_ptr(), 1.0, 2, x2.data_ptr(), B, Cc, h * w, s),
                   "sdb_latents_to_nhwc_bf16")
        pred = net.forward(x2, tsteps, ctx, forward_only=True)   # no tape: activations die behind the pass
        net.graph = None
        Co = pred.v.shape[-1]
        _lib.check(l.sdb_cfg_combine(pred.v.data_ptr(), gs, out.data_ptr(), B, Co, h * w, s), "sdb_cfg_combine")
        return out

    def _captured_body(self, net, lat, tsteps, ctx, gs: float, out):
        """The loop body as one CUDA graph over static buffers, captured once per (executor, shapes, guidance)."""
        key = (id(net), tuple(lat.shape), tuple(ctx.shape), gs)
        e = self._graphs.get(key)
        if e is None:
            ops.freeze_workspaces()
            e = {"lat": torch.empty_like(lat), "t": torch.empty_like(tsteps), "ctx": torch.empty_like(ctx),
                 "out": torch.empty_like(out), "net": net}
            torch.cuda.synchronize(self.device)
            n0 = _lib.lib().sdb_launch_count()
            g = torch.cuda.CUDAGraph()
            stream = torch.cuda.Stream(device=self.device)
            stream.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.graph(g, stream=stream, capture_error_mode="thread_local"):
                self._loop_body(net, e["lat"], e["t"], e["ctx"], gs, e["out"])
            e["graph"] = g
            e["launches"] = _lib.lib().sdb_launch_count() - n0
            self._captured_launches += e["launches"]   # enqueued into the graph, not onto the device
            self._graphs[key] = e
        e["lat"].copy_(lat)
        e["t"].copy_(tsteps)
        if e.get("ctx_src") is not ctx:
            e["ctx"].copy_(ctx)
            e["ctx_src"] = ctx
        e["graph"].replay()
        out.copy_(e["out"])
        return out, e["launches"]

    def _generate(self, prompt_ids, params: dict, prng_seed, num_inference_steps: int, height: int, width: int,
                  guidance_scale, latents=None, neg_prompt_ids=None, jit: bool = False):
        if height % 8 != 0 or width % 8 != 0:
            raise ValueError(f"`height` and `width` have to be divisible by 8 but are {height} and {width}.")
        dev = self.device
        n0 = _lib.lib().sdb_launch_count()
        self._captured_launches = 0
        ids = self._ids(prompt_ids)
        batch_size, max_length = ids.shape
        if neg_prompt_ids is None:
            uncond = self.tokenizer([""] * batch_size, padding="max_length", max_length=max_length,
                                    return_tensors="np").input_ids
        else:
            uncond = neg_prompt_ids
        uncond = self._ids(uncond)
        if tuple(uncond.shape) != tuple(ids.shape):
            raise ValueError(f"negative prompt ids {tuple(uncond.shape)} do not match prompt ids {tuple(ids.shape)}")
        text = self._executor("text_encoder", params["text_encoder"])
        hidden = text.forward(torch.cat([uncond, ids]), forward_only=True)   # concatenate([negative, prompt]) (:186)
        text.graph = None
        context = hidden.v.view(2 * batch_size, max_length, -1)
        gs = self._scalar_guidance(guidance_scale)

        cin = self.unet.config.in_channels
        latents_shape = (batch_size, cin, height // self.vae_scale_factor, width // self.vae_scale_factor)
        if latents is None:
            latents = JR.sample_noise(prng_seed, latents_shape, dev)
        else:
            latents = torch.as_tensor(np.asarray(latents) if not isinstance(latents, torch.Tensor) else latents)
            if tuple(latents.shape) != latents_shape:
                raise ValueError(f"Unexpected latents shape, got {tuple(latents.shape)}, expected {latents_shape}")
            latents = latents.to(device=dev, dtype=torch.float32).contiguous().clone()

        sched_state = params.get("scheduler")
        if sched_state is None:
            sched_state = self.scheduler.create_state()
        sched_state = self.scheduler.set_timesteps(sched_state, num_inference_steps=num_inference_steps,
                                                   shape=latents.shape)
        # scale the initial noise by the standard deviation required by the scheduler (:241)
        sigma = float(sched_state.init_noise_sigma)
        if sigma != 1.0:
            latents = latents * sigma
        net = self._executor("unet", params["unet"])
        timesteps = [int(t) for t in sched_state.timesteps.tolist()]
        tbuf = torch.empty((2 * batch_size,), dtype=torch.int32, device=dev)
        noise_pred = torch.empty(latents_shape, dtype=torch.float32, device=dev)
        replayed = 0
        for i, t in enumerate(timesteps[:num_inference_steps]):
            tbuf.fill_(t)
            x = self.scheduler.scale_model_input(sched_state, latents, t)
            if jit and i > 0:   # step 0 runs eagerly (lazy one-time setup must not happen under capture)
                _, n = self._captured_body(net, x, tbuf, context, gs, noise_pred)
                replayed += n
            else:
                self._loop_body(net, x, tbuf, context, gs, noise_pred)
            latents, sched_state = self.scheduler.step(sched_state, noise_pred, t, latents, return_dict=False)

        # scale and decode the image latents with vae (:246-250)
        dec = self.vae.decoder(params["vae"], dev)
        image = dec.decode_image(latents, 1.0 / float(self.vae.config.scaling_factor))
        self.launches_last_call = _lib.lib().sdb_launch_count() - n0 - self._captured_launches + replayed
        self.last_latents = latents
        return image

    def __call__(self, prompt_ids, params: dict, prng_seed, num_inference_steps: int = 50, height: Optional[int] = None,
                 width: Optional[int] = None, guidance_scale=7.5, latents=None, neg_prompt_ids=None,
                 return_dict: bool = True, jit: bool = False, output_type: str = "np"):
        """pipeline :256-352. `jit` selects the CUDA-graph loop body (the reference's pmap switch); `output_type="pt"`
        keeps the images on the device instead of the reference's np.asarray."""
        sample_size = getattr(self.unet.config, "sample_size", 64)
        height = height or sample_size * self.vae_scale_factor
        width = width or sample_size * self.vae_scale_factor
        images = self._generate(prompt_ids, params, prng_seed, num_inference_steps, height, width, guidance_scale,
                                latents, neg_prompt_ids, jit=jit)
        if output_type == "np":
            images = images.cpu().numpy()
        elif output_type != "pt":
            raise ValueError(f"output_type {output_type!r} must be 'np' or 'pt'")
        has_nsfw_concept = False
        if not return_dict:
            return (images, has_nsfw_concept)
        return {"images": images, "nsfw_content_detected": has_nsfw_concept}
