"""The denoising train step behind the reference's call surface (/root/reference/training_utils.py).

  TrainState               flax.training.train_state.TrainState as built at training_utils.py:383-387 (step, apply_fn,
                           params, tx, opt_state; .create(), .apply_gradients())
  FrozenModel              training_utils.py:40-49 ((call, params) pair: the frozen noise scheduler)
  create_mask              training_utils.py:116-131
  create_lion_optimizer    the optax chain of training_utils.py:355-382: clip_by_global_norm(1) -> lion_8bit(lr/7, .9,
                           .99, wd*7, block 16, masks)
  train_step(state, batch, rng) -> (state, metrics)
                           training_utils.py:504-762 restricted to the UNet (north-star surface): key split + noise /
                           timestep sampling with jax's Threefry generator (530, 590-624; offset / perturbation noise),
                           add_noise, UNet forward, epsilon / v target, MSE (+ min-SNR-gamma, weights computed on the
                           device), backward, [data-parallel gradient mean], clip + 8-bit Lion + apply (+ EMA of the
                           new parameters fused into the same launches, 537-544 / 735-739), step + 1. The state and the
                           EMA tree are DONATED (updated in place), as the reference's jit does
                           (training_utils.py:846-851); the loss stays on the device; `return_rng=True` appends
                           `new_train_rng` to the result like the reference's tuple (755-762).

Parameters, gradients, bf16 GEMM operands and optimizer state live in flat arenas so that the optimizer is three kernel
launches and the gradient all-reduce runs over contiguous buckets; `state.params` / `state.opt_state` are zero-copy
views with the reference's tree structure and the (int8 codes, fp32 reciprocal-absmax) leaf layout.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib, ops
from . import jax_random as JR
from . import lion_quant as LQ
from .lion_quant import create_mask, flatten_dict, unflatten_dict
from .schedulers import FlaxDDPMScheduler
from . import clip as CL
from .unet import SD15, UNet, UNetConfig, init_params_numpy, layer_program, param_shapes


@dataclass
class FrozenModel:
    """mutable wrapper around (call, params) (training_utils.py:40-49)."""
    call: Any
    params: Any


@dataclass
class TrainingConfig:
    """The hyper-parameters of training_utils.py:52-113 that reach the UNet train step (defaults of
    model_properties_example.json / on_device_model_training_state, training_utils.py:430-442)."""
    u_net_learning_rate: float = 1e-6
    adam_to_lion_scale_factor: int = 7
    lion_8bit_block_size: int = 16
    quantize_unet_state: bool = True
    excluded_layer_pattern_from_weight_decay: Tuple[str, ...] = ("bias", "scale", "embedding")
    excluded_layer_from_quantization: Tuple[str, ...] = ("bias", "scale", "embedding", "conv_in", "conv_out",
                                                         "time_embedding", "embeddings", "time_emb_proj")
    beta_schedule: str = "scaled_linear"
    prediction_type: str = "epsilon"
    offset_noise_magnitude: float = 0.0
    min_snr_gamma_magnitude: float = 0.0
    perturbation_noise_magnitude: float = 0.0
    accumulate_unet_ema: bool = False   # model_properties_example.json:38
    ema_rate: float = 0.0               # model_properties_example.json:40 uses 0.99998


def create_lion_optimizer(params: dict, cfg: TrainingConfig = TrainingConfig()) -> LQ.GradientTransformation:
    """training_utils.py:281-382 for the UNet: masks by path component, lr / 7, weight decay 1e-2 * 7, clip 1."""
    wd_mask = create_mask(params, list(cfg.excluded_layer_pattern_from_weight_decay))
    if cfg.quantize_unet_state:
        q_mask = create_mask(params, list(cfg.excluded_layer_from_quantization))
    else:
        q_mask = LQ.tree_map(lambda _p: False, params)
    lion = LQ.lion_8bit(learning_rate=LQ.constant_schedule(cfg.u_net_learning_rate / cfg.adam_to_lion_scale_factor),
                        b1=0.9, b2=0.99, weight_decay=1e-2 * cfg.adam_to_lion_scale_factor, mask=wd_mask,
                        block_size=cfg.lion_8bit_block_size, excluded_layer_mask=q_mask)
    return LQ.chain(LQ.clip_by_global_norm(1), lion)


def create_lion_optimizer_states(models: dict, train_unet: bool = True, train_text_encoder: bool = True,
                                 adam_to_lion_scale_factor: int = 7, u_net_learning_rate: float = 1e-6,
                                 text_encoder_learning_rate: float = 1e-6,
                                 excluded_layer_pattern_from_weight_decay: Sequence[str] = (),
                                 excluded_layer_from_quantization: Sequence[str] = (),
                                 lion_8bit_block_size: Optional[int] = None, quantize_unet_state: bool = False,
                                 quantize_text_encoder_state: bool = False, device="cuda") -> dict:
    """create_lion_optimizer_states (training_utils.py:281-428), same arguments and result:
    models = {"unet": {"unet_model", "unet_params"}, "text_encoder": {"text_encoder_model", "text_encoder_params"}} ->
    {"unet_state": TrainState | None, "text_encoder_state": TrainState | None}, each over
    clip_by_global_norm(1) -> Lion(lr / scale, b1 .9, b2 .99, weight_decay 1e-2 * scale, decay mask by path component),
    8-bit block-quantised momentum for the leaves the quantisation list does not exclude when quantize_*_state is set
    (else fp32 momentum everywhere: the reference's optax.lion branch). `device`: where the states live (the reference
    builds them on the host and device_puts them later, training_utils.py:446-460)."""
    out = {"unet_state": None, "text_encoder_state": None}

    def build(params, lr, quantize):
        wd_mask = create_mask(params, list(excluded_layer_pattern_from_weight_decay)) \
            if excluded_layer_pattern_from_weight_decay else None
        q_mask = create_mask(params, list(excluded_layer_from_quantization)) \
            if (quantize and lion_8bit_block_size is not None) else LQ.tree_map(lambda _p: False, params)
        lion = LQ.lion_8bit(learning_rate=LQ.constant_schedule(lr / adam_to_lion_scale_factor), b1=0.9, b2=0.99,
                            weight_decay=1e-2 * adam_to_lion_scale_factor, mask=wd_mask,
                            block_size=lion_8bit_block_size or 16, excluded_layer_mask=q_mask)
        return LQ.chain(LQ.clip_by_global_norm(1), lion)

    if train_unet:
        p = models["unet"]["unet_params"]
        out["unet_state"] = TrainState.create(apply_fn=models["unet"]["unet_model"].apply, params=p,
                                              tx=build(p, u_net_learning_rate, quantize_unet_state), device=device)
    if train_text_encoder:
        p = models["text_encoder"]["text_encoder_params"]
        out["text_encoder_state"] = TrainState.create(
            apply_fn=models["text_encoder"]["text_encoder_model"].__call__, params=p,
            tx=build(p, text_encoder_learning_rate, quantize_text_encoder_state), device=device)
    return out


# ---------------------------------------------------------------------------------------------------------------
class ParamArena:
    """Flat device storage of one model's parameters and optimizer state.

    Leaves are grouped in four segments by (quantised state?, weight decay?) and, inside a segment, ordered as the
    backward pass produces their gradients, so that gradient buckets complete front to back. Every leaf starts on a
    boundary of max(16, block_size) elements (quantisation blocks never straddle leaves); padding stays zero forever
    (zero gradient -> zero update)."""

    def __init__(self, shapes: Dict[str, tuple], quant: Dict[str, bool], decay: Dict[str, bool], order: List[str],
                 device, block_size: int = 16):
        self.device = torch.device(device)
        self.block_size = block_size
        self.ALIGN = max(16, int(block_size))
        self.shapes = dict(shapes)
        rank = {n: i for i, n in enumerate(order)}
        names = sorted(shapes, key=lambda n: rank.get(n, len(rank)))
        self.segments = []  # (quant, decay, start, end)
        self.offset: Dict[str, int] = {}
        off = 0
        for q in (True, False):
            for wd in (True, False):
                start = off
                for n in names:
                    if bool(quant[n]) == q and bool(decay[n]) == wd:
                        numel = int(np.prod(shapes[n]))
                        if q and numel % block_size != 0:
                            raise ValueError(f"cannot reshape array of size {numel} into shape (-1, {block_size})")
                        self.offset[n] = off
                        off += -(-numel // self.ALIGN) * self.ALIGN
                if off > start:
                    self.segments.append((q, wd, start, off))
        self.n_total = off
        self.n_quant = sum(e - s for q, _, s, e in self.segments if q)
        dev = self.device
        self.master = torch.zeros(off, dtype=torch.float32, device=dev)
        self.grad = torch.zeros(off, dtype=torch.float32, device=dev)
        self.shadow = torch.zeros(off, dtype=torch.bfloat16, device=dev)
        # quantised segments come first, so codes/scales index the arena prefix directly
        self.codes = torch.full((max(self.n_quant, 1),), 3, dtype=torch.int8, device=dev)
        self.inv_scales = torch.ones((max(self.n_quant // block_size, 1),), dtype=torch.float32, device=dev)
        self.mu = torch.zeros((max(off - self.n_quant, 1),), dtype=torch.float32, device=dev)
        self.quant = dict(quant)
        self.decay = dict(decay)
        self.p = {n: self._view(self.master, n) for n in shapes}
        self.g = {n: self._view(self.grad, n) for n in shapes}
        self.w = {n: self._view(self.shadow, n) for n in shapes}
        l = _lib.lib()
        self.norm_ws = torch.empty(l.sdb_sumsq_workspace_bytes(), dtype=torch.uint8, device=dev)
        self.gnorm = torch.zeros(1, dtype=torch.float32, device=dev)
        self.ema: Optional[torch.Tensor] = None   # flat fp32 EMA of the masters (create_ema)
        self.e: Dict[str, torch.Tensor] = {}

    def create_ema(self) -> Dict[str, torch.Tensor]:
        """EMA copy of the parameters (training_utils.py:466-483 copies the loaded params) as one flat buffer with the
        arena's layout, so the moving average rides in the optimizer launches."""
        if self.ema is None:
            self.ema = self.master.clone()
            self.e = {n: self._view(self.ema, n) for n in self.shapes}
        return self.e

    def _view(self, flat, name):
        o = self.offset[name]
        shape = self.shapes[name]
        return flat[o:o + int(np.prod(shape))].view(shape)

    def load(self, flat_params: Dict[str, Any]) -> None:
        for n, v in flat_params.items():
            t = torch.as_tensor(np.asarray(v) if not isinstance(v, torch.Tensor) else v)
            self.p[n].copy_(t.to(torch.float32).reshape(self.shapes[n]))
        self.refresh_shadow()

    def refresh_shadow(self) -> None:
        ops.cast_bf16(self.master, out=self.shadow)

    def mu_quant_tree(self) -> dict:
        """state.mu_quant with the reference's leaf layout (lion_quant.py:66-80, 115-131), as views."""
        bs = self.block_size
        out = {}
        for n, shape in self.shapes.items():
            o, numel = self.offset[n], int(np.prod(shape))
            if self.quant[n]:
                out[n] = (self.codes[o:o + numel].view(numel // bs, bs), self.inv_scales[o // bs:(o + numel) // bs].view(-1, 1))
            else:
                out[n] = self.mu[o - self.n_quant:o - self.n_quant + numel].view(shape)
        return unflatten_dict(out)

    def optimizer_step(self, *, lr: float, b1: float, b2: float, wd: float, max_norm: Optional[float],
                       ema_rate: float = 0.0) -> None:
        """clip_by_global_norm -> lion_8bit -> apply_updates over the whole arena (+ bf16 operand refresh, + the EMA of
        the new parameters when `ema_rate` is set and create_ema() was called)."""
        l = _lib.lib()
        st = _lib.stream_ptr(self.device)
        ema0 = self.ema.data_ptr() if (ema_rate and self.ema is not None) else 0
        gn = 0
        if max_norm is not None:
            _lib.check(l.sdb_global_norm(self.grad.data_ptr(), self.n_total, self.norm_ws.data_ptr(),
                                         self.norm_ws.numel(), self.gnorm.data_ptr(), 0, st), "sdb_global_norm")
            gn = self.gnorm.data_ptr()
        es4, es2 = 4, 2
        for q, wdf, s, e in self.segments:
            n = e - s
            g, p, sh = self.grad.data_ptr() + s * es4, self.master.data_ptr() + s * es4, self.shadow.data_ptr() + s * es2
            em = ema0 + s * es4 if ema0 else 0
            if q:
                _lib.check(l.sdb_lion8_update(g, p, p, sh, self.codes.data_ptr() + s,
                                              self.inv_scales.data_ptr() + (s // self.block_size) * 4, n,
                                              self.block_size, b1, b2, lr, wd, int(wdf), gn,
                                              1.0 if max_norm is None else max_norm, 0, em, float(ema_rate), st),
                           "sdb_lion8_update")
            else:
                _lib.check(l.sdb_lion_fp32_update(g, p, p, sh, self.mu.data_ptr() + (s - self.n_quant) * 4, n, b1, b2,
                                                  lr, wd, int(wdf), gn, 1.0 if max_norm is None else max_norm, 0,
                                                  em, float(ema_rate), st), "sdb_lion_fp32_update")


def backward_param_order(cfg: UNetConfig) -> List[str]:
    """Parameter names in the order the backward pass finishes their gradients (reverse forward order)."""
    shapes = param_shapes(cfg)
    fwd: List[str] = [n for n in shapes if n.startswith("time_embedding/")]
    for op in layer_program(cfg):
        if op[0] in ("conv_in", "down", "up", "resnet", "xformer"):
            fwd += [n for n in shapes if n.startswith(op[1] + "/")]
        elif op[0] == "out":
            fwd += [n for n in shapes if n.startswith(op[1] + "/") or n.startswith(op[2] + "/")]
    seen, uniq = set(), []
    for n in fwd:
        if n not in seen:
            seen.add(n)
            uniq.append(n)
    assert len(uniq) == len(shapes)
    # the time-embedding MLP is used by every ResNet, so its gradient completes last
    return list(reversed(uniq))


class UNetModel:
    """Stand-in for the FlaxUNet2DConditionModel module object: `.config` and `.apply` (forward only)."""

    def __init__(self, config: UNetConfig = SD15):
        self.config = config
        self._state: Optional["TrainState"] = None

    def init_weights(self, seed: int = 0) -> dict:
        return unflatten_dict(init_params_numpy(self.config, seed))

    def apply(self, variables, sample, timesteps, encoder_hidden_states, train: bool = False):
        """unet.apply({"params": p}, sample NCHW, timesteps, encoder_hidden_states).sample -> NCHW fp32."""
        st = self._state
        if st is None or variables.get("params") is not st.params:
            raise ValueError("UNetModel.apply: pass {'params': state.params} of the TrainState created with this model")
        B, Cc, H, W = sample.shape
        x = ops.nchw_to_nhwc_bf16(sample.to(torch.float32).contiguous())
        ctx = ops.cast_bf16(encoder_hidden_states.to(torch.float32).contiguous())
        pred = st.net.forward(x, timesteps.to(torch.int32), ctx, forward_only=True)
        st.net.graph = None
        Co = pred.v.shape[-1]
        out = torch.empty((B, Co, H, W), dtype=torch.float32, device=pred.v.device)
        _lib.check(_lib.lib().sdb_nhwc_bf16_to_nchw_f32(pred.v.data_ptr(), out.data_ptr(), B, Co, H * W,
                                                        _lib.stream_ptr(out.device)), "sdb_nhwc_bf16_to_nchw_f32")
        from types import SimpleNamespace
        return SimpleNamespace(sample=out)


class TrainState:
    """flax.training.train_state.TrainState over a ParamArena (training_utils.py:383-387)."""

    def __init__(self, step, apply_fn, params, tx, opt_state, arena: ParamArena, net: UNet, hyper: dict):
        self.step = step
        self.apply_fn = apply_fn
        self.params = params
        self.tx = tx
        self.opt_state = opt_state
        self.arena = arena
        self.net = net
        self.hyper = hyper
        self.dp = None  # data-parallel reducer (parallel.GradReducer), attached by parallel.attach()
        self._host_step = 0
        self.ema_params: Optional[dict] = None

    def create_ema(self) -> dict:
        """unet_ema_params: a copy of the parameters to be averaged (training_utils.py:466-483), laid out like the
        arena so that train_step(..., unet_ema_params=<this tree>, ema_rate=r) fuses the average into the optimizer
        launches. Returns the tree (views)."""
        if self.ema_params is None:
            self.ema_params = unflatten_dict(self.arena.create_ema())
        return self.ema_params

    @classmethod
    def create(cls, *, apply_fn, params: dict, tx: LQ.GradientTransformation, device="cuda",
               reference_attention_quirk: bool = True) -> "TrainState":
        model = getattr(apply_fn, "__self__", None)
        is_text = isinstance(model, CL.CLIPTextModel)
        if not isinstance(model, UNetModel) and not is_text:
            raise TypeError("TrainState.create: apply_fn must be UNetModel(...).apply or CLIPTextModel(...).__call__")
        cfg = model.config
        flat = flatten_dict(params)
        shapes = CL.param_shapes(cfg) if is_text else param_shapes(cfg)
        if set(flat) != set(shapes):
            missing = sorted(set(shapes) - set(flat))[:3]
            extra = sorted(set(flat) - set(shapes))[:3]
            raise ValueError(f"params do not match the model config (missing {missing}, unexpected {extra})")
        hyper = _parse_chain(tx.hyper)
        ones = {n: True for n in shapes}
        quant = flatten_dict(LQ._expand_mask(hyper["excluded_layer_mask"], params)) if hyper["quantize"] else \
            {n: False for n in shapes}
        decay = flatten_dict(LQ._expand_mask(hyper["mask"], params)) if hyper["mask"] is not None else ones
        order = CL.backward_param_order(cfg) if is_text else backward_param_order(cfg)
        arena = ParamArena(shapes, quant, decay, order, device, hyper["block_size"])
        arena.load(flat)
        tree = unflatten_dict(arena.p)
        net = CL.ClipText(cfg, arena.p, arena.w, arena.g) if is_text else \
            UNet(cfg, arena.p, arena.w, arena.g, reference_attention_quirk)
        dev = arena.device
        lion_state = LQ.ScaleBy8bitLionState(count=torch.zeros([], dtype=torch.int32, device=dev),
                                             mu_quant=arena.mu_quant_tree(),
                                             mu_quant_flag=hyper["excluded_layer_mask"])
        chain_state = LQ.LionChainState(lion_state, LQ.MaskedState(LQ.EmptyState()),
                                        LQ.ScaleByScheduleState(torch.zeros([], dtype=torch.int32, device=dev)))
        opt_state = (LQ.EmptyState(), chain_state) if hyper["max_norm"] is not None else chain_state
        st = cls(torch.zeros([], dtype=torch.int32, device=dev), apply_fn, tree, tx, opt_state, arena, net, hyper)
        model._state = st
        return st

    def apply_gradients(self, *, grads: Optional[dict] = None, ema_rate: float = 0.0) -> "TrainState":
        """tx.update + optax.apply_updates + step+1, in place (the state is donated). `grads=None` or the state's own
        gradient tree uses the gradients already in the arena (what train_step produces). `ema_rate` (with
        create_ema()) also updates the EMA tree from the new parameters in the same launches."""
        a = self.arena
        if grads is not None and grads is not getattr(self, "_grad_tree", None):
            for n, g in flatten_dict(grads).items():
                a.g[n].copy_(g)
        h = self.hyper
        lr = h["learning_rate"]
        lion = self.opt_state[1] if isinstance(self.opt_state, tuple) and not isinstance(self.opt_state, LQ.LionChainState) \
            else self.opt_state
        if callable(lr):
            lr = lr(self._host_step)  # host mirror of ScaleByScheduleState.count: no device sync per step
        self._host_step += 1
        a.optimizer_step(lr=float(lr), b1=h["b1"], b2=h["b2"], wd=h["weight_decay"], max_norm=h["max_norm"],
                         ema_rate=ema_rate)
        _lib.check(_lib.lib().sdb_bump_counters(self.step.data_ptr(), lion.lion.count.data_ptr(),
                                                lion.schedule.count.data_ptr(), _lib.stream_ptr(a.device)),
                   "sdb_bump_counters")
        return self

    @property
    def grads(self) -> dict:
        """The gradient tree of the last backward pass (views into the arena)."""
        if getattr(self, "_grad_tree", None) is None:
            self._grad_tree = unflatten_dict(self.arena.g)
        return self._grad_tree


def _parse_chain(h: dict) -> dict:
    parts = h["parts"] if h["kind"] == "chain" else [h]
    out = {"max_norm": None, "quantize": True}
    lion = None
    for p in parts:
        if p["kind"] == "clip":
            out["max_norm"] = p["max_norm"]
        elif p["kind"] == "lion8":
            lion = p
        else:
            raise NotImplementedError(f"TrainState: optimizer part {p['kind']!r} has no fused arena path")
    if lion is None:
        raise NotImplementedError("TrainState: the optimizer chain must contain lion_8bit")
    out.update(lion)
    if out["block_size"] not in (16, 32, 64):
        raise NotImplementedError("the optimizer kernels are built for block_size 16 (model_properties_example.json:41), "
                                  "32 and 64 (lion_8bit's default)")
    return out


# ---------------------------------------------------------------------------------------------------------------
_KEY = {}


def _dp_rows(state: "TrainState", B: int) -> Tuple[int, int]:
    """(global batch rows, first row of this rank): the reference draws noise / timesteps for the GLOBAL batch and
    shards it on dim 0 (training_utils.py:835-840); rank r owns rows [r*B, (r+1)*B)."""
    dp = state.dp
    if dp is None or dp.world_size == 1:
        return B, 0
    return B * dp.world_size, B * dp.rank


def draw_noise_and_timesteps(state: "TrainState", shape, rng, num_train_timesteps: int, *, need_noise: bool,
                             need_timesteps: bool, offset_noise_magnitude: float = 0.0,
                             perturbation_noise_magnitude: float = 0.0, noise_out=None, timesteps_out=None):
    """training_utils.py:530, 590-624 on the device: returns (noise | None, timesteps | None, new_train_rng).
    `rng` is the step's `train_rng` (a key, an int seed or a jax_random.RngStream, which is advanced in place)."""
    keys = JR.train_step_keys(rng)
    dev = state.arena.device
    B = int(shape[0])
    Bg, b0 = _dp_rows(state, B)
    noise = timesteps = None
    if need_noise:
        noise = JR.sample_noise(keys["noise"], shape, dev, offset_key=keys["offset"],
                                offset_noise_magnitude=offset_noise_magnitude, perturb_key=keys["perturb"],
                                perturbation_noise_magnitude=perturbation_noise_magnitude, batch_global=Bg,
                                batch_begin=b0, out=noise_out)
    if need_timesteps:
        timesteps = JR.randint(keys["timestep"], B, 0, num_train_timesteps, dev, n_global=Bg, begin=b0,
                               out=timesteps_out)
    if isinstance(rng, JR.RngStream):
        rng.key = keys["new_train"]
    return noise, timesteps, keys["new_train"]


def _check_rng(rng, return_rng: bool, what: str) -> None:
    if rng is None:
        raise ValueError(f"{what}: the batch has no 'noise' / 'timesteps', so a train_rng is needed (a jax_random key, "
                         "an int seed or a jax_random.RngStream)")
    if not isinstance(rng, JR.RngStream) and not return_rng:
        raise ValueError(f"{what}: sampling from a raw key without return_rng=True would draw the same noise and "
                         "timesteps every step; pass return_rng=True and feed new_train_rng back "
                         "(training_utils.py:530, 755-762) or pass a jax_random.RngStream")


_CTX_INDEX: Dict[tuple, torch.Tensor] = {}


def _context_index(device, batch: int, chunks: int, strip: bool) -> torch.Tensor:
    key = (str(device), batch, chunks, strip)
    if key not in _CTX_INDEX:
        _CTX_INDEX[key] = torch.from_numpy(CL.context_row_index(batch, chunks, strip).reshape(-1).copy()).to(device)
    return _CTX_INDEX[key]


def encode_pixels(state: TrainState, frozen_vae_state: FrozenModel, pixel_values: torch.Tensor, rng, return_rng: bool,
                  what: str) -> torch.Tensor:
    """training_utils.py:574-586: frozen VAE encode, `latent_dist.sample(sample_rng)`, NCHW, * 0.18215. The sample key is
    the step's `sample_rng` (split(train_rng, 3)[1]) — the same key the noise keys are split from, as in the reference."""
    _check_rng(rng, return_rng, what)
    dev = state.arena.device
    enc = frozen_vae_state.call.encoder(frozen_vae_state.params, dev)
    B, _, H, W = pixel_values.shape
    mom = enc.encode(pixel_values)
    Bg, b0 = _dp_rows(state, B)
    down = 2 ** (len(enc.cfg.block_out_channels) - 1)
    return enc.sample(mom, (H // down, W // down), key=JR.train_step_keys(rng)["sample"], batch_global=Bg, batch_begin=b0)


def train_step(state: TrainState, batch: dict, rng=None, frozen_noise_scheduler_state: Optional[FrozenModel] = None, *,
               frozen_vae_state: Optional[FrozenModel] = None,
               text_encoder_state: Optional[TrainState] = None, strip_bos_eos_token: bool = True,
               unet_ema_params: Optional[dict] = None, text_encoder_ema_params: Optional[dict] = None,
               offset_noise_magnitude: float = 0.0, min_snr_gamma_magnitude: float = 0.0,
               perturbation_noise_magnitude: float = 0.0, ema_rate: float = 0.0, return_rng: bool = False):
    """One denoising train step (training_utils.py:504-762, UNet part). batch: {"latents" [B,4,H,W],
    "encoder_hidden_states" [B,L,D], optional "noise" [B,4,H,W], "timesteps" int [B]} (device tensors; floating
    inputs of any width are promoted to fp32 like jnp would). Whatever the batch does not supply is drawn from `rng`
    exactly as the reference draws it (split(train_rng, 3) / split(sample_rng, 4), jax.random.normal / randint; offset
    and perturbation noise apply to drawn noise only, as in the reference).
    With `frozen_vae_state` (FrozenModel(call=vae.AutoencoderKL, params=tree), training_utils.py:218-222) and
    "pixel_values" fp32 [B, 3, H, W] in the batch instead of "latents", the frozen VAE encodes the pixels and the latents
    are sampled from its posterior with the step's `sample_rng`, scaled by 0.18215 (574-586).
    With `text_encoder_state` (a TrainState over clip.CLIPTextModel, training_utils.py:389-425) the batch carries
    "input_ids" int [B * chunks, 77] instead of "encoder_hidden_states": the text encoder runs inside the step, its
    hidden states are assembled into the context as the reference does (641-674, `strip_bos_eos_token`), the UNet's
    context gradient flows back through it and both optimizers step (732-733); that state is donated too.
    Returns (state, {"loss": 0-d device tensor}) and, with return_rng=True, new_train_rng as third element. `state`
    and `unet_ema_params` (state.create_ema(); updated when ema_rate != 0, training_utils.py:735-739) are donated;
    `text_encoder_ema_params` (text_encoder_state.create_ema(), training_utils.py:740-746) likewise.
    Raises ValueError for an unknown prediction type (training_utils.py:697-701)."""
    if frozen_noise_scheduler_state is None:
        frozen_noise_scheduler_state = default_scheduler(state.arena.device)
    sched: FlaxDDPMScheduler = frozen_noise_scheduler_state.call
    sstate = frozen_noise_scheduler_state.params
    ptype = sched.config.prediction_type
    if ptype not in ("epsilon", "v_prediction"):
        raise ValueError(f"Unknown prediction type {ptype}")
    dev = state.arena.device
    if torch.cuda.current_device() != dev.index:
        raise RuntimeError(f"train_step: the state lives on {dev} but the current CUDA device is "
                           f"cuda:{torch.cuda.current_device()} (call torch.cuda.set_device first)")
    used_rng = False
    if "latents" not in batch and frozen_vae_state is not None and "pixel_values" in batch:
        lat = encode_pixels(state, frozen_vae_state, batch["pixel_values"], rng, return_rng, "train_step")
        used_rng = True
    else:
        lat = batch["latents"]
    if lat.device != dev:
        raise ValueError(f"train_step: batch['latents'] is on {lat.device}, the state on {dev}")
    if lat.dtype != torch.float32 or not lat.is_contiguous():
        lat = lat.to(torch.float32).contiguous()
    B = lat.shape[0]
    noise = batch.get("noise")
    timesteps = batch.get("timesteps")
    new_rng = None
    if noise is None or timesteps is None:
        _check_rng(rng, return_rng, "train_step")
        n_, t_, new_rng = draw_noise_and_timesteps(
            state, lat.shape, rng, sched.config.num_train_timesteps, need_noise=noise is None,
            need_timesteps=timesteps is None, offset_noise_magnitude=offset_noise_magnitude,
            perturbation_noise_magnitude=perturbation_noise_magnitude)
        noise = n_ if noise is None else noise
        timesteps = t_ if timesteps is None else timesteps
    elif (return_rng or used_rng) and rng is not None:
        new_rng = JR.train_step_keys(rng)["new_train"]  # the key still advances (training_utils.py:530)
        if isinstance(rng, JR.RngStream):
            rng.key = new_rng
    if noise.dtype != torch.float32 or not noise.is_contiguous():
        noise = noise.to(torch.float32).contiguous()
    timesteps = timesteps.to(torch.int32)

    noisy = sched.add_noise(sstate, lat, noise, timesteps)
    target = noise if ptype == "epsilon" else sched.get_velocity(sstate, lat, noise, timesteps)
    weight = None
    if min_snr_gamma_magnitude:
        weight = min_snr_weight(sstate.common.alphas_cumprod, timesteps, float(min_snr_gamma_magnitude),
                                ptype == "v_prediction")
    x = ops.nchw_to_nhwc_bf16(noisy)
    tnet = None
    if text_encoder_state is not None:
        if not isinstance(text_encoder_state.net, CL.ClipText):
            raise TypeError("train_step: text_encoder_state must be a TrainState over clip.CLIPTextModel")
        ids = batch["input_ids"]
        if ids.device != dev:
            raise ValueError(f"train_step: batch['input_ids'] is on {ids.device}, the state on {dev}")
        ids = ids.to(torch.int32).contiguous().view(-1, ids.shape[-1])
        if ids.shape[0] % B != 0 or ids.shape[1] != 77:
            raise ValueError(f"train_step: input_ids {tuple(ids.shape)} must be [B * chunks, 77] with B = {B}")
        tnet = text_encoder_state.net
        hidden = tnet.forward(ids)                                            # training_utils.py:635-640
        idx = _context_index(dev, B, ids.shape[0] // B, bool(strip_bos_eos_token))
        ctx_t = tnet.graph.gather_rows(hidden, idx)                            # training_utils.py:641-674
        ctx = (ctx_t, B)
    else:
        ctx = batch["encoder_hidden_states"]
        if ctx.dtype != torch.bfloat16:
            ctx = ops.cast_bf16(ctx.to(torch.float32).contiguous())

    net = state.net
    on_grad = state.dp.on_param_grad if state.dp is not None else None
    pred = net.forward(x, timesteps, ctx)
    loss, dpred = mse_loss_and_grad(pred.v, target, weight, 1.0 / (state.dp.world_size if state.dp else 1))
    if state.dp is not None:
        state.dp.begin(net.graph)
        state.dp.before_issue = net.graph.join
    net.backward(pred, dpred, on_grad)
    if state.dp is not None:
        loss = state.dp.finish(loss)
    if tnet is not None:  # the text encoder's backward starts from the context gradient the UNet left in ctx_t.g
        tdp = text_encoder_state.dp
        if tdp is not None:
            tdp.begin(tnet.graph)
            tdp.before_issue = tnet.graph.join
        tnet.backward(tdp.on_param_grad if tdp is not None else None)
        if tdp is not None:
            tdp.finish(None)
        t_ema = bool(ema_rate) and text_encoder_ema_params is not None
        t_fused = t_ema and text_encoder_ema_params is text_encoder_state.ema_params
        text_encoder_state.apply_gradients(grads=None, ema_rate=float(ema_rate) if t_fused else 0.0)
        if t_ema and not t_fused:
            LQ.compute_model_ema(text_encoder_ema_params, text_encoder_state.params, float(ema_rate))
    fused_ema = bool(ema_rate) and unet_ema_params is not None and unet_ema_params is state.ema_params
    state.apply_gradients(grads=None, ema_rate=float(ema_rate) if fused_ema else 0.0)
    if ema_rate and unet_ema_params is not None and not fused_ema:
        # a foreign EMA tree (not laid out like the arena): leaf-by-leaf compute_model_ema, training_utils.py:537-544
        LQ.compute_model_ema(unet_ema_params, state.params, float(ema_rate))
    metrics = {"loss": loss}
    return (state, metrics, new_rng) if return_rng else (state, metrics)


def min_snr_weight(alphas_cumprod: torch.Tensor, timesteps: torch.Tensor, gamma: float, v_pred: bool) -> torch.Tensor:
    """training_utils.py:546-568 — one weight per sample, computed on the device (no host round trip: the step stays
    asynchronous and CUDA-graph capturable)."""
    B = timesteps.numel()
    w = torch.empty((B,), dtype=torch.float32, device=alphas_cumprod.device)
    _lib.check(_lib.lib().sdb_min_snr_weight(alphas_cumprod.data_ptr(), alphas_cumprod.numel(), timesteps.data_ptr(), B,
                                             float(gamma), int(v_pred), w.data_ptr(),
                                             _lib.stream_ptr(w.device)), "sdb_min_snr_weight")
    return w


def mse_loss_and_grad(pred_nhwc: torch.Tensor, target_nchw: torch.Tensor, weight, grad_scale: float = 1.0):
    """loss = mean(w_b (target - pred)^2) (training_utils.py:704-709) and d loss / d pred * grad_scale (bf16 NHWC)."""
    l = _lib.lib()
    B, Cc, H, W = target_nchw.shape
    dpred = torch.empty_like(pred_nhwc)
    loss = torch.empty(1, dtype=torch.float32, device=pred_nhwc.device)
    ws = ops.workspace(pred_nhwc.device)
    _lib.check(l.sdb_mse_fwd_bwd(pred_nhwc.data_ptr(), target_nchw.data_ptr(), _lib.ptr(weight), dpred.data_ptr(),
                                 loss.data_ptr(), B, Cc, H * W, grad_scale, ws.data_ptr(), ws.numel(),
                                 _lib.stream_ptr()), "sdb_mse_fwd_bwd")
    return loss.view(()), dpred


def default_scheduler(device, beta_schedule: str = "scaled_linear", prediction_type: str = "epsilon") -> FrozenModel:
    """training_utils.py:223-230: FlaxDDPMScheduler(beta 0.00085 -> 0.012, 1000 steps)."""
    key = (str(device), beta_schedule, prediction_type)
    if key not in _KEY:
        s = FlaxDDPMScheduler(beta_start=0.00085, beta_end=0.012, beta_schedule=beta_schedule,
                              num_train_timesteps=1000, prediction_type=prediction_type, device=device)
        _KEY[key] = FrozenModel(call=s, params=s.create_state())
    return _KEY[key]


# ---------------------------------------------------------------------------------------------------------------
class CompiledTrainStep:
    """train_step captured once per batch shape into a CUDA graph and replayed — the counterpart of the reference's
    AOT `jax.jit(...).lower(...).compile()` per resolution bucket (training_utils.py:765-983: one executable per
    pixel shape, looked up by `batch["pixel_values"].shape` at training.py:214-234). A step is ~2 000 kernel launches;
    replaying them from a graph removes the per-launch host cost and keeps the GPU back-to-back.

    Per shape: the first call runs eagerly (fills the lazily built caches: workspaces, key-bias tables, function
    attributes), the second call captures and replays, later calls only copy the batch into the graph's static input
    buffers and replay. Noise / timesteps missing from the batch are drawn OUTSIDE the graph into those buffers (one
    or two Threefry kernels, training_utils.py:590-624); min-SNR weights and the EMA update are inside the graph.
    The loss returned in `metrics` is a static device tensor, overwritten by the next call (async, like the reference's
    futures: the host blocks only when it reads it)."""

    def __init__(self, state: TrainState, frozen_noise_scheduler_state: Optional[FrozenModel] = None, *,
                 frozen_vae_state: Optional[FrozenModel] = None,
                 text_encoder_state: Optional[TrainState] = None, strip_bos_eos_token: bool = True,
                 unet_ema_params: Optional[dict] = None, text_encoder_ema_params: Optional[dict] = None,
                 offset_noise_magnitude: float = 0.0, min_snr_gamma_magnitude: float = 0.0,
                 perturbation_noise_magnitude: float = 0.0, ema_rate: float = 0.0):
        self.vae = frozen_vae_state   # the frozen encode runs eagerly in front of the graph (its latents are an input)
        self.text_state = text_encoder_state
        self.strip = bool(strip_bos_eos_token)
        if text_encoder_state is not None:
            tlr = text_encoder_state.hyper["learning_rate"]
            if callable(tlr) and tlr(0) != tlr(1_000_000):
                raise NotImplementedError("CompiledTrainStep: only a constant learning-rate schedule can be baked in")
        lr = state.hyper["learning_rate"]
        if callable(lr) and lr(0) != lr(1_000_000):
            # the learning rate is a kernel argument baked into the graph; the reference only ever uses
            # optax.constant_schedule (training_utils.py:355-357)
            raise NotImplementedError("CompiledTrainStep: only a constant learning-rate schedule can be baked in")
        if ema_rate and unet_ema_params is not None and unet_ema_params is not state.ema_params:
            raise ValueError("CompiledTrainStep: unet_ema_params must be the tree returned by state.create_ema()")
        if ema_rate and text_encoder_ema_params is not None and (
                text_encoder_state is None or text_encoder_ema_params is not text_encoder_state.ema_params):
            raise ValueError("CompiledTrainStep: text_encoder_ema_params must be the tree returned by "
                             "text_encoder_state.create_ema()")
        self.text_ema_params = text_encoder_ema_params
        self.state = state
        self.sched = frozen_noise_scheduler_state or default_scheduler(state.arena.device)
        self.offset = float(offset_noise_magnitude)
        self.perturb = float(perturbation_noise_magnitude)
        self.min_snr = float(min_snr_gamma_magnitude)
        self.ema_params = unet_ema_params
        self.ema_rate = float(ema_rate)
        self._entries: Dict[tuple, dict] = {}
        self.launches_per_replay = 0

    def _flags(self) -> dict:
        return dict(unet_ema_params=self.ema_params, text_encoder_ema_params=self.text_ema_params,
                    ema_rate=self.ema_rate, min_snr_gamma_magnitude=self.min_snr,
                    text_encoder_state=self.text_state, strip_bos_eos_token=self.strip)

    def _with_latents(self, batch, rng, return_rng):
        if "latents" in batch or self.vae is None or "pixel_values" not in batch:
            return batch
        lat = encode_pixels(self.state, self.vae, batch["pixel_values"], rng, return_rng, "CompiledTrainStep")
        return dict({k: v for k, v in batch.items() if k != "pixel_values"}, latents=lat)

    def _eager(self, batch, rng, return_rng):
        return train_step(self.state, batch, rng, self.sched, offset_noise_magnitude=self.offset,
                          perturbation_noise_magnitude=self.perturb, return_rng=return_rng, **self._flags())

    def _fill(self, e, batch, rng, return_rng):
        """Copy / draw the step's inputs into the graph's static buffers (outside the graph). Returns new_train_rng."""
        buf = e["buf"]
        buf["latents"].copy_(batch["latents"], non_blocking=True)
        cond = "input_ids" if self.text_state is not None else "encoder_hidden_states"
        buf[cond].copy_(batch[cond], non_blocking=True)
        have_n, have_t = batch.get("noise") is not None, batch.get("timesteps") is not None
        if have_n:
            buf["noise"].copy_(batch["noise"], non_blocking=True)
        if have_t:
            buf["timesteps"].copy_(batch["timesteps"], non_blocking=True)
        new_rng = None
        if not (have_n and have_t):
            _check_rng(rng, return_rng, "CompiledTrainStep")
            _, _, new_rng = draw_noise_and_timesteps(
                self.state, buf["latents"].shape, rng, self.sched.call.config.num_train_timesteps,
                need_noise=not have_n, need_timesteps=not have_t, offset_noise_magnitude=self.offset,
                perturbation_noise_magnitude=self.perturb, noise_out=buf["noise"], timesteps_out=buf["timesteps"])
        elif return_rng and rng is not None:
            new_rng = JR.train_step_keys(rng)["new_train"]
            if isinstance(rng, JR.RngStream):
                rng.key = new_rng
        return new_rng

    def __call__(self, state: TrainState, batch: dict, rng=None, return_rng: bool = False):
        if state is not self.state:
            raise ValueError("CompiledTrainStep: called with a different TrainState than it was compiled for")
        batch = self._with_latents(batch, rng, return_rng)
        cond = "input_ids" if self.text_state is not None else "encoder_hidden_states"
        key = (tuple(batch["latents"].shape), tuple(batch[cond].shape))
        e = self._entries.get(key)
        if e is None:
            self._entries[key] = {"graph": None}
            return self._eager(batch, rng, return_rng)  # also the warm-up of every lazily built cache
        if e["graph"] is None:
            dev = state.arena.device
            e["buf"] = {"latents": torch.empty(key[0], dtype=torch.float32, device=dev),
                        "noise": torch.empty(key[0], dtype=torch.float32, device=dev),
                        "timesteps": torch.empty((key[0][0],), dtype=torch.int32, device=dev),
                        cond: torch.empty(key[1], dtype=torch.int32 if cond == "input_ids" else batch[cond].dtype,
                                          device=dev)}
            new_rng = self._fill(e, batch, rng, return_rng)
            ops.freeze_workspaces()  # the graphs bake the scratch pointers: growth would leave them dangling
            self._capture(e)
        else:
            new_rng = self._fill(e, batch, rng, return_rng)
        self._replay(e)
        state._host_step += 1
        return (state, e["metrics"], new_rng) if return_rng else (state, e["metrics"])

    def _capture(self, e) -> None:
        """Capture the step as ONE CUDA graph (single GPU, or data parallel with the direct NCCL communicator), else as
        a chain of CUDA-graph segments sharing one memory pool: with torch.distributed's collectives, NCCL stays OUT of
        the graphs: the reducer closes a segment whenever
        enough gradient buckets are complete, and at replay time those buckets are all-reduced eagerly on the
        communication stream right after the segment is enqueued, overlapping the following segments."""
        import gc
        state = self.state
        dev = state.arena.device
        dp = state.dp
        gc.collect()
        torch.cuda.synchronize(dev)
        pool = torch.cuda.graph_pool_handle()
        stream = torch.cuda.Stream(device=dev)
        stream.wait_stream(torch.cuda.current_stream(dev))
        segs = []   # (graph, bucket ids reduced after it, is_last_backward_segment)
        cur = {"g": None}
        n0 = _lib.lib().sdb_launch_count()
        host_step = state._host_step

        def begin():
            cur["g"] = torch.cuda.CUDAGraph()
            # thread_local: other threads (torch's NCCL watchdog polling events) must not invalidate the capture
            cur["g"].capture_begin(pool=pool, capture_error_mode="thread_local")

        def cut(final: bool):
            cur["g"].capture_end()
            segs.append((cur["g"], list(dp._deferred), final))
            dp._deferred.clear()
            begin()

        # With the direct NCCL communicator the collectives are captured like any kernel (forked onto the communication
        # stream by events, joined before the optimizer): ONE graph per step. Otherwise (SDB_DP_NCCL=torch) the step is
        # cut into segments and the collectives are issued eagerly between their replays.
        one_graph = dp is None or dp.world_size == 1 or (dp.graph_capturable and os.environ.get("SDB_DP_SEGMENTS") != "1")
        if self.text_state is not None and not one_graph:
            raise NotImplementedError("CompiledTrainStep: a trained text encoder under data parallelism needs the "
                                      "capturable collectives (peer-memory or direct NCCL), not graph segments")
        with torch.cuda.stream(stream):
            if dp is not None and not one_graph:
                dp.capture = cut
            try:
                begin()
                _, metrics = train_step(state, e["buf"], None, self.sched, **self._flags())
                cur["g"].capture_end()
                segs.append((cur["g"], [], False))
            finally:
                if dp is not None:
                    dp.capture = None
        torch.cuda.current_stream(dev).wait_stream(stream)
        state._host_step = host_step  # capture enqueued nothing: the replay that follows is the real step
        e["launches"] = _lib.lib().sdb_launch_count() - n0
        e["graph"], e["segs"], e["metrics"] = segs[0][0], segs, metrics
        self.launches_per_replay = e["launches"]

    def _replay(self, e) -> None:
        dp = self.state.dp
        for g, buckets, final in e["segs"]:
            g.replay()
            if dp is not None:
                if final:
                    dp.replay_finish(buckets, e["metrics"]["loss"])
                else:
                    dp.replay_segment_buckets(buckets)


def compile_train_step(state: TrainState, frozen_noise_scheduler_state: Optional[FrozenModel] = None, **flags
                       ) -> CompiledTrainStep:
    return CompiledTrainStep(state, frozen_noise_scheduler_state, **flags)
