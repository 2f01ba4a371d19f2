"""The denoising train step behind the reference's call surface (/root/reference/training_utils.py).

  TrainState               flax.training.train_state.TrainState as built at training_utils.py:383-387 (step, apply_fn,
                           params, tx, opt_state; .create(), .apply_gradients())
  FrozenModel              training_utils.py:40-49 ((call, params) pair: the frozen noise scheduler)
  create_mask              training_utils.py:116-131
  create_lion_optimizer    the optax chain of training_utils.py:355-382: clip_by_global_norm(1) -> lion_8bit(lr/7, .9,
                           .99, wd*7, block 16, masks)
  train_step(state, batch, rng) -> (state, metrics)
                           training_utils.py:504-762 restricted to the UNet (north-star surface): add_noise, UNet
                           forward, epsilon / v target, MSE (+ min-SNR-gamma), backward, [data-parallel gradient mean],
                           clip + 8-bit Lion + apply, step + 1. The state is DONATED (updated in place), as the
                           reference's jit does (training_utils.py:846-851); the loss stays on the device.

Parameters, gradients, bf16 GEMM operands and optimizer state live in flat arenas so that the optimizer is three kernel
launches and the gradient all-reduce runs over contiguous buckets; `state.params` / `state.opt_state` are zero-copy
views with the reference's tree structure and the (int8 codes, fp32 reciprocal-absmax) leaf layout.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, List, Optional, Tuple

import numpy as np
import torch

from . import _lib, ops
from . import lion_quant as LQ
from .lion_quant import create_mask, flatten_dict, unflatten_dict
from .schedulers import FlaxDDPMScheduler
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


# ---------------------------------------------------------------------------------------------------------------
class ParamArena:
    """Flat device storage of one model's parameters and optimizer state.

    Leaves are grouped in four segments by (quantised state?, weight decay?) and, inside a segment, ordered as the
    backward pass produces their gradients, so that gradient buckets complete front to back. Every leaf starts on a
    16-element boundary; padding stays zero forever (zero gradient -> zero update)."""

    ALIGN = 16

    def __init__(self, shapes: Dict[str, tuple], quant: Dict[str, bool], decay: Dict[str, bool], order: List[str],
                 device, block_size: int = 16):
        self.device = torch.device(device)
        self.block_size = block_size
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

    def optimizer_step(self, *, lr: float, b1: float, b2: float, wd: float, max_norm: Optional[float]) -> None:
        """clip_by_global_norm -> lion_8bit -> apply_updates over the whole arena (+ bf16 operand refresh)."""
        l = _lib.lib()
        st = _lib.stream_ptr()
        gn = 0
        if max_norm is not None:
            _lib.check(l.sdb_global_norm(self.grad.data_ptr(), self.n_total, self.norm_ws.data_ptr(),
                                         self.norm_ws.numel(), self.gnorm.data_ptr(), 0, st), "sdb_global_norm")
            gn = self.gnorm.data_ptr()
        es4, es2 = 4, 2
        for q, wdf, s, e in self.segments:
            n = e - s
            g, p, sh = self.grad.data_ptr() + s * es4, self.master.data_ptr() + s * es4, self.shadow.data_ptr() + s * es2
            if q:
                _lib.check(l.sdb_lion8_update(g, p, p, sh, self.codes.data_ptr() + s,
                                              self.inv_scales.data_ptr() + (s // self.block_size) * 4, n,
                                              self.block_size, b1, b2, lr, wd, int(wdf), gn,
                                              1.0 if max_norm is None else max_norm, 0, st), "sdb_lion8_update")
            else:
                _lib.check(l.sdb_lion_fp32_update(g, p, p, sh, self.mu.data_ptr() + (s - self.n_quant) * 4, n, b1, b2,
                                                  lr, wd, int(wdf), gn, 1.0 if max_norm is None else max_norm, 0, st),
                           "sdb_lion_fp32_update")


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
        pred = st.net.forward(x, timesteps.to(torch.int32), ctx)
        st.net.graph = None
        out = pred.v.view(B, H, W, -1).permute(0, 3, 1, 2).float()  # layout change of the result only
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

    @classmethod
    def create(cls, *, apply_fn, params: dict, tx: LQ.GradientTransformation, device="cuda",
               reference_attention_quirk: bool = True) -> "TrainState":
        model = getattr(apply_fn, "__self__", None)
        if not isinstance(model, UNetModel):
            raise TypeError("TrainState.create: apply_fn must be UNetModel(...).apply")
        cfg = model.config
        flat = flatten_dict(params)
        shapes = param_shapes(cfg)
        if set(flat) != set(shapes):
            missing = sorted(set(shapes) - set(flat))[:3]
            extra = sorted(set(flat) - set(shapes))[:3]
            raise ValueError(f"params do not match the UNet config (missing {missing}, unexpected {extra})")
        hyper = _parse_chain(tx.hyper)
        ones = {n: True for n in shapes}
        quant = flatten_dict(LQ._expand_mask(hyper["excluded_layer_mask"], params)) if hyper["quantize"] else \
            {n: False for n in shapes}
        decay = flatten_dict(LQ._expand_mask(hyper["mask"], params)) if hyper["mask"] is not None else ones
        arena = ParamArena(shapes, quant, decay, backward_param_order(cfg), device, hyper["block_size"])
        arena.load(flat)
        tree = unflatten_dict(arena.p)
        net = UNet(cfg, arena.p, arena.w, arena.g, reference_attention_quirk)
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

    def apply_gradients(self, *, grads: Optional[dict] = None) -> "TrainState":
        """tx.update + optax.apply_updates + step+1, in place (the state is donated). `grads=None` or the state's own
        gradient tree uses the gradients already in the arena (what train_step produces)."""
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
        a.optimizer_step(lr=float(lr), b1=h["b1"], b2=h["b2"], wd=h["weight_decay"], max_norm=h["max_norm"])
        lion.lion.count.add_(1)
        lion.schedule.count.add_(1)
        self.step.add_(1)
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
    if out["block_size"] != 16:
        raise NotImplementedError("the fused optimizer path is built for block_size 16 (model_properties_example.json:41)")
    return out


# ---------------------------------------------------------------------------------------------------------------
_KEY = {}


def _generator(rng, device):
    if isinstance(rng, torch.Generator):
        return rng
    g = torch.Generator(device=device)
    g.manual_seed(int(rng) if rng is not None else 0)
    return g


def train_step(state: TrainState, batch: dict, rng, frozen_noise_scheduler_state: Optional[FrozenModel] = None, *,
               offset_noise_magnitude: float = 0.0, min_snr_gamma_magnitude: float = 0.0,
               perturbation_noise_magnitude: float = 0.0):
    """One denoising train step (training_utils.py:504-762, UNet part). batch: {"latents" f32 [B,4,H,W],
    "encoder_hidden_states" [B,L,D], optional "noise" f32 [B,4,H,W], "timesteps" int [B]} (device tensors).
    Returns (state, {"loss": 0-d device tensor}); `rng` (a torch.Generator or seed) is consumed for whatever the batch
    does not supply. Raises ValueError for an unknown prediction type (training_utils.py:697-701)."""
    if frozen_noise_scheduler_state is None:
        frozen_noise_scheduler_state = default_scheduler(state.arena.device)
    sched: FlaxDDPMScheduler = frozen_noise_scheduler_state.call
    sstate = frozen_noise_scheduler_state.params
    ptype = sched.config.prediction_type
    if ptype not in ("epsilon", "v_prediction"):
        raise ValueError(f"Unknown prediction type {ptype}")
    lat = batch["latents"]
    dev = lat.device
    B = lat.shape[0]
    gen = None
    noise = batch.get("noise")
    if noise is None:
        gen = _generator(rng, dev)
        noise = torch.randn(lat.shape, generator=gen, device=dev, dtype=torch.float32)
        if offset_noise_magnitude:
            noise = noise + torch.randn((B, lat.shape[1], 1, 1), generator=gen, device=dev) * offset_noise_magnitude
        if perturbation_noise_magnitude:
            noise = noise + perturbation_noise_magnitude * torch.randn(lat.shape, generator=gen, device=dev)
    timesteps = batch.get("timesteps")
    if timesteps is None:
        gen = gen or _generator(rng, dev)
        timesteps = torch.randint(0, sched.config.num_train_timesteps, (B,), generator=gen, device=dev)
    timesteps = timesteps.to(torch.int32)

    noisy = sched.add_noise(sstate, lat, noise, timesteps)
    target = noise if ptype == "epsilon" else sched.get_velocity(sstate, lat, noise, timesteps)
    weight = None
    if min_snr_gamma_magnitude:
        weight = _min_snr_weight(sstate.common.alphas_cumprod, timesteps, float(min_snr_gamma_magnitude),
                                 ptype == "v_prediction")
    x = ops.nchw_to_nhwc_bf16(noisy)
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
    state.apply_gradients(grads=None)
    return state, {"loss": loss}


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


def _min_snr_weight(alphas_cumprod, timesteps, gamma: float, v_pred: bool):
    """training_utils.py:546-568 — B scalars; host arithmetic in fp32 on the gathered alphas."""
    ac = alphas_cumprod[timesteps.long()].cpu().numpy().astype(np.float32)
    snr = ac / (np.float32(1) - ac)
    m = np.minimum(snr, np.float32(gamma))
    w = m / (snr + np.float32(1)) if v_pred else m / snr
    return torch.from_numpy(w.astype(np.float32)).to(alphas_cumprod.device)


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
    buffers and replay. Noise / timesteps missing from the batch are drawn OUTSIDE the graph into those buffers.
    The loss returned in `metrics` is a static device tensor, overwritten by the next call (async, like the reference's
    futures: the host blocks only when it reads it)."""

    def __init__(self, state: TrainState, frozen_noise_scheduler_state: Optional[FrozenModel] = None, *,
                 offset_noise_magnitude: float = 0.0, min_snr_gamma_magnitude: float = 0.0,
                 perturbation_noise_magnitude: float = 0.0):
        if min_snr_gamma_magnitude:
            raise NotImplementedError("CompiledTrainStep: min-SNR weighting gathers on the host; use train_step")
        lr = state.hyper["learning_rate"]
        if callable(lr) and lr(0) != lr(1_000_000):
            raise NotImplementedError("CompiledTrainStep: only a constant learning-rate schedule can be baked in")
        self.state = state
        self.sched = frozen_noise_scheduler_state or default_scheduler(state.arena.device)
        self.offset = float(offset_noise_magnitude)
        self.perturb = float(perturbation_noise_magnitude)
        self._entries: Dict[tuple, dict] = {}
        self.launches_per_replay = 0

    def _eager(self, batch, rng):
        return train_step(self.state, batch, rng, self.sched, offset_noise_magnitude=self.offset,
                          perturbation_noise_magnitude=self.perturb)

    def _fill(self, e, batch, rng):
        dev = self.state.arena.device
        buf = e["buf"]
        buf["latents"].copy_(batch["latents"], non_blocking=True)
        ctx = batch["encoder_hidden_states"]
        buf["encoder_hidden_states"].copy_(ctx, non_blocking=True)
        B = buf["latents"].shape[0]
        gen = None
        if batch.get("noise") is not None:
            buf["noise"].copy_(batch["noise"], non_blocking=True)
        else:
            gen = _generator(rng, dev)
            shape = buf["latents"].shape
            noise = torch.randn(shape, generator=gen, device=dev, dtype=torch.float32)
            if self.offset:
                noise = noise + torch.randn((B, shape[1], 1, 1), generator=gen, device=dev) * self.offset
            if self.perturb:
                noise = noise + self.perturb * torch.randn(shape, generator=gen, device=dev)
            buf["noise"].copy_(noise)
        if batch.get("timesteps") is not None:
            buf["timesteps"].copy_(batch["timesteps"], non_blocking=True)
        else:
            gen = gen or _generator(rng, dev)
            buf["timesteps"].copy_(torch.randint(0, self.sched.call.config.num_train_timesteps, (B,), generator=gen,
                                                 device=dev).to(torch.int32))

    def __call__(self, state: TrainState, batch: dict, rng=None):
        if state is not self.state:
            raise ValueError("CompiledTrainStep: called with a different TrainState than it was compiled for")
        key = (tuple(batch["latents"].shape), tuple(batch["encoder_hidden_states"].shape))
        e = self._entries.get(key)
        if e is None:
            self._entries[key] = {"graph": None}
            return self._eager(batch, rng)  # also the warm-up of every lazily built cache
        if e["graph"] is None:
            dev = state.arena.device
            e["buf"] = {"latents": torch.empty(key[0], dtype=torch.float32, device=dev),
                        "noise": torch.empty(key[0], dtype=torch.float32, device=dev),
                        "timesteps": torch.empty((key[0][0],), dtype=torch.int32, device=dev),
                        "encoder_hidden_states": torch.empty(key[1], dtype=batch["encoder_hidden_states"].dtype,
                                                             device=dev)}
            self._fill(e, batch, rng)
            self._capture(e)
        else:
            self._fill(e, batch, rng)
        self._replay(e)
        state._host_step += 1
        return state, e["metrics"]

    def _capture(self, e) -> None:
        """Capture the step as a chain of CUDA-graph segments sharing one memory pool. Without data parallelism that
        is one segment. With it, the NCCL collectives stay OUT of the graphs: the reducer closes a segment whenever
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
            cur["g"].capture_begin(pool=pool)

        def cut(final: bool):
            cur["g"].capture_end()
            segs.append((cur["g"], list(dp._deferred), final))
            dp._deferred.clear()
            begin()

        with torch.cuda.stream(stream):
            if dp is not None:
                dp.capture = cut
            try:
                begin()
                _, metrics = train_step(state, e["buf"], None, self.sched)
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
