"""DDPM scheduler behind the reference's surface (/root/reference/schedulers/scheduling_ddpm_flax.py:36-300,
scheduling_utils_flax.py:188-343): FlaxDDPMScheduler with create_state / add_noise / get_velocity / step /
set_timesteps / scale_model_input / __len__, the immutable CommonSchedulerState / DDPMSchedulerState dataclasses and
FlaxDDPMSchedulerOutput. Same argument names, defaults and exceptions.

State creation (1000-element schedules) is init-time host arithmetic in fp32 NumPy in the reference's op order; the
per-sample arithmetic (add_noise, get_velocity, step) runs in libsdb200.so on CUDA tensors.
"""
from __future__ import annotations

import dataclasses
import math
from dataclasses import dataclass
from types import SimpleNamespace
from typing import Optional, Tuple, Union

import numpy as np
import torch

from . import _lib

_F = np.float32


def broadcast_to_shape_from_left(x: torch.Tensor, shape) -> torch.Tensor:
    assert len(shape) >= x.dim()
    return x.reshape(tuple(x.shape) + (1,) * (len(shape) - x.dim())).expand(*shape)


def betas_for_alpha_bar(num_diffusion_timesteps: int, max_beta=0.999) -> np.ndarray:
    def alpha_bar(time_step):
        return math.cos((time_step + 0.008) / 1.008 * math.pi / 2) ** 2

    betas = []
    for i in range(num_diffusion_timesteps):
        t1 = i / num_diffusion_timesteps
        t2 = (i + 1) / num_diffusion_timesteps
        betas.append(min(1 - alpha_bar(t2) / alpha_bar(t1), max_beta))
    return np.array(betas, dtype=_F)


def _linspace(start, stop, num) -> np.ndarray:
    # jnp.linspace in fp32: start*(1-s) + stop*s, s = iota(num-1)/(num-1), endpoint appended exactly
    s = np.arange(num - 1, dtype=_F) / _F(num - 1)
    body = _F(start) * (_F(1) - s) + _F(stop) * s
    return np.append(body, _F(stop)).astype(_F)


def rescale_betas(betas: np.ndarray) -> np.ndarray:
    """Rescales betas to have zero terminal SNR (scheduling_utils_flax.py:222-263, arXiv 2305.08891 alg. 1)."""
    alphas = _F(1) - betas
    alphas_bar = np.cumprod(alphas, dtype=_F)
    alphas_bar_sqrt = np.sqrt(alphas_bar)
    a0 = alphas_bar_sqrt[0].copy()
    aT = alphas_bar_sqrt[-1].copy()
    alphas_bar_sqrt = alphas_bar_sqrt - aT
    alphas_bar_sqrt = alphas_bar_sqrt * a0 / (a0 - aT)
    alphas_bar = alphas_bar_sqrt * alphas_bar_sqrt
    alphas = alphas_bar[1:] / alphas_bar[:-1]
    alphas = np.concatenate([alphas_bar[0:1], alphas])
    return (_F(1) - alphas).astype(_F)


@dataclass(frozen=True)
class CommonSchedulerState:
    alphas: torch.Tensor
    betas: torch.Tensor
    alphas_cumprod: torch.Tensor

    @classmethod
    def create(cls, scheduler):
        config = scheduler.config
        n = config.num_train_timesteps
        if config.trained_betas is not None:
            betas = np.asarray(config.trained_betas, dtype=_F)
        elif config.beta_schedule == "linear":
            betas = _linspace(config.beta_start, config.beta_end, n)
        elif config.beta_schedule == "scaled_linear":
            b = _linspace(config.beta_start ** 0.5, config.beta_end ** 0.5, n)
            betas = (b * b).astype(_F)
        elif config.beta_schedule == "zero_snr_scaled_linear":
            b = _linspace(config.beta_start ** 0.5, config.beta_end ** 0.5, n)
            betas = rescale_betas((b * b).astype(_F))
        elif config.beta_schedule == "squaredcos_cap_v2":
            betas = betas_for_alpha_bar(n)
        else:
            raise NotImplementedError(
                f"beta_schedule {config.beta_schedule} is not implemented for scheduler {scheduler.__class__.__name__}")
        alphas = (_F(1.0) - betas).astype(_F)
        alphas_cumprod = np.cumprod(alphas, dtype=_F)
        dev = scheduler.device
        return cls(alphas=torch.from_numpy(alphas).to(dev), betas=torch.from_numpy(betas).to(dev),
                   alphas_cumprod=torch.from_numpy(alphas_cumprod).to(dev))


@dataclass(frozen=True)
class DDPMSchedulerState:
    common: CommonSchedulerState
    init_noise_sigma: torch.Tensor
    timesteps: torch.Tensor
    num_inference_steps: Optional[int] = None

    @classmethod
    def create(cls, common, init_noise_sigma, timesteps):
        return cls(common=common, init_noise_sigma=init_noise_sigma, timesteps=timesteps)

    def replace(self, **kw):
        return dataclasses.replace(self, **kw)


@dataclass
class FlaxSchedulerOutput:
    prev_sample: torch.Tensor


@dataclass
class FlaxDDPMSchedulerOutput(FlaxSchedulerOutput):
    state: DDPMSchedulerState


def _as_f32(t: torch.Tensor, device, what: str) -> torch.Tensor:
    """The kernels read fp32: bf16 / fp16 / fp64 inputs (the reference's VAE runs in bf16, training_utils.py:209-222)
    are promoted like jnp would promote them against the fp32 schedule; anything on another device is an error."""
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{what}: expected a torch tensor, got {type(t).__name__}")
    if t.device != device:
        raise ValueError(f"{what}: tensor is on {t.device}, the scheduler state on {device}")
    if not t.dtype.is_floating_point:
        raise TypeError(f"{what}: expected a floating-point tensor, got {t.dtype}")
    return t.to(torch.float32).contiguous()


def _noise_mix(fn_name, state: CommonSchedulerState, a, noise, timesteps):
    l = _lib.lib()
    ac = state.alphas_cumprod
    a = _as_f32(a, ac.device, fn_name)
    noise = _as_f32(noise, ac.device, fn_name)
    if noise.shape != a.shape:
        raise ValueError(f"{fn_name}: sample {tuple(a.shape)} and noise {tuple(noise.shape)} shapes differ")
    if timesteps.device != ac.device:
        raise ValueError(f"{fn_name}: timesteps are on {timesteps.device}, the scheduler state on {ac.device}")
    t = timesteps.to(torch.int32).contiguous().view(-1)
    B = a.shape[0]
    if t.numel() != B:
        raise ValueError(f"{fn_name}: {t.numel()} timesteps for a batch of {B}")
    out = torch.empty_like(a)
    # out-of-range timesteps are wrapped / clamped inside the kernel like jnp's gather (never read out of bounds)
    _lib.check(getattr(l, fn_name)(a.data_ptr(), noise.data_ptr(), t.data_ptr(), ac.data_ptr(), ac.numel(),
                                   out.data_ptr(), B, a.numel() // B, _lib.stream_ptr()), fn_name)
    return out


def add_noise_common(state: CommonSchedulerState, original_samples, noise, timesteps):
    """sqrt(ac[t]) * x0 + sqrt(1 - ac[t]) * noise   (scheduling_utils_flax.py:332-337)."""
    return _noise_mix("sdb_add_noise", state, original_samples, noise, timesteps)


def get_velocity_common(state: CommonSchedulerState, sample, noise, timesteps):
    """sqrt(ac[t]) * noise - sqrt(1 - ac[t]) * sample   (scheduling_utils_flax.py:340-343)."""
    return _noise_mix("sdb_get_velocity", state, sample, noise, timesteps)


class FlaxDDPMScheduler:
    """scheduling_ddpm_flax.py:55-300."""

    has_state = True

    def __init__(self, num_train_timesteps: int = 1000, beta_start: float = 0.0001, beta_end: float = 0.02,
                 beta_schedule: str = "linear", trained_betas=None, variance_type: str = "fixed_small",
                 clip_sample: bool = True, prediction_type: str = "epsilon", dtype=torch.float32, device="cuda"):
        if dtype not in (torch.float32, "float32"):
            raise NotImplementedError("FlaxDDPMScheduler: only dtype float32 (the reference default) is built")
        self.dtype = torch.float32
        self.device = torch.device(device)
        self.config = SimpleNamespace(num_train_timesteps=num_train_timesteps, beta_start=beta_start, beta_end=beta_end,
                                      beta_schedule=beta_schedule, trained_betas=trained_betas,
                                      variance_type=variance_type, clip_sample=clip_sample,
                                      prediction_type=prediction_type)
        self._host = None  # host copies of the schedules for the scalar arithmetic of step()

    def create_state(self, common: Optional[CommonSchedulerState] = None) -> DDPMSchedulerState:
        if common is None:
            common = CommonSchedulerState.create(self)
        init_noise_sigma = torch.tensor(1.0, dtype=self.dtype, device=self.device)
        timesteps = torch.arange(0, self.config.num_train_timesteps, device=self.device).flip(0)
        return DDPMSchedulerState.create(common=common, init_noise_sigma=init_noise_sigma, timesteps=timesteps)

    def scale_model_input(self, state, sample, timestep: Optional[int] = None):
        return sample

    def set_timesteps(self, state: DDPMSchedulerState, num_inference_steps: int, shape: Tuple = ()):
        step_ratio = self.config.num_train_timesteps // num_inference_steps
        timesteps = (torch.arange(0, num_inference_steps, device=self.device) * step_ratio).flip(0)
        return state.replace(num_inference_steps=num_inference_steps, timesteps=timesteps)

    def _host_tables(self, state):
        if self._host is None or self._host[0] is not state.common:
            c = state.common
            self._host = (c, c.alphas.cpu().numpy(), c.betas.cpu().numpy(), c.alphas_cumprod.cpu().numpy())
        return self._host[1:]

    def _get_variance(self, state, t, predicted_variance=None, variance_type=None):
        alphas, betas, ac = self._host_tables(state)
        alpha_prod_t = ac[t]
        alpha_prod_t_prev = ac[t - 1] if t > 0 else _F(1.0)
        variance = _F(_F(_F(1) - alpha_prod_t_prev) / _F(_F(1) - alpha_prod_t)) * betas[t]
        if variance_type is None:
            variance_type = self.config.variance_type
        if variance_type == "fixed_small":
            variance = np.maximum(variance, _F(1e-20))
        elif variance_type == "fixed_small_log":
            variance = np.log(np.maximum(variance, _F(1e-20)))
        elif variance_type == "fixed_large":
            variance = betas[t]
        elif variance_type == "fixed_large_log":
            variance = np.log(betas[t])
        elif variance_type == "learned":
            return predicted_variance
        elif variance_type == "learned_range":
            raise NotImplementedError("learned_range variance needs the model's variance head (not on the SD path)")
        return _F(variance)

    def step(self, state: DDPMSchedulerState, model_output, timestep: int, sample, key=None, return_dict: bool = True
             ) -> Union[FlaxDDPMSchedulerOutput, Tuple]:
        """Reverse-diffusion step (scheduling_ddpm_flax.py:196-279). `key`: None (seed 0), an int seed or a
        torch.Generator for the N(0,1) draw (jax.random keys do not exist here); a float32 tensor of the sample's
        shape is used AS the draw (parity tests inject the reference's noise that way)."""
        t = int(timestep)
        ptype = self.config.prediction_type
        if ptype not in ("epsilon", "sample", "v_prediction"):
            raise ValueError(f"prediction_type given as {ptype} must be one of `epsilon`, `sample` "
                             " for the FlaxDDPMScheduler.")
        if self.config.variance_type in ("learned", "learned_range"):
            raise NotImplementedError("learned variance types are not on the SD training path")
        alphas, betas, ac = self._host_tables(state)
        alpha_prod_t = ac[t]
        alpha_prod_t_prev = ac[t - 1] if t > 0 else _F(1.0)
        beta_prod_t = _F(_F(1) - alpha_prod_t)
        beta_prod_t_prev = _F(_F(1) - alpha_prod_t_prev)
        x0_coeff = _F(_F(np.sqrt(alpha_prod_t_prev) * betas[t]) / beta_prod_t)
        sample_coeff = _F(_F(np.sqrt(alphas[t]) * beta_prod_t_prev) / beta_prod_t)
        model_output = model_output.contiguous()
        sample = sample.contiguous()
        noise = None
        sigma = 0.0
        if t > 0:
            if isinstance(key, torch.Tensor) and key.shape == model_output.shape:
                noise = key.to(device=sample.device, dtype=torch.float32).contiguous()
            else:
                gen = key
                if not isinstance(gen, torch.Generator):
                    gen = torch.Generator(device=sample.device)
                    gen.manual_seed(0 if key is None else int(key))
                noise = torch.randn(model_output.shape, generator=gen, device=sample.device, dtype=torch.float32)
            sigma = float(np.sqrt(self._get_variance(state, t)))
        prev = torch.empty_like(sample)
        _lib.check(_lib.lib().sdb_ddpm_step(model_output.data_ptr(), sample.data_ptr(), _lib.ptr(noise), prev.data_ptr(),
                                            sample.numel(), float(np.sqrt(alpha_prod_t)), float(np.sqrt(beta_prod_t)),
                                            float(x0_coeff), float(sample_coeff), sigma,
                                            ("epsilon", "sample", "v_prediction").index(ptype),
                                            int(bool(self.config.clip_sample)), _lib.stream_ptr()), "sdb_ddpm_step")
        if not return_dict:
            return (prev, state)
        return FlaxDDPMSchedulerOutput(prev_sample=prev, state=state)

    def add_noise(self, state: DDPMSchedulerState, original_samples, noise, timesteps):
        return add_noise_common(state.common, original_samples, noise, timesteps)

    def get_velocity(self, state: DDPMSchedulerState, sample, noise, timesteps):
        return get_velocity_common(state.common, sample, noise, timesteps)

    def __len__(self):
        return self.config.num_train_timesteps


# ---------------------------------------------------------------------------------------------------------------
@dataclass(frozen=True)
class DDIMSchedulerState:
    """schedulers/scheduling_ddim_flax.py:36-57."""
    common: CommonSchedulerState
    final_alpha_cumprod: torch.Tensor
    init_noise_sigma: torch.Tensor
    timesteps: torch.Tensor
    num_inference_steps: Optional[int] = None

    @classmethod
    def create(cls, common, final_alpha_cumprod, init_noise_sigma, timesteps):
        return cls(common=common, final_alpha_cumprod=final_alpha_cumprod, init_noise_sigma=init_noise_sigma,
                   timesteps=timesteps)

    def replace(self, **kw):
        return dataclasses.replace(self, **kw)


@dataclass
class FlaxDDIMSchedulerOutput(FlaxSchedulerOutput):
    state: DDIMSchedulerState


class FlaxDDIMScheduler:
    """schedulers/scheduling_ddim_flax.py:66-305 — the scheduler the reference writes into saved pipelines
    (training_utils.py:997-1004) and the one its validation sampling uses. Same constructor, state and methods; the
    elementwise part of step() is sdb_ddim_step, the scalars are fp32 host arithmetic in the reference's order."""

    has_state = True

    def __init__(self, num_train_timesteps: int = 1000, beta_start: float = 0.0001, beta_end: float = 0.02,
                 beta_schedule: str = "linear", trained_betas=None, set_alpha_to_one: bool = True,
                 steps_offset: int = 0, prediction_type: str = "epsilon", dtype=torch.float32, device="cuda"):
        if dtype not in (torch.float32, "float32"):
            raise NotImplementedError("FlaxDDIMScheduler: only dtype float32 (the reference default) is built")
        self.dtype = torch.float32
        self.device = torch.device(device)
        self.config = SimpleNamespace(num_train_timesteps=num_train_timesteps, beta_start=beta_start, beta_end=beta_end,
                                      beta_schedule=beta_schedule, trained_betas=trained_betas,
                                      set_alpha_to_one=set_alpha_to_one, steps_offset=steps_offset,
                                      prediction_type=prediction_type)
        self._host = None

    def create_state(self, common: Optional[CommonSchedulerState] = None) -> DDIMSchedulerState:
        if common is None:
            common = CommonSchedulerState.create(self)
        final = torch.tensor(1.0, dtype=self.dtype, device=self.device) if self.config.set_alpha_to_one \
            else common.alphas_cumprod[0]
        init_noise_sigma = torch.tensor(1.0, dtype=self.dtype, device=self.device)
        timesteps = torch.arange(0, self.config.num_train_timesteps, device=self.device).flip(0)
        return DDIMSchedulerState.create(common=common, final_alpha_cumprod=final, init_noise_sigma=init_noise_sigma,
                                         timesteps=timesteps)

    def scale_model_input(self, state, sample, timestep: Optional[int] = None):
        return sample

    def set_timesteps(self, state: DDIMSchedulerState, num_inference_steps: int, shape: Tuple = ()):
        step_ratio = self.config.num_train_timesteps // num_inference_steps
        timesteps = (torch.arange(0, num_inference_steps, device=self.device) * step_ratio).flip(0) \
            + self.config.steps_offset
        return state.replace(num_inference_steps=num_inference_steps, timesteps=timesteps)

    def _tables(self, state):
        if self._host is None or self._host[0] is not state:
            self._host = (state, state.common.alphas_cumprod.cpu().numpy(),
                          _F(state.final_alpha_cumprod.cpu().numpy()))
        return self._host[1], self._host[2]

    def _alpha_pair(self, state, timestep: int, prev_timestep: int):
        ac, final = self._tables(state)
        n = ac.shape[0]
        idx = lambda t: min(max(t + n if t < 0 else t, 0), n - 1)  # jnp gather: wrap a negative index once, clamp
        return ac[idx(timestep)], (ac[idx(prev_timestep)] if prev_timestep >= 0 else final)

    def _get_variance(self, state, timestep: int, prev_timestep: int):
        a_t, a_prev = self._alpha_pair(state, int(timestep), int(prev_timestep))
        beta_t, beta_prev = _F(_F(1) - a_t), _F(_F(1) - a_prev)
        return _F(_F(beta_prev / beta_t) * _F(_F(1) - _F(a_t / a_prev)))

    def step(self, state: DDIMSchedulerState, model_output, timestep: int, sample, eta: float = 0.0,
             return_dict: bool = True):
        if state.num_inference_steps is None:
            raise ValueError("Number of inference steps is 'None', you need to run 'set_timesteps' after creating "
                             "the scheduler")
        ptype = self.config.prediction_type
        if ptype not in ("epsilon", "sample", "v_prediction"):
            raise ValueError(f"prediction_type given as {ptype} must be one of `epsilon`, `sample`, or `v_prediction`")
        t = int(timestep)
        prev_t = t - self.config.num_train_timesteps // state.num_inference_steps
        a_t, a_prev = self._alpha_pair(state, t, prev_t)
        beta_t = _F(_F(1) - a_t)
        variance = self._get_variance(state, t, prev_t)
        std = _F(_F(eta) * _F(np.sqrt(variance)))
        cdir = _F(np.sqrt(_F(_F(_F(1) - a_prev) - _F(std * std))))
        model_output = _as_f32(model_output, self.device, "FlaxDDIMScheduler.step")
        sample = _as_f32(sample, self.device, "FlaxDDIMScheduler.step")
        prev = torch.empty_like(sample)
        _lib.check(_lib.lib().sdb_ddim_step(model_output.data_ptr(), sample.data_ptr(), prev.data_ptr(), sample.numel(),
                                            float(np.sqrt(a_t)), float(np.sqrt(beta_t)), float(np.sqrt(a_prev)),
                                            float(cdir), ("epsilon", "sample", "v_prediction").index(ptype),
                                            _lib.stream_ptr(self.device)), "sdb_ddim_step")
        if not return_dict:
            return (prev, state)
        return FlaxDDIMSchedulerOutput(prev_sample=prev, state=state)

    def add_noise(self, state: DDIMSchedulerState, original_samples, noise, timesteps):
        return add_noise_common(state.common, original_samples, noise, timesteps)

    def get_velocity(self, state: DDIMSchedulerState, sample, noise, timesteps):
        return get_velocity_common(state.common, sample, noise, timesteps)

    def __len__(self):
        return self.config.num_train_timesteps
