"""ORACLE (test infrastructure only — never imported by the product path).

CPU restatement, in NumPy fp32, of the reference's scheduler arithmetic on the train-step path:

  /root/reference/schedulers/scheduling_utils_flax.py:188-190   broadcast_to_shape_from_left
  /root/reference/schedulers/scheduling_utils_flax.py:193-219   betas_for_alpha_bar (squaredcos_cap_v2)
  /root/reference/schedulers/scheduling_utils_flax.py:222-263   rescale_betas (zero terminal SNR)
  /root/reference/schedulers/scheduling_utils_flax.py:266-313   CommonSchedulerState.create
  /root/reference/schedulers/scheduling_utils_flax.py:316-343   get_sqrt_alpha_prod / add_noise_common / get_velocity_common
  /root/reference/schedulers/scheduling_ddpm_flax.py:111-124    create_state
  /root/reference/schedulers/scheduling_ddpm_flax.py:163-194    _get_variance
  /root/reference/schedulers/scheduling_ddpm_flax.py:196-279    step
  /root/reference/training_utils.py:546-568                     min_snr_gamma_loss_rescale
  /root/reference/training_utils.py:688-709                     target selection + MSE

Every op is an individually rounded fp32 op in the reference's order. `x ** 0.5` is the correctly rounded fp32 sqrt
(XLA lowers pow(x, 0.5) to sqrt); `jnp.linspace` is restated as start + i * ((stop-start)/(n-1)) in fp32 with the
endpoint pinned, `cumprod` as the sequential fp32 product.

Parity pin: tests/golden/sched_ref_*.npz are produced by running the UNMODIFIED reference scheduler files under a
NumPy-backed `jax.numpy` shim (oracle/jax_shim.py, oracle/make_golden.py); tests/test_oracle_golden.py checks this
restatement against them.
"""
from __future__ import annotations

import math

import numpy as np

F32 = np.float32


def linspace_f32(start: float, stop: float, num: int) -> np.ndarray:
    """jnp.linspace(start, stop, num, dtype=float32) [upstream jax]: computed in fp32 as
    start*(1-t) + stop*t with t = arange(num-1)/ (num-1), endpoint appended."""
    start32, stop32 = F32(start), F32(stop)
    div = num - 1
    step = np.arange(div, dtype=F32) / F32(div)
    out = start32 * (F32(1) - step) + stop32 * step
    return np.concatenate([out, np.array([stop32], F32)]).astype(F32)


def cumprod_f32(a: np.ndarray) -> np.ndarray:
    out = np.empty_like(a, dtype=F32)
    acc = F32(1)
    for i, v in enumerate(a.astype(F32)):
        acc = F32(acc * v)
        out[i] = acc
    return out


def betas_for_alpha_bar(n: int, max_beta: float = 0.999) -> np.ndarray:
    def alpha_bar(t):
        return math.cos((t + 0.008) / 1.008 * math.pi / 2) ** 2

    return np.array([min(1 - alpha_bar((i + 1) / n) / alpha_bar(i / n), max_beta) for i in range(n)], F32)


def rescale_betas(betas: np.ndarray) -> np.ndarray:
    """scheduling_utils_flax.py:222-263."""
    alphas = F32(1) - betas
    alphas_bar = cumprod_f32(alphas)
    abs_ = np.sqrt(alphas_bar).astype(F32)
    a0, aT = abs_[0], abs_[-1]
    abs_ = abs_ - aT
    abs_ = abs_ * a0 / (a0 - aT)
    alphas_bar = abs_ * abs_  # integer_pow 2
    alphas = alphas_bar[1:] / alphas_bar[:-1]
    alphas = np.concatenate([alphas_bar[0:1], alphas]).astype(F32)
    return (F32(1) - alphas).astype(F32)


def make_betas(beta_schedule: str, beta_start: float, beta_end: float, n: int, trained_betas=None) -> np.ndarray:
    """CommonSchedulerState.create, scheduling_utils_flax.py:266-303."""
    if trained_betas is not None:
        return np.asarray(trained_betas, F32)
    if beta_schedule == "linear":
        return linspace_f32(beta_start, beta_end, n)
    if beta_schedule in ("scaled_linear", "zero_snr_scaled_linear"):
        b = linspace_f32(beta_start ** 0.5, beta_end ** 0.5, n)
        b = (b * b).astype(F32)
        return rescale_betas(b) if beta_schedule == "zero_snr_scaled_linear" else b
    if beta_schedule == "squaredcos_cap_v2":
        return betas_for_alpha_bar(n)
    raise NotImplementedError(f"beta_schedule {beta_schedule} is not implemented")


def common_state(beta_schedule="linear", beta_start=0.0001, beta_end=0.02, n=1000, trained_betas=None):
    betas = make_betas(beta_schedule, beta_start, beta_end, n, trained_betas)
    alphas = (F32(1.0) - betas).astype(F32)
    return {"betas": betas, "alphas": alphas, "alphas_cumprod": cumprod_f32(alphas)}


def _coef(alphas_cumprod, timesteps, ndim):
    ac = alphas_cumprod[np.asarray(timesteps)]
    a = np.sqrt(ac).astype(F32)
    s = np.sqrt((F32(1) - ac).astype(F32)).astype(F32)
    shp = a.shape + (1,) * (ndim - a.ndim)
    return a.reshape(shp), s.reshape(shp)


def add_noise(alphas_cumprod, original_samples, noise, timesteps):
    a, s = _coef(alphas_cumprod, timesteps, original_samples.ndim)
    return (a * original_samples.astype(F32) + s * noise.astype(F32)).astype(F32)


def get_velocity(alphas_cumprod, sample, noise, timesteps):
    a, s = _coef(alphas_cumprod, timesteps, sample.ndim)
    return (a * noise.astype(F32) - s * sample.astype(F32)).astype(F32)


def get_variance(state, t: int, variance_type="fixed_small", predicted_variance=None):
    """scheduling_ddpm_flax.py:163-194 (scalar t)."""
    ac, betas = state["alphas_cumprod"], state["betas"]
    ap = ac[t]
    ap_prev = ac[t - 1] if t > 0 else F32(1.0)
    variance = F32(F32(F32(1) - ap_prev) / F32(F32(1) - ap)) * betas[t]
    if variance_type == "fixed_small":
        variance = np.maximum(variance, F32(1e-20))
    elif variance_type == "fixed_small_log":
        variance = np.log(np.maximum(variance, F32(1e-20))).astype(F32)
    elif variance_type == "fixed_large":
        variance = betas[t]
    elif variance_type == "fixed_large_log":
        variance = np.log(betas[t]).astype(F32)
    elif variance_type == "learned":
        return predicted_variance
    elif variance_type == "learned_range":
        frac = (predicted_variance + F32(1)) / F32(2)
        variance = frac * betas[t] + (F32(1) - frac) * variance
    return F32(variance)


def step_coefficients(state, t: int):
    """Scalars of scheduling_ddpm_flax.py:229-264 in fp32."""
    ac, betas, alphas = state["alphas_cumprod"], state["betas"], state["alphas"]
    ap = ac[t]
    ap_prev = ac[t - 1] if t > 0 else F32(1.0)
    bp = F32(F32(1) - ap)
    bp_prev = F32(F32(1) - ap_prev)
    return {
        "sqrt_alpha_prod": np.sqrt(ap).astype(F32), "sqrt_beta_prod": np.sqrt(bp).astype(F32),
        "x0_coeff": F32(F32(np.sqrt(ap_prev).astype(F32) * betas[t]) / bp),
        "sample_coeff": F32(F32(np.sqrt(alphas[t]).astype(F32) * bp_prev) / bp),
    }


def ddpm_step(state, model_output, t: int, sample, noise=None, prediction_type="epsilon", clip_sample=True,
              variance_type="fixed_small"):
    """scheduling_ddpm_flax.py:196-279 with the N(0,1) draw passed in (`noise`), fixed_* variance types."""
    c = step_coefficients(state, t)
    mo, x = model_output.astype(F32), sample.astype(F32)
    if prediction_type == "epsilon":
        x0 = (x - c["sqrt_beta_prod"] * mo) / c["sqrt_alpha_prod"]
    elif prediction_type == "sample":
        x0 = mo
    elif prediction_type == "v_prediction":
        x0 = c["sqrt_alpha_prod"] * x - c["sqrt_beta_prod"] * mo
    else:
        raise ValueError(f"prediction_type given as {prediction_type} must be one of `epsilon`, `sample` "
                         " for the FlaxDDPMScheduler.")
    if clip_sample:
        x0 = np.clip(x0, F32(-1), F32(1))
    prev = c["x0_coeff"] * x0 + c["sample_coeff"] * x
    if t > 0:
        sigma = np.sqrt(get_variance(state, t, variance_type)).astype(F32)
        prev = prev + sigma * noise.astype(F32)
    else:
        prev = prev + np.zeros_like(prev)
    return prev.astype(F32)


def ddim_variance(state, t: int, prev_t: int, final_alpha_cumprod=F32(1.0)):
    """scheduling_ddim_flax.py:186-197 (_get_variance), fp32."""
    ac = state["alphas_cumprod"]
    ap = ac[t]
    ap_prev = ac[prev_t] if prev_t >= 0 else F32(final_alpha_cumprod)
    bp, bp_prev = F32(F32(1) - ap), F32(F32(1) - ap_prev)
    return F32(F32(bp_prev / bp) * F32(F32(1) - F32(ap / ap_prev)))


def ddim_step(state, model_output, t: int, sample, num_inference_steps: int, prediction_type="epsilon", eta=0.0,
              num_train_timesteps=1000, final_alpha_cumprod=F32(1.0)):
    """scheduling_ddim_flax.py:199-284: x0 / epsilon from the model output, sigma_t(eta), direction, x_{t-1}. fp32,
    one rounded op per reference op (x ** 0.5 == correctly rounded sqrt)."""
    ac = state["alphas_cumprod"]
    prev_t = t - num_train_timesteps // num_inference_steps
    ap = ac[t]
    ap_prev = ac[prev_t] if prev_t >= 0 else F32(final_alpha_cumprod)
    bp = F32(F32(1) - ap)
    sa, sb = np.sqrt(ap).astype(F32), np.sqrt(bp).astype(F32)
    mo, x = model_output.astype(F32), sample.astype(F32)
    if prediction_type == "epsilon":
        x0 = ((x - sb * mo) / sa).astype(F32)
        eps = mo
    elif prediction_type == "sample":
        x0 = mo
        eps = ((x - sa * x0) / sb).astype(F32)
    elif prediction_type == "v_prediction":
        x0 = (sa * x - sb * mo).astype(F32)
        eps = (sa * mo + sb * x).astype(F32)
    else:
        raise ValueError(f"prediction_type given as {prediction_type} must be one of `epsilon`, `sample`, or"
                         " `v_prediction`")
    var = ddim_variance(state, t, prev_t, final_alpha_cumprod)
    std = F32(F32(eta) * np.sqrt(var).astype(F32))
    cdir = np.sqrt(F32(F32(F32(1) - ap_prev) - F32(std * std))).astype(F32)
    return (np.sqrt(ap_prev).astype(F32) * x0 + cdir * eps).astype(F32)


def min_snr_weight(alphas_cumprod, timesteps, gamma: float, v_prediction: bool):
    """training_utils.py:546-568: snr = ac/(1-ac); w = min(snr, gamma)/snr (eps) or /(snr+1) (v)."""
    ac = alphas_cumprod[np.asarray(timesteps)]
    snr = (ac / (F32(1) - ac)).astype(F32)
    m = np.minimum(snr, F32(gamma))
    return (m / (snr + F32(1)) if v_prediction else m / snr).astype(F32)


def mse_loss(pred, target, weight=None):
    """training_utils.py:704-709: mean((target - pred)^2 [* w_b])."""
    d = (target.astype(F32) - pred.astype(F32))
    l = d * d
    if weight is not None:
        l = l * weight.reshape((-1,) + (1,) * (l.ndim - 1)).astype(F32)
    return np.mean(l, dtype=np.float64)
