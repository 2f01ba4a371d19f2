"""ORACLE (test infrastructure only — never imported by the product path).

CPU restatement, in NumPy fp32, of the reference's block-quantised 8-bit Lion:

  /root/reference/lion_quant.py:47-64    offset / _quantize / _dequantize
  /root/reference/lion_quant.py:66-92    _block_quantize / _block_dequantize
  /root/reference/lion_quant.py:97-113   _update_moment_quant
  /root/reference/lion_quant.py:115-154  init_fn / update_fn
  /root/reference/lion_quant.py:201-211  chain(scale_by_lion_8bit, add_decayed_weights, scale_by_learning_rate)
  /root/reference/training_utils.py:355-382  clip_by_global_norm(1) -> lion_8bit(lr/7, .9, .99, wd*7)

Third-party pieces restated from their published behaviour (not vendored under /root/reference):
optax (unpinned, ~0.1.7) clip_by_global_norm / add_decayed_weights / scale_by_learning_rate /
apply_updates; jax integer_pow lowering (x**5 = x*((x*x)*(x*x))).

Every op is an individually rounded IEEE fp32 op in the reference's order (no FMA contraction).
`pow(a, 0.2f)` is *defined* here as fl32(pow_f64(a, (double)0.2f)); the CUDA kernel reproduces it through
the monotone threshold table `code_thresholds()` below.

Parity pin: tests/golden/lion_ref_*.npz are produced by running the UNMODIFIED reference file
/root/reference/lion_quant.py under a NumPy-backed `jax.numpy` shim (oracle/jax_shim.py,
oracle/make_golden.py); tests/test_oracle_golden.py checks this restatement against them bit for bit.
"""
from __future__ import annotations

import numpy as np

F32 = np.float32
OFFSET = F32(3.7398995e-09)  # lion_quant.py:49
EXPONENT = F32(1 / 5)        # lion_quant.py:56: python float 0.2 -> weak-typed -> fp32


def _pow_fifth_root(a: np.ndarray) -> np.ndarray:
    """fl32(pow_f64(a, (double)0.2f)) — the oracle's definition of jnp.power(a, 1/5) in fp32."""
    return np.power(a.astype(np.float64), np.float64(EXPONENT)).astype(F32)


def quantize(x: np.ndarray) -> np.ndarray:
    """lion_quant.py:52-59."""
    x = x.astype(F32, copy=False)
    v = x + OFFSET
    q_sign = np.sign(v)
    q = _pow_fifth_root(np.abs(v))
    q = (q * q_sign) * F32(127)
    return np.rint(q).astype(np.int8)  # jnp.round = half-to-even


def dequant_lut() -> np.ndarray:
    """256-entry table indexed by the uint8 view of the int8 code. lion_quant.py:61-64.

    (q / 127) ** 5 with jax integer_pow lowering: acc = x; x2 = x*x; x4 = x2*x2; acc*x4.
    """
    q = np.arange(256, dtype=np.uint8).view(np.int8).astype(F32)
    x = q / F32(127)
    x2 = x * x
    x4 = x2 * x2
    return (x * x4) - OFFSET


_LUT = None


def dequantize(codes: np.ndarray) -> np.ndarray:
    global _LUT
    if _LUT is None:
        _LUT = dequant_lut()
    return _LUT[codes.view(np.uint8)]


def block_quantize(leaf: np.ndarray, block_size: int = 16):
    """lion_quant.py:66-80 -> (codes int8[n/bs, bs], inv_scale f32[n/bs, 1]) — the scale is the RECIPROCAL absmax."""
    leaf = leaf.astype(F32, copy=False).reshape(-1, block_size)
    amax = np.max(np.abs(leaf), axis=-1, keepdims=True)
    inv = F32(1) / np.where(amax <= F32(0.0), F32(1), amax).astype(F32)
    return quantize(leaf * inv), inv.astype(F32)


def block_dequantize(shape, codes: np.ndarray, inv: np.ndarray) -> np.ndarray:
    """lion_quant.py:82-92."""
    return (dequantize(codes) / inv).reshape(shape)


def init_state(shape, quantized: bool, block_size: int = 16):
    """lion_quant.py:115-131. Quantised leaves start as quantize(0) == code 3, inv == 1.0."""
    z = np.zeros(shape, F32)
    return block_quantize(z, block_size) if quantized else z


def clip_by_global_norm(grads, max_norm: float = 1.0):
    """optax.clip_by_global_norm [upstream]: g if gn < max else (g / gn) * max. Returns (grads, gn)."""
    gn = np.sqrt(sum(np.sum(np.square(g.astype(F32)), dtype=F32) for g in grads)).astype(F32)
    return [clip_leaf(g, gn, max_norm) for g in grads], gn


def clip_leaf(g: np.ndarray, gn, max_norm: float = 1.0) -> np.ndarray:
    gn = F32(gn)
    if gn < F32(max_norm):
        return g.astype(F32, copy=False)
    return (g.astype(F32) / gn) * F32(max_norm)


def lion_leaf_update(g, p, state, *, b1=0.9, b2=0.99, lr=1e-6 / 7, wd=0.07, wd_flag=True, block_size=16):
    """One leaf of clip-free  lion_8bit.update + apply_updates.

    g: clipped gradient (fp32), p: fp32 parameter, state: (codes, inv) tuple or fp32 momentum.
    Returns (new_p, new_state, scaled_update).  lion_quant.py:133-154, 201-211; optax apply_updates.
    """
    g = g.astype(F32, copy=False)
    p = p.astype(F32, copy=False)
    quantized = isinstance(state, tuple)
    m = block_dequantize(g.shape, *state) if quantized else state.astype(F32, copy=False)
    omb1, fb1 = F32(1.0 - b1), F32(b1)
    omb2, fb2 = F32(1.0 - b2), F32(b2)
    u = np.sign(omb1 * g + fb1 * m)
    m2 = omb2 * g + fb2 * m
    new_state = block_quantize(m2, block_size) if quantized else m2
    upd = u + F32(wd) * p if wd_flag else u
    upd = F32(-1.0 * lr) * upd
    return p + upd, new_state, upd


def ema_update(ema: np.ndarray, p: np.ndarray, ema_rate: float) -> np.ndarray:
    """compute_model_ema (training_utils.py:537-544): `ema_rate * leaf_ema + (1 - ema_rate) * leaf_model` with a python
    float rate — jnp weak typing rounds `ema_rate` and the python-double `1 - ema_rate` to fp32 once each, then two fp32
    products and one fp32 sum, individually rounded."""
    r = np.float32(ema_rate)
    omr = np.float32(1.0 - float(ema_rate))
    a = (r * ema.astype(np.float32)).astype(np.float32)
    b = (omr * p.astype(np.float32)).astype(np.float32)
    return (a + b).astype(np.float32)


def min_snr_weight(alphas_cumprod: np.ndarray, timesteps: np.ndarray, gamma: float, v_prediction: bool) -> np.ndarray:
    """min_snr_gamma_loss_rescale (training_utils.py:546-568), the per-sample weight: snr = ac / (1 - ac) over the
    whole table, gather by timestep, min(snr, gamma) / snr  (or / (snr + 1) for v_prediction). fp32."""
    ac = np.asarray(alphas_cumprod, dtype=np.float32)
    with np.errstate(divide="ignore", invalid="ignore"):
        snrs = (ac / (np.float32(1) - ac).astype(np.float32)).astype(np.float32)
        snr = snrs[np.asarray(timesteps).astype(np.int64)]
        m = np.minimum(snr, np.float32(gamma))
        w = m / (snr + np.float32(1)).astype(np.float32) if v_prediction else m / snr
    return w.astype(np.float32)


def code_magnitude(a: np.ndarray) -> np.ndarray:
    """|code| as a function of a = |y + offset| >= 0 (monotone non-decreasing)."""
    return np.rint(_pow_fifth_root(a.astype(F32)) * F32(127)).astype(np.int32)


def code_thresholds() -> np.ndarray:
    """T[k], k = 0..128 (fp32): |code| = #{k >= 1 : a >= T[k]}; T[0] = 0, T[128] = +inf.

    T[k] is the smallest non-negative fp32 `a` whose code magnitude is >= k, found by bisection over float
    bit patterns (positive floats order like their uint32 patterns).
    """
    thr = np.zeros(129, F32)
    thr[128] = np.inf
    for k in range(1, 128):
        lo, hi = 0, int(np.array(1.0, F32).view(np.uint32))  # code(0) = 0 < k <= 127 = code(1.0)
        while hi - lo > 1:
            mid = (lo + hi) // 2
            a = np.array([mid], np.uint32).view(F32)
            if code_magnitude(a)[0] >= k:
                hi = mid
            else:
                lo = mid
        thr[k] = np.array([hi], np.uint32).view(F32)[0]
    return thr


def write_threshold_include(path: str) -> None:
    """Emit the table the CUDA kernel embeds (checked in; tests compare it with code_thresholds())."""
    thr = code_thresholds()
    bits = thr.view(np.uint32)
    with open(path, "w") as f:
        f.write("// GENERATED by oracle/lion.py:write_threshold_include — do not edit.\n")
        f.write("// T[k] (fp32 bit patterns), k = 0..128: |code| = #{k>=1 : |y+offset| >= T[k]}\n")
        for i in range(0, 129, 6):
            f.write("  " + ", ".join(f"0x{int(b):08x}u" for b in bits[i:i + 6]) + ",\n")


if __name__ == "__main__":
    import os, sys
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(
        os.path.dirname(__file__), "..", "stable_diffusion_training_b200", "csrc", "lion_thresholds.inc")
    write_threshold_include(out)
    t = code_thresholds()
    print("T1=%.10g T64=%.10g T127=%.10g" % (t[1], t[64], t[127]))
