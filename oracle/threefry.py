"""ORACLE (test infrastructure only — never imported by the product path).

NumPy restatement of the jax.random pieces the reference's train_step uses for in-step sampling
(/root/reference/training_utils.py:530 `jax.random.split(train_rng, num=3)`, 590-592 `split(sample_rng, num=4)`,
593-615 `jax.random.normal`, 619-624 `jax.random.randint`). jax is a third-party dependency that is absent here
(requirements: unpinned; `jax.random.KeyArray` at schedulers/scheduling_ddpm_flax.py:202 bounds it to <= 0.4.23), so the
published algorithm of that era is restated: default PRNG implementation threefry2x32, `jax_threefry_partitionable`
off.

  threefry2x32            Threefry-2x32, 20 rounds (Salmon et al., Random123), rotations (13,15,26,6 | 17,29,16,24),
                          key schedule constant 0x1BD11BDA — exactly jax._src.prng.threefry2x32_p
  threefry_2x32(key, c)   jax's array form: odd-length counts are zero-padded, the array is split in two halves that
                          form the two words of each block, outputs are concatenated again
  split(key, num)         threefry_2x32(key, iota(2*num)).reshape(num, 2)
  random_bits(key, n)     threefry_2x32(key, iota(n)) (32-bit)
  uniform / normal        mantissa trick ((bits >> 9) | 0x3f800000) - 1, scaled to [nextafter(-1,0), 1); normal =
                          sqrt(2) * erf_inv(u) with XLA's single-precision erf_inv (Giles' polynomial)
  randint                 two draws of bits, (hi % span * ((2^16 % span)^2 % span) + lo % span) % span

Pins: the Random123 known-answer vectors for Threefry-2x32-20 (also asserted by jax's own test-suite) and the key /
sample values printed in jax's public documentation (tests/test_threefry.py). The integer paths are bit-exact by
construction; `normal` goes through log1p / sqrt whose last-bit rounding on a TPU is not knowable here, so it is
compared with a few-ulp tolerance and that is stated where it is used.
"""
from __future__ import annotations

import numpy as np

_ROT = ((13, 15, 26, 6), (17, 29, 16, 24))
U32 = np.uint32


def _rotl(x, r):
    return (x << U32(r)) | (x >> U32(32 - r))


def threefry2x32(k0, k1, x0, x1):
    """One Threefry-2x32-20 block per element. All inputs uint32 arrays / scalars; returns (y0, y1)."""
    with np.errstate(over="ignore"):
        k0, k1 = U32(k0), U32(k1)
        x0 = np.asarray(x0, dtype=U32).copy()
        x1 = np.asarray(x1, dtype=U32).copy()
        ks = (k0, k1, U32(k0 ^ k1 ^ U32(0x1BD11BDA)))
        x0 += ks[0]
        x1 += ks[1]
        for i in range(5):
            for r in _ROT[i % 2]:
                x0 += x1
                x1 = _rotl(x1, r)
                x1 ^= x0
            x0 += ks[(i + 1) % 3]
            x1 += ks[(i + 2) % 3] + U32(i + 1)
    return x0, x1


def threefry_2x32(key, count):
    """jax._src.prng.threefry_2x32: key (k0, k1), count uint32 [n] -> uint32 [n]."""
    count = np.asarray(count, dtype=U32).ravel()
    n = count.size
    if n % 2:
        count = np.concatenate([count, np.zeros(1, U32)])
    half = count.size // 2
    y0, y1 = threefry2x32(key[0], key[1], count[:half], count[half:])
    return np.concatenate([y0, y1])[:n]


def PRNGKey(seed: int):
    seed = int(seed)
    return np.array([(seed >> 32) & 0xFFFFFFFF, seed & 0xFFFFFFFF], dtype=U32)


def split(key, num: int = 2):
    return threefry_2x32(key, np.arange(2 * num, dtype=U32)).reshape(num, 2)


def random_bits(key, n: int):
    return threefry_2x32(key, np.arange(n, dtype=U32))


def uniform_pm1(bits):
    """jax.random.uniform(minval=nextafter(-1, 0), maxval=1) from raw bits, fp32, individually rounded ops."""
    lo = np.nextafter(np.float32(-1.0), np.float32(0.0))
    f = ((bits >> U32(9)) | U32(0x3F800000)).view(np.float32) - np.float32(1.0)
    scale = np.float32(np.float32(1.0) - lo)  # == 2.0f (ties-to-even)
    return np.maximum(lo, (f * scale).astype(np.float32) + lo).astype(np.float32)


_ALPHA = np.array([2.81022636e-08, 3.43273939e-07, -3.5233877e-06, -4.39150654e-06, 0.00021858087, -0.00125372503,
                   -0.00417768164, 0.246640727, 1.50140941], dtype=np.float32)
_BETA = np.array([-0.000200214257, 0.000100950558, 0.00134934322, -0.00367342844, 0.00573950773, -0.0076224613,
                  0.00943887047, 1.00167406, 2.83297682], dtype=np.float32)


def erf_inv_f32(x):
    """XLA ErfInv for F32 (xla/client/lib/math.cc: Giles, "Approximating the erfinv function"): w = -log1p(-x*x);
    two degree-8 polynomials in (w - 2.5) / (sqrt(w) - 3) selected at w < 5; result p * x; +-inf at |x| == 1."""
    x = np.asarray(x, dtype=np.float32)
    with np.errstate(divide="ignore"):
        w = -np.log1p(-(x * x).astype(np.float32)).astype(np.float32)
    lt = w < np.float32(5.0)
    with np.errstate(invalid="ignore"):
        wv = np.where(lt, w - np.float32(2.5), np.sqrt(w).astype(np.float32) - np.float32(3.0)).astype(np.float32)
    p = np.where(lt, _ALPHA[0], _BETA[0]).astype(np.float32)
    for i in range(1, 9):
        p = (np.where(lt, _ALPHA[i], _BETA[i]).astype(np.float32) + (p * wv).astype(np.float32)).astype(np.float32)
    out = (p * x).astype(np.float32)
    return np.where(np.abs(x) == np.float32(1.0), np.float32(np.inf) * x, out).astype(np.float32)


def normal(key, shape):
    n = int(np.prod(shape)) if len(shape) else 1
    u = uniform_pm1(random_bits(key, n))
    return (np.float32(np.sqrt(2)) * erf_inv_f32(u)).astype(np.float32).reshape(shape)


def randint(key, shape, minval: int, maxval: int):
    """jax.random.randint for int32 (jax 0.4.x): unbiased-enough double-draw modulo construction."""
    n = int(np.prod(shape)) if len(shape) else 1
    k1, k2 = split(key, 2)
    hi, lo = random_bits(k1, n), random_bits(k2, n)
    span = U32(maxval - minval) if maxval > minval else U32(1)
    with np.errstate(over="ignore"):
        mult = U32(U32(1 << 16) % span)
        mult = U32((mult * mult) % span)
        off = ((hi % span) * mult + (lo % span)) % span
    return (np.int32(minval) + off.astype(np.int32)).reshape(shape)


def train_step_keys(train_rng):
    """training_utils.py:530, 590-592: returns dict(dropout, sample, new_train, offset, noise, perturb, timestep)."""
    dropout, sample, new = split(train_rng, 3)
    off, noise, perturb, tstep = split(sample, 4)
    return {"dropout": dropout, "sample": sample, "new_train": new, "offset": off, "noise": noise, "perturb": perturb,
            "timestep": tstep}


def sample_noise_and_timesteps(train_rng, shape, num_train_timesteps=1000, offset_noise_magnitude=0.0,
                               perturbation_noise_magnitude=0.0):
    """training_utils.py:590-624: (noise f32 [B,C,H,W], timesteps int32 [B], new_train_rng)."""
    k = train_step_keys(train_rng)
    noise = normal(k["noise"], shape)
    if offset_noise_magnitude:
        noise = (noise + (normal(k["offset"], (shape[0], shape[1], 1, 1)) * np.float32(offset_noise_magnitude))
                 .astype(np.float32)).astype(np.float32)
    if perturbation_noise_magnitude:
        noise = (noise + (np.float32(perturbation_noise_magnitude) * normal(k["perturb"], shape)).astype(np.float32)
                 ).astype(np.float32)
    t = randint(k["timestep"], (shape[0],), 0, num_train_timesteps)
    return noise, t, k["new_train"]
