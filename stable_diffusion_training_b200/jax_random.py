"""jax.random as the reference's train step uses it (/root/reference/training_utils.py:530, 590-624), B200 side.

  PRNGKey(seed)                 uint32[2] key, (hi, lo) words of the seed
  split(key, num)               jax.random.split — Threefry-2x32-20 over iota(2*num); a handful of integer blocks, done
                                on the host in Python ints (no device work, no sync)
  train_step_keys(train_rng)    the step's key tree: split(train_rng, 3) -> dropout, sample, new_train_rng (530);
                                split(sample_rng, 4) -> offset, noise, perturb, timestep keys (590-592)
  sample_noise(...)             noise = normal(noise_key, shape) [+ normal(offset_key, (B,C,1,1)) * m_o]
                                [+ m_p * normal(perturb_key, shape)] (593-615), one kernel (csrc/rng.cu)
  randint(...)                  jax.random.randint(timestep_key, (B,), 0, T) (619-624), one kernel
  RngStream(seed | key)         mutable holder of `train_rng`: train_step replaces `.key` by `new_train_rng`

The draws are the ones jax 0.4.x makes with its default threefry2x32 implementation (restated and pinned in
oracle/threefry.py / tests/test_threefry.py): bits and integers bit-exact, normals through XLA's F32 erf_inv
polynomial. `batch_global` / `batch_begin` let a data-parallel rank draw only its rows of the global batch, so N ranks
see exactly the samples a single process would draw for the concatenated batch (the reference shards one global draw:
training_utils.py:835-840).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import torch

from . import _lib

_M32 = 0xFFFFFFFF
_ROT = ((13, 15, 26, 6), (17, 29, 16, 24))

Key = Tuple[int, int]


def _rotl(x: int, r: int) -> int:
    return ((x << r) | (x >> (32 - r))) & _M32


def _threefry2x32(key: Key, x0: int, x1: int) -> Tuple[int, int]:
    k0, k1 = key
    ks = (k0, k1, k0 ^ k1 ^ 0x1BD11BDA)
    x0 = (x0 + ks[0]) & _M32
    x1 = (x1 + ks[1]) & _M32
    for i in range(5):
        for r in _ROT[i % 2]:
            x0 = (x0 + x1) & _M32
            x1 = _rotl(x1, r) ^ x0
        x0 = (x0 + ks[(i + 1) % 3]) & _M32
        x1 = (x1 + ks[(i + 2) % 3] + i + 1) & _M32
    return x0, x1


def as_key(key) -> Key:
    """Accepts (hi, lo) tuples, uint32[2] numpy arrays / torch tensors, or an int seed."""
    if isinstance(key, RngStream):
        return key.key
    if isinstance(key, int):
        return PRNGKey(key)
    if isinstance(key, torch.Tensor):
        key = key.detach().cpu().tolist()
    k = [int(v) & _M32 for v in key]
    if len(k) != 2:
        raise ValueError(f"a PRNG key has two uint32 words, got {len(k)}")
    return k[0], k[1]


def PRNGKey(seed: int) -> Key:
    seed = int(seed)
    return (seed >> 32) & _M32, seed & _M32


def split(key, num: int = 2) -> Sequence[Key]:
    """jax.random.split: threefry_2x32(key, iota(2 * num)).reshape(num, 2) — counts split in two halves."""
    key = as_key(key)
    n = 2 * num
    half = n // 2
    out = [0] * n
    for i in range(half):
        out[i], out[i + half] = _threefry2x32(key, i, i + half)
    return [(out[2 * i], out[2 * i + 1]) for i in range(num)]


def train_step_keys(train_rng) -> dict:
    dropout, sample, new = split(train_rng, 3)
    off, noise, perturb, tstep = split(sample, 4)
    return {"dropout": dropout, "sample": sample, "new_train": new, "offset": off, "noise": noise, "perturb": perturb,
            "timestep": tstep}


class RngStream:
    """The training loop's `train_rng`, advanced in place by every train_step that draws from it (the reference
    threads `new_train_rng` through its loop: training.py:214-234, training_utils.py:530, 755-762)."""

    def __init__(self, seed_or_key=0):
        self.key: Key = as_key(seed_or_key)

    def __repr__(self):
        return f"RngStream(key=({self.key[0]}, {self.key[1]}))"


def _key_arr(key: Optional[Key]):
    if key is None:
        return None
    return (C.c_uint32 * 2)(key[0], key[1])


def sample_noise(noise_key, shape, device, *, offset_key=None, offset_noise_magnitude: float = 0.0, perturb_key=None,
                 perturbation_noise_magnitude: float = 0.0, batch_global: Optional[int] = None, batch_begin: int = 0,
                 out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """fp32 [B_local, C, H, W] noise as the reference draws it (rows batch_begin.. of a batch_global-row draw)."""
    B, Cc = int(shape[0]), int(shape[1])
    hw = 1
    for d in shape[2:]:
        hw *= int(d)
    if out is None:
        out = torch.empty(tuple(shape), dtype=torch.float32, device=device)
    kn, ko, kp = _key_arr(as_key(noise_key)), None, None
    if offset_noise_magnitude:
        ko = _key_arr(as_key(offset_key))
    if perturbation_noise_magnitude:
        kp = _key_arr(as_key(perturb_key))
    _lib.check(_lib.lib().sdb_sample_noise(kn, ko, kp, float(offset_noise_magnitude), float(perturbation_noise_magnitude),
                                           out.data_ptr(), int(batch_global or B), int(batch_begin), B, Cc, hw,
                                           _lib.stream_ptr(out.device)), "sdb_sample_noise")
    return out


def randint(key, n: int, minval: int, maxval: int, device, *, n_global: Optional[int] = None, begin: int = 0,
            out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """int32 [n] = jax.random.randint(key, (n_global,), minval, maxval)[begin : begin + n]."""
    k = as_key(key)
    if out is None:
        out = torch.empty((n,), dtype=torch.int32, device=device)
    _lib.check(_lib.lib().sdb_threefry_randint(k[0], k[1], out.data_ptr(), int(n_global or n), int(begin), int(n),
                                               int(minval), int(maxval), _lib.stream_ptr(out.device)),
               "sdb_threefry_randint")
    return out


def random_bits(key, n: int, device, *, n_global: Optional[int] = None, begin: int = 0) -> torch.Tensor:
    """uint32 bits (returned as int32 storage) = threefry_2x32(key, iota(n_global))[begin : begin + n]."""
    k = as_key(key)
    out = torch.empty((n,), dtype=torch.int32, device=device)
    _lib.check(_lib.lib().sdb_threefry_bits(k[0], k[1], out.data_ptr(), int(n_global or n), int(begin), int(n),
                                            _lib.stream_ptr(out.device)), "sdb_threefry_bits")
    return out
