"""GPU parity of the in-step sampling kernels (csrc/rng.cu through stable_diffusion_training_b200/jax_random.py)
against oracle/threefry.py — the restatement of jax.random as /root/reference/training_utils.py:590-624 uses it.
Bits and randint: bit-exact. normal: |gpu - oracle| <= 4 ulp-ish (rel 5e-7 + abs 1e-7): log1p / sqrt of two different
libms may round the last bit differently, everything else is individually rounded fp32 in the same order."""
import numpy as np
import pytest
import torch

from oracle import threefry as T
from stable_diffusion_training_b200 import jax_random as JR

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n", [1, 2, 3, 8, 1001, 131072])
def test_bits_bit_exact(cuda, n):
    key = JR.PRNGKey(5)
    got = JR.random_bits(key, n, cuda).cpu().numpy().view(np.uint32)
    want = T.random_bits(T.PRNGKey(5), n)
    assert np.array_equal(got, want)
    if n >= 8:  # a sub-range is the slice of the global draw
        sub = JR.random_bits(key, n // 4, cuda, n_global=n, begin=n // 2).cpu().numpy().view(np.uint32)
        assert np.array_equal(sub, want[n // 2:n // 2 + n // 4])


@pytest.mark.parametrize("n,lo,hi", [(8, 0, 1000), (64, 0, 1000), (7, 3, 17), (1000, 0, 1)])
def test_randint_bit_exact(cuda, n, lo, hi):
    key = JR.PRNGKey(9)
    got = JR.randint(key, n, lo, hi, cuda).cpu().numpy()
    want = T.randint(T.PRNGKey(9), (n,), lo, hi)
    assert got.dtype == np.int32 and np.array_equal(got, want)
    if n >= 8:
        sub = JR.randint(key, n // 2, lo, hi, cuda, n_global=n, begin=n // 4).cpu().numpy()
        assert np.array_equal(sub, want[n // 4:n // 4 + n // 2])


def _close(a, b):
    return np.all(np.abs(a.astype(np.float64) - b.astype(np.float64)) <= 5e-7 * np.abs(b) + 1e-7)


@pytest.mark.parametrize("shape,off,pert", [((8, 4, 64, 64), 0.0, 0.0), ((2, 4, 32, 32), 0.1, 0.0),
                                            ((3, 4, 8, 8), 0.05, 0.1), ((1, 4, 3, 3), 0.0, 0.2)])
def test_sample_noise_matches_oracle(cuda, shape, off, pert):
    rng = JR.PRNGKey(123)
    keys = JR.train_step_keys(rng)
    got = JR.sample_noise(keys["noise"], shape, cuda, offset_key=keys["offset"], offset_noise_magnitude=off,
                          perturb_key=keys["perturb"], perturbation_noise_magnitude=pert).cpu().numpy()
    want, want_t, new = T.sample_noise_and_timesteps(T.PRNGKey(123), shape, 1000, off, pert)
    assert _close(got, want), float(np.abs(got - want).max())
    assert np.mean(got.view(np.uint32) == want.view(np.uint32)) > 0.9  # nearly all values are bit-identical
    t = JR.randint(keys["timestep"], shape[0], 0, 1000, cuda).cpu().numpy()
    assert np.array_equal(t, want_t)
    assert list(keys["new_train"]) == new.tolist()
    # data-parallel rows: rank r of 2 draws rows [r*B/2, (r+1)*B/2) of the same global array
    if shape[0] % 2 == 0:
        h = shape[0] // 2
        for r in range(2):
            part = JR.sample_noise(keys["noise"], (h,) + shape[1:], cuda, offset_key=keys["offset"],
                                   offset_noise_magnitude=off, perturb_key=keys["perturb"],
                                   perturbation_noise_magnitude=pert, batch_global=shape[0], batch_begin=r * h)
            assert torch.equal(part.cpu(), torch.from_numpy(got[r * h:(r + 1) * h]))
