"""CPU pins of the jax.random restatement (oracle/threefry.py) and of the host-side key arithmetic of the product
(stable_diffusion_training_b200/jax_random.py) that replaces /root/reference/training_utils.py:530, 590-592.

jax is an absent third-party dependency (no wheel, no network), so the pins are published values:
  * the Random123 known-answer vectors of Threefry-2x32-20 (Salmon et al.; the same three vectors jax's own
    tests/random_test.py::testThreefry2x32 asserts);
  * the keys and samples printed in jax's public documentation ("JAX - The Sharp Bits", "Pseudo Random Numbers in JAX")
    for the default threefry2x32 implementation: split(PRNGKey(0)), the second split, normal(PRNGKey(0), (1,)),
    normal of the two sub-keys, normal(PRNGKey(42)).
"""
import numpy as np

from oracle import threefry as T
from stable_diffusion_training_b200 import jax_random as JR


def test_random123_known_answers():
    kat = [((0x0, 0x0), (0x0, 0x0), (0x6B200159, 0x99BA4EFE)),
           ((0xFFFFFFFF, 0xFFFFFFFF), (0xFFFFFFFF, 0xFFFFFFFF), (0x1CB996FC, 0xBB002BE7)),
           ((0x13198A2E, 0x03707344), (0x243F6A88, 0x85A308D3), (0xC4923A9C, 0x483DF7A0))]
    for key, ctr, want in kat:
        y0, y1 = T.threefry2x32(key[0], key[1], [ctr[0]], [ctr[1]])
        assert (int(y0[0]), int(y1[0])) == want
        assert JR._threefry2x32(key, ctr[0], ctr[1]) == want


def test_documented_keys_and_samples():
    k = T.PRNGKey(0)
    assert k.tolist() == [0, 0]
    s = T.split(k)
    assert s.tolist() == [[4146024105, 967050713], [2718843009, 1272950319]]
    assert T.normal(k, (1,))[0] == np.float32(-0.20584226)
    assert T.normal(s[1], (1,))[0] == np.float32(-1.2515389)
    s2 = T.split(s[0])
    assert s2.tolist() == [[2384771982, 3928867769], [1278412471, 2182328957]]
    assert T.normal(s2[1], (1,))[0] == np.float32(-0.58665055)
    assert T.PRNGKey(42).tolist() == [0, 42]
    assert T.normal(T.PRNGKey(42), ())[()] == np.float32(-0.18471177)


def test_host_split_equals_oracle():
    for seed in (0, 1, 42, 2 ** 32 + 5, 2 ** 63 - 1):
        assert list(JR.PRNGKey(seed)) == T.PRNGKey(seed).tolist()
        for num in (2, 3, 4, 7):
            got = [list(k) for k in JR.split(JR.PRNGKey(seed), num)]
            assert got == T.split(T.PRNGKey(seed), num).tolist()
    a = JR.train_step_keys(JR.PRNGKey(7))
    b = T.train_step_keys(T.PRNGKey(7))
    assert {k: list(v) for k, v in a.items()} == {k: v.tolist() for k, v in b.items()}
    # an RngStream advances to new_train_rng
    st = JR.RngStream(7)
    assert st.key == JR.PRNGKey(7)


def test_odd_lengths_pad_like_jax():
    """threefry_2x32 zero-pads an odd count array: element i of an n-draw is NOT a prefix of an (n+1)-draw."""
    k = T.PRNGKey(3)
    b3, b4 = T.random_bits(k, 3), T.random_bits(k, 4)
    assert b3[0] == b4[0] and b3[1] != b4[1]
    # the pad block is (count[1], 0)
    y0, y1 = T.threefry2x32(k[0], k[1], [0, 1], [2, 0])
    assert b3.tolist() == [int(y0[0]), int(y0[1]), int(y1[0])]


def test_uniform_range_and_randint():
    k = T.PRNGKey(11)
    u = T.uniform_pm1(T.random_bits(k, 100000))
    assert u.min() >= np.nextafter(np.float32(-1), np.float32(0)) and u.max() < 1.0
    z = T.normal(k, (100000,))
    assert abs(float(z.mean())) < 0.02 and abs(float(z.std()) - 1.0) < 0.02 and np.isfinite(z).all()
    t = T.randint(k, (100000,), 0, 1000)
    assert t.min() == 0 and t.max() == 999 and t.dtype == np.int32
    assert abs(float(t.mean()) - 499.5) < 5
