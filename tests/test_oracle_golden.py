"""Pin the oracle (oracle/lion.py, oracle/sched.py) against tests/golden/*.npz — the outputs of the UNMODIFIED reference
files /root/reference/lion_quant.py and /root/reference/schedulers/scheduling_{utils,ddpm}_flax.py executed under
oracle/jax_shim.py by oracle/make_golden.py. Bit-exact (fp32 bit patterns, int8 codes). CPU only."""
import os

import numpy as np
import pytest

from oracle import lion as OL
from oracle import sched as OS

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _bits(a):
    a = np.ascontiguousarray(np.asarray(a))
    return a.view(np.uint32) if a.dtype == np.float32 else a


@pytest.fixture(scope="module")
def lion():
    return np.load(os.path.join(G, "lion_ref.npz"))


@pytest.fixture(scope="module")
def sched():
    return np.load(os.path.join(G, "sched_ref.npz"))


def _state(z, tag, n):
    if f"{tag}/codes/{n}" in z:
        return (z[f"{tag}/codes/{n}"], z[f"{tag}/inv/{n}"])
    return z[f"{tag}/mu/{n}"]


def test_lion_init_state(lion):
    for n in lion["names"]:
        st = _state(lion, "s0", n)
        if isinstance(st, tuple):
            assert st[0].dtype == np.int8 and (st[0] == 3).all() and (st[1] == 1.0).all()
            want = OL.init_state((st[0].size,), True)
            assert np.array_equal(want[0], st[0]) and np.array_equal(_bits(want[1]), _bits(st[1]))
        else:
            assert (st == 0).all()
    assert int(lion["s0/count"]) == 0


def test_lion_chain_steps_bit_exact(lion):
    lr = float(lion["lr"])
    for step in range(1, int(lion["nsteps"]) + 1):
        for n in lion["names"]:
            g = lion[f"g{step}/{n}"]
            p = lion[f"p{step - 1}/{n}"]
            st = _state(lion, f"s{step - 1}", n)
            quant = isinstance(st, tuple)
            st_in = (st[0], st[1]) if quant else st.reshape(-1)
            new_p, new_s, upd = OL.lion_leaf_update(g.reshape(-1), p.reshape(-1), st_in, lr=lr, wd=0.07,
                                                    wd_flag=n.endswith("kernel"))
            assert np.array_equal(_bits(upd), _bits(lion[f"u{step}/{n}"].reshape(-1))), (step, n)
            assert np.array_equal(_bits(new_p), _bits(lion[f"p{step}/{n}"].reshape(-1))), (step, n)
            want = _state(lion, f"s{step}", n)
            if quant:
                assert np.array_equal(new_s[0], want[0]), (step, n)
                assert np.array_equal(_bits(new_s[1]), _bits(want[1])), (step, n)
            else:
                assert np.array_equal(_bits(new_s), _bits(want.reshape(-1))), (step, n)
        assert int(lion[f"s{step}/count"]) == step


def test_scale_by_lion_alone(lion):
    g = lion["sb/g"]
    st = OL.init_state((256,), True)
    # update = sign(c): run the oracle leaf with lr = -1, no decay
    _, st1, u1 = OL.lion_leaf_update(g, np.zeros_like(g), st, lr=-1.0, wd=0.0, wd_flag=False)
    assert np.array_equal(u1, lion["sb/u1"]) and np.array_equal(st1[0], lion["sb/codes1"])
    assert np.array_equal(_bits(st1[1]), _bits(lion["sb/inv1"]))
    _, st2, u2 = OL.lion_leaf_update((-0.5 * g).astype(np.float32), np.zeros_like(g), st1, lr=-1.0, wd=0.0,
                                     wd_flag=False)
    assert np.array_equal(u2, lion["sb/u2"]) and np.array_equal(st2[0], lion["sb/codes2"])
    assert np.array_equal(_bits(st2[1]), _bits(lion["sb/inv2"]))


def test_clip_by_global_norm(lion):
    for tag in ("clip_hi", "clip_lo"):
        gn = lion[f"{tag}/gn"]
        out, gn2 = OL.clip_by_global_norm([lion[f"{tag}/a"], lion[f"{tag}/b"]], 1.0)
        assert abs(float(gn2) - float(gn)) <= 2e-7 * float(gn)
        for got, key, src in ((out[0], "oa", "a"), (out[1], "ob", "b")):
            # given the reference's own norm the clipped leaves are bit-exact
            assert np.array_equal(_bits(OL.clip_leaf(lion[f"{tag}/{src}"], gn, 1.0)), _bits(lion[f"{tag}/{key}"]))
    assert float(lion["clip_hi/gn"]) > 1 > float(lion["clip_lo/gn"])


@pytest.mark.parametrize("name", ["linear", "scaled_linear", "zero_snr_scaled_linear", "squaredcos_cap_v2"])
def test_beta_schedules(sched, name):
    st = OS.common_state(name, 0.00085, 0.012, 1000)
    for k in ("betas", "alphas", "alphas_cumprod"):
        assert np.array_equal(_bits(st[k]), _bits(sched[f"{name}/{k}"])), (name, k)
    assert np.array_equal(sched[f"{name}/timesteps"], np.arange(1000)[::-1])
    assert float(sched[f"{name}/init_noise_sigma"]) == 1.0


def test_known_answers_from_survey(sched):
    ac = sched["scaled_linear/alphas_cumprod"]
    assert ac[0] == np.float32(0.99915) and abs(ac[500] - 0.27633256) < 1e-6 and abs(ac[999] - 0.0046600895) < 1e-8
    z = sched["zero_snr_scaled_linear/alphas_cumprod"]
    assert z[999] == 0.0 and sched["zero_snr_scaled_linear/betas"][999] == 1.0 and abs(z[500] - 0.24101873) < 1e-6


@pytest.mark.parametrize("name", ["scaled_linear", "zero_snr_scaled_linear"])
def test_add_noise_get_velocity(sched, name):
    ac = sched[f"{name}/alphas_cumprod"]
    x0, nz, t = sched["x0"], sched["noise"], sched["t"]
    assert np.array_equal(_bits(OS.add_noise(ac, x0, nz, t)), _bits(sched[f"{name}/noisy"]))
    assert np.array_equal(_bits(OS.get_velocity(ac, x0, nz, t)), _bits(sched[f"{name}/velocity"]))


@pytest.mark.parametrize("ptype", ["epsilon", "sample", "v_prediction"])
@pytest.mark.parametrize("vt", ["fixed_small", "fixed_large"])
def test_ddpm_step(sched, ptype, vt):
    st = OS.common_state("scaled_linear", 0.00085, 0.012, 1000)
    mo, smp, vn = sched["step/model_output"], sched["step/sample"], sched["step/noise"]
    for t in (0, 500, 999):
        got = OS.ddpm_step(st, mo, t, smp, noise=vn, prediction_type=ptype, clip_sample=(ptype != "sample"),
                           variance_type=vt)
        assert np.array_equal(_bits(got), _bits(sched[f"step/{ptype}/{vt}/{t}"])), (ptype, vt, t)
        assert np.array_equal(_bits(np.float32(OS.get_variance(st, t, vt))), _bits(sched[f"var/{vt}/{t}"]))


def test_set_timesteps(sched):
    assert np.array_equal(sched["set_timesteps_50"], (np.arange(50) * 20)[::-1])
    assert int(sched["len"]) == 1000


@pytest.mark.parametrize("ptype", ["epsilon", "sample", "v_prediction"])
@pytest.mark.parametrize("one", [True, False])
def test_ddim_step_bit_exact(ptype, one):
    """oracle/sched.py:ddim_step == the UNMODIFIED schedulers/scheduling_ddim_flax.py:199-284 run under the shim
    (tests/golden/ddim_ref.npz), bit for bit: 3 prediction types, set_alpha_to_one on / off, eta 0 and 0.5, first / middle /
    last steps of a 50-step schedule with steps_offset 1."""
    z = np.load(os.path.join(G, "ddim_ref.npz"))
    st = OS.common_state("scaled_linear", 0.00085, 0.012)
    final = np.float32(1.0) if one else st["alphas_cumprod"][0]
    assert z[f"final/{int(one)}"] == final
    assert np.array_equal(z[f"timesteps/{int(one)}"], (np.arange(50) * 20)[::-1] + 1)
    for ti in (981, 501, 21, 1):
        assert _bits(OS.ddim_variance(st, ti, ti - 20, final)) == _bits(z[f"var/{int(one)}/{ti}"])
        for eta in (0.0, 0.5):
            got = OS.ddim_step(st, z["model_output"], ti, z["sample"], 50, prediction_type=ptype, eta=eta,
                               final_alpha_cumprod=final)
            want = z[f"step/{ptype}/{int(one)}/{ti}/{eta}"]
            assert np.array_equal(_bits(got), _bits(want)), (ptype, one, ti, eta)
