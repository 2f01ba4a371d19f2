"""The CUDA path, called through the reference's own API surface (lion_8bit(...).init/.update + apply_updates,
FlaxDDPMScheduler.create_state / add_noise / get_velocity / step), against tests/golden/*.npz — outputs of the
UNMODIFIED reference files run under oracle/jax_shim.py (oracle/make_golden.py). Bit-exact: int8 codes, fp32
reciprocal-absmax scales, fp32 updates / parameters / samples compared as bit patterns."""
import os

import numpy as np
import pytest
import torch

from stable_diffusion_training_b200 import lion_quant as LQ
from stable_diffusion_training_b200.schedulers import FlaxDDPMScheduler

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _bits(a):
    a = np.ascontiguousarray(np.asarray(a))
    return a.view(np.uint32) if a.dtype == np.float32 else a


def _np(t):
    return t.detach().cpu().numpy()


def test_lion_8bit_api_matches_reference_run(cuda):
    z = np.load(os.path.join(G, "lion_ref.npz"))
    names = [str(n) for n in z["names"]]
    dev = cuda
    params = LQ.unflatten_dict({n: torch.from_numpy(z[f"p0/{n}"]).to(dev) for n in names})
    qmask = LQ.unflatten_dict({n: n.endswith("kernel") and "time_emb_proj" not in n for n in names})
    wmask = LQ.unflatten_dict({n: n.endswith("kernel") for n in names})
    tx = LQ.lion_8bit(LQ.constant_schedule(float(z["lr"])), b1=0.9, b2=0.99, block_size=16, weight_decay=1e-2 * 7,
                      mask=wmask, excluded_layer_mask=qmask)
    state = tx.init(params)
    for n, m in LQ._flatten_state(state.lion.mu_quant).items():
        if isinstance(m, tuple):
            assert np.array_equal(_np(m[0]), z[f"s0/codes/{n}"]) and np.array_equal(_np(m[1]), z[f"s0/inv/{n}"])
        else:
            assert np.array_equal(_np(m), z[f"s0/mu/{n}"])
    for step in range(1, int(z["nsteps"]) + 1):
        grads = LQ.unflatten_dict({n: torch.from_numpy(z[f"g{step}/{n}"]).to(dev) for n in names})
        updates, state = tx.update(grads, state, params)
        params = LQ.apply_updates(params, updates)
        torch.cuda.synchronize()
        fu, fp = LQ.flatten_dict(updates), LQ.flatten_dict(params)
        fm = LQ._flatten_state(state.lion.mu_quant)
        for n in names:
            assert np.array_equal(_bits(_np(fu[n])), _bits(z[f"u{step}/{n}"])), (step, n, "update")
            assert np.array_equal(_bits(_np(fp[n])), _bits(z[f"p{step}/{n}"])), (step, n, "param")
            if isinstance(fm[n], tuple):
                assert fm[n][0].dtype == torch.int8
                assert np.array_equal(_np(fm[n][0]), z[f"s{step}/codes/{n}"]), (step, n, "codes")
                assert np.array_equal(_bits(_np(fm[n][1])), _bits(z[f"s{step}/inv/{n}"])), (step, n, "scales")
            else:
                assert np.array_equal(_bits(_np(fm[n])), _bits(z[f"s{step}/mu/{n}"])), (step, n, "mu")
        assert int(state.lion.count) == step == int(z[f"s{step}/count"])


def test_scale_by_lion_8bit_alone(cuda):
    z = np.load(os.path.join(G, "lion_ref.npz"))
    tx = LQ.scale_by_lion_8bit(b1=0.9, b2=0.99, block_size=16, excluded_layer_mask={"w": True})
    st = tx.init({"w": torch.zeros(256, device=cuda)})
    g = torch.from_numpy(z["sb/g"]).to(cuda)
    u1, st = tx.update({"w": g}, st)
    assert np.array_equal(_np(u1["w"]), z["sb/u1"]) and np.array_equal(_np(st.mu_quant["w"][0]), z["sb/codes1"])
    assert np.array_equal(_bits(_np(st.mu_quant["w"][1])), _bits(z["sb/inv1"]))
    u2, st = tx.update({"w": torch.from_numpy((-0.5 * z["sb/g"]).astype(np.float32)).to(cuda)}, st)
    assert np.array_equal(_np(u2["w"]), z["sb/u2"]) and np.array_equal(_np(st.mu_quant["w"][0]), z["sb/codes2"])
    assert np.array_equal(_bits(_np(st.mu_quant["w"][1])), _bits(z["sb/inv2"]))


def test_clip_by_global_norm(cuda):
    z = np.load(os.path.join(G, "lion_ref.npz"))
    clip = LQ.clip_by_global_norm(1)
    for tag in ("clip_hi", "clip_lo"):
        gr = {"a": torch.from_numpy(z[f"{tag}/a"]).to(cuda), "b": torch.from_numpy(z[f"{tag}/b"]).to(cuda)}
        gn = float(LQ.global_norm(gr))
        assert abs(gn - float(z[f"{tag}/gn"])) <= 2e-7 * gn
        out, _ = clip.update(gr, clip.init(gr))
        for k in ("a", "b"):
            # the norm is accumulated in fp64 on the GPU, fp32 in the reference: allow the 1-ulp-of-norm difference
            np.testing.assert_allclose(_np(out[k]), z[f"{tag}/o{k}"], rtol=3e-7, atol=0)


@pytest.mark.parametrize("name", ["linear", "scaled_linear", "zero_snr_scaled_linear", "squaredcos_cap_v2"])
def test_scheduler_state(cuda, name):
    z = np.load(os.path.join(G, "sched_ref.npz"))
    s = FlaxDDPMScheduler(beta_start=0.00085, beta_end=0.012, beta_schedule=name, num_train_timesteps=1000, device=cuda)
    st = s.create_state()
    for k in ("betas", "alphas", "alphas_cumprod"):
        assert np.array_equal(_bits(_np(getattr(st.common, k))), _bits(z[f"{name}/{k}"])), (name, k)
    assert np.array_equal(_np(st.timesteps), z[f"{name}/timesteps"])
    assert float(st.init_noise_sigma) == 1.0 and len(s) == int(z["len"])
    assert np.array_equal(_np(s.set_timesteps(st, 50).timesteps), z["set_timesteps_50"])


@pytest.mark.parametrize("name", ["scaled_linear", "zero_snr_scaled_linear"])
def test_add_noise_get_velocity(cuda, name):
    z = np.load(os.path.join(G, "sched_ref.npz"))
    s = FlaxDDPMScheduler(beta_start=0.00085, beta_end=0.012, beta_schedule=name, num_train_timesteps=1000, device=cuda)
    st = s.create_state()
    x0, nz, t = (torch.from_numpy(z[k]).to(cuda) for k in ("x0", "noise", "t"))
    assert np.array_equal(_bits(_np(s.add_noise(st, x0, nz, t))), _bits(z[f"{name}/noisy"]))
    assert np.array_equal(_bits(_np(s.get_velocity(st, x0, nz, t))), _bits(z[f"{name}/velocity"]))


@pytest.mark.parametrize("ptype", ["epsilon", "sample", "v_prediction"])
@pytest.mark.parametrize("vt", ["fixed_small", "fixed_large"])
def test_ddpm_step(cuda, ptype, vt):
    z = np.load(os.path.join(G, "sched_ref.npz"))
    s = FlaxDDPMScheduler(beta_start=0.00085, beta_end=0.012, beta_schedule="scaled_linear", num_train_timesteps=1000,
                          prediction_type=ptype, variance_type=vt, clip_sample=(ptype != "sample"), device=cuda)
    st = s.create_state()
    mo, smp, vn = (torch.from_numpy(z[k]).to(cuda) for k in ("step/model_output", "step/sample", "step/noise"))
    for t in (0, 500, 999):
        out = s.step(st, mo, t, smp, key=vn)
        assert np.array_equal(_bits(_np(out.prev_sample)), _bits(z[f"step/{ptype}/{vt}/{t}"])), (ptype, vt, t)
        assert np.array_equal(_bits(np.float32(s._get_variance(st, t))), _bits(z[f"var/{vt}/{t}"]))


@pytest.mark.parametrize("ptype", ["epsilon", "sample", "v_prediction"])
@pytest.mark.parametrize("one", [True, False])
def test_ddim_scheduler_matches_reference_run(cuda, ptype, one):
    """FlaxDDIMScheduler (sdb_ddim_step) == tests/golden/ddim_ref.npz, the output of the unmodified
    schedulers/scheduling_ddim_flax.py: create_state / set_timesteps / _get_variance / step, bit for bit."""
    from stable_diffusion_training_b200.schedulers import FlaxDDIMScheduler
    z = np.load(os.path.join(G, "ddim_ref.npz"))
    s = FlaxDDIMScheduler(beta_start=0.00085, beta_end=0.012, beta_schedule="scaled_linear", num_train_timesteps=1000,
                          set_alpha_to_one=one, steps_offset=1, prediction_type=ptype, device=cuda)
    st0 = s.create_state()
    mo, smp = torch.from_numpy(z["model_output"]).to(cuda), torch.from_numpy(z["sample"]).to(cuda)
    with pytest.raises(ValueError):
        s.step(st0, mo, 981, smp)
    st = s.set_timesteps(st0, 50)
    assert np.array_equal(_np(st.timesteps), z[f"timesteps/{int(one)}"])
    assert np.array_equal(_bits(_np(st.final_alpha_cumprod)), _bits(z[f"final/{int(one)}"]))
    assert len(s) == 1000 and s.scale_model_input(st, smp) is smp
    for ti in (981, 501, 21, 1):
        assert np.array_equal(_bits(s._get_variance(st, ti, ti - 20)), _bits(z[f"var/{int(one)}/{ti}"]))
        for eta in (0.0, 0.5):
            out = s.step(st, mo, ti, smp, eta=eta)
            assert np.array_equal(_bits(_np(out.prev_sample)), _bits(z[f"step/{ptype}/{int(one)}/{ti}/{eta}"])), (ti, eta)
    prev, st2 = s.step(st, mo, 501, smp, return_dict=False)
    assert st2 is st and prev.shape == smp.shape
