"""GPU parity of the frozen VAE encode (/root/reference/training_utils.py:574-586) against
oracle/vae.py (hardened by tests/test_oracle_vae.py): helper kernels, the encoder at toy and REAL widths (128 / 256 / 512
/ 512 channels, single-head 512-wide mid attention) vs the fp64 oracle with the bf16-emulation yardstick, the latent
sample with the step's `sample_rng`, and the train step consuming `pixel_values`."""
import math

import numpy as np
import pytest
import torch

from oracle import threefry as OT
from oracle import vae as OV
from stable_diffusion_training_b200 import _lib
from stable_diffusion_training_b200 import jax_random as JR
from stable_diffusion_training_b200 import lion_quant as LQ
from stable_diffusion_training_b200 import training_utils as TU
from stable_diffusion_training_b200 import unet as PU
from stable_diffusion_training_b200 import vae as PV

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = np.asarray(a, np.float64).ravel(), np.asarray(b, np.float64).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def test_softmax_rows_and_sample_kernels(cuda):
    l = _lib.lib()
    x = (torch.randn(300, 1000) * 3).to(torch.bfloat16).to(cuda)
    out = torch.empty_like(x)
    _lib.check(l.sdb_softmax_rows_bf16(x.data_ptr(), out.data_ptr(), 300, 1000, 1000, 0.25, _lib.stream_ptr()), "softmax")
    ref = torch.softmax(x.float().cpu() * 0.25, dim=-1)
    assert (out.float().cpu() - ref).abs().max().item() < 2e-3
    B, L, HW = 2, 4, 96
    mom = (torch.randn(B, HW, 2 * L) * 2).to(torch.bfloat16)
    mom[0, 0, L:] = 50.0   # logvar clipped to 20
    mom[0, 1, L:] = -50.0  # and to -30
    eps = torch.randn(B, HW, L)
    lat = torch.empty(B, L, HW, device=cuda)
    mom_d, eps_d = mom.to(cuda), eps.to(cuda)
    _lib.check(l.sdb_vae_sample(mom_d.data_ptr(), eps_d.data_ptr(), lat.data_ptr(), B, L, HW, 0.18215,
                                _lib.stream_ptr()), "vae_sample")
    torch.cuda.synchronize()
    want = OV.sample_latents(OV.SD_VAE, mom.float().view(B, 1, HW, 2 * L), eps.view(B, 1, HW, L)).reshape(B, L, HW)
    assert _rel(lat.cpu().numpy(), want.numpy()) < 1e-5


def _encoder(cuda, pcfg, seed=2):
    params_np = PV.init_params_numpy(pcfg, seed)
    tree = LQ.unflatten_dict(params_np)
    model = PV.AutoencoderKL(pcfg)
    return params_np, model, tree, model.encoder(tree, cuda)


@pytest.mark.parametrize("name,B,H,W", [("tiny", 2, 64, 72), ("sd", 1, 128, 128)])
def test_vae_encode_matches_oracle(cuda, name, B, H, W):
    pcfg, ocfg = (PV.TINY_VAE, OV.TINY_VAE) if name == "tiny" else (PV.SD_VAE, OV.SD_VAE)
    assert PV.param_shapes(pcfg) == OV.param_shapes(ocfg)
    params_np, model, tree, enc = _encoder(cuda, pcfg)
    g = torch.Generator().manual_seed(7)
    px = (torch.rand(B, 3, H, W, generator=g) * 2 - 1)
    mom = enc.encode(px.to(cuda))
    torch.cuda.synchronize()
    op = {k: torch.from_numpy(v) for k, v in params_np.items()}
    m64 = OV.encode_moments(ocfg, op, px, mode="fp64")
    m16 = OV.encode_moments(ocfg, op, px, mode="bf16")
    down = 2 ** (len(pcfg.block_out_channels) - 1)
    assert mom.shape == (B, (H // down) * (W // down), 8)
    got = mom.float().cpu().view(m64.shape)
    err, yard = _rel(got.numpy(), m64.numpy()), _rel(m16.numpy(), m64.numpy())
    print(f"\nVAE encode {name} {H}x{W}: moments err {err:.3e} (bf16-emu yardstick {yard:.3e})")
    assert err <= 3 * yard
    # latent sample with a jax key: the draw has the NHWC mean shape
    key = JR.PRNGKey(11)
    lat = enc.sample(mom, (H // down, W // down), key=key)
    eps = torch.from_numpy(OT.normal(OT.PRNGKey(11), (B, H // down, W // down, 4)))
    want = OV.sample_latents(ocfg, got, eps)
    assert lat.shape == want.shape and _rel(lat.cpu().numpy(), want.numpy()) < 1e-4
    mode = enc.sample(mom, (H // down, W // down), key=None)
    assert _rel(mode.cpu().numpy(), (got[..., :4].permute(0, 3, 1, 2) * 0.18215).numpy()) < 1e-5


def test_train_step_from_pixel_values(cuda):
    """The reference's batch surface: {"pixel_values", ...} + frozen VAE (training_utils.py:574-586, 801-830)."""
    ucfg = PU.TINY
    up = LQ.unflatten_dict(PU.init_params_numpy(ucfg, 0))
    s1 = TU.TrainState.create(apply_fn=TU.UNetModel(ucfg).apply, params=up, tx=TU.create_lion_optimizer(up), device=cuda)
    up2 = LQ.unflatten_dict(PU.init_params_numpy(ucfg, 0))
    s2 = TU.TrainState.create(apply_fn=TU.UNetModel(ucfg).apply, params=up2, tx=TU.create_lion_optimizer(up2), device=cuda)
    _, model, tree, enc = _encoder(cuda, PV.TINY_VAE)
    frozen_vae = TU.FrozenModel(call=model, params=tree)
    sched = TU.default_scheduler(cuda, "scaled_linear", "epsilon")
    g = torch.Generator().manual_seed(3)
    B = 2
    px = (torch.rand(B, 3, 32, 32, generator=g) * 2 - 1).to(cuda)   # TINY_VAE: /2 -> 16x16 latents
    ctx = torch.randn(B, 77, ucfg.cross_attention_dim, generator=g).to(cuda)
    batch = {"pixel_values": px, "encoder_hidden_states": ctx}
    with pytest.raises(ValueError):
        TU.train_step(s1, batch, None, sched, frozen_vae_state=frozen_vae)
    stream = JR.RngStream(21)
    s1, m1 = TU.train_step(s1, batch, stream, sched, frozen_vae_state=frozen_vae)
    keys = JR.train_step_keys(JR.PRNGKey(21))
    assert stream.key == keys["new_train"]
    # the same step fed explicitly: latents sampled with sample_rng, noise / timesteps drawn from the same train_rng
    lat = enc.sample(enc.encode(px), (16, 16), key=keys["sample"])
    s2, m2 = TU.train_step(s2, {"latents": lat, "encoder_hidden_states": ctx}, JR.RngStream(21), sched)
    torch.cuda.synchronize()
    assert abs(float(m1["loss"]) - float(m2["loss"])) <= 1e-6 * abs(float(m2["loss"]))
    # compiled: the frozen encode runs in front of the graph
    step = TU.compile_train_step(s1, sched, frozen_vae_state=frozen_vae)
    for _ in range(3):
        s1, m = step(s1, batch, stream)
    assert int(s1.step) == 4 and math.isfinite(float(m["loss"]))
