"""GPU parity of add_noise / get_velocity / ddpm_step / MSE against oracle/sched.py (bit-exact for the fp32
elementwise ops; the loss is a mean over fp64 partial sums -> 1e-6 relative)."""
import numpy as np
import pytest
import torch

from oracle import sched as O
from stable_diffusion_training_b200 import _lib

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("schedule", ["scaled_linear", "zero_snr_scaled_linear", "linear"])
@pytest.mark.parametrize("shape", [(2, 4, 32, 32), (8, 4, 64, 64), (3, 4, 96, 96)])
def test_add_noise_velocity_bit_exact(cuda, schedule, shape):
    l = _lib.lib()
    st = O.common_state(schedule, 0.00085, 0.012)
    rng = np.random.default_rng(1)
    x0 = rng.standard_normal(shape).astype(np.float32)
    eps = rng.standard_normal(shape).astype(np.float32)
    t = rng.integers(0, 1000, shape[0]).astype(np.int32)
    t[0] = 999
    t[-1] = 0
    ac = torch.from_numpy(st["alphas_cumprod"]).to(cuda)
    tx, te, tt = torch.from_numpy(x0).to(cuda), torch.from_numpy(eps).to(cuda), torch.from_numpy(t).to(cuda)
    out = torch.empty_like(tx)
    per = int(np.prod(shape[1:]))
    for fn, ref in ((l.sdb_add_noise, O.add_noise), (l.sdb_get_velocity, O.get_velocity)):
        _lib.check(fn(tx.data_ptr(), te.data_ptr(), tt.data_ptr(), ac.data_ptr(), ac.numel(), out.data_ptr(), shape[0],
                      per, _lib.stream_ptr()), "noise")
        torch.cuda.synchronize()
        want = ref(st["alphas_cumprod"], x0, eps, t)
        assert np.array_equal(out.cpu().numpy().view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("ptype", ["epsilon", "sample", "v_prediction"])
@pytest.mark.parametrize("t", [0, 1, 500, 999])
def test_ddpm_step_bit_exact(cuda, ptype, t):
    l = _lib.lib()
    st = O.common_state("scaled_linear", 0.00085, 0.012)
    rng = np.random.default_rng(2)
    shape = (2, 4, 16, 16)
    mo, x, nz = (rng.standard_normal(shape).astype(np.float32) for _ in range(3))
    c = O.step_coefficients(st, t)
    sigma = float(np.sqrt(O.get_variance(st, t)).astype(np.float32)) if t > 0 else 0.0
    tm, tx, tn = (torch.from_numpy(a).to(cuda) for a in (mo, x, nz))
    out = torch.empty_like(tx)
    _lib.check(l.sdb_ddpm_step(tm.data_ptr(), tx.data_ptr(), tn.data_ptr() if t > 0 else 0, out.data_ptr(), mo.size,
                               float(c["sqrt_alpha_prod"]), float(c["sqrt_beta_prod"]), float(c["x0_coeff"]),
                               float(c["sample_coeff"]), sigma, ["epsilon", "sample", "v_prediction"].index(ptype), 1,
                               _lib.stream_ptr()), "step")
    torch.cuda.synchronize()
    want = O.ddpm_step(st, mo, t, x, nz, prediction_type=ptype)
    assert np.array_equal(out.cpu().numpy().view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("weighted", [False, True])
def test_mse_fwd_bwd(cuda, weighted):
    l = _lib.lib()
    B, Cc, H, W = 8, 4, 64, 64
    rng = np.random.default_rng(4)
    pred_nchw = rng.standard_normal((B, Cc, H, W)).astype(np.float32)
    target = rng.standard_normal((B, Cc, H, W)).astype(np.float32)
    w = rng.uniform(0.1, 1.0, B).astype(np.float32) if weighted else None
    pred = torch.from_numpy(pred_nchw).to(cuda).permute(0, 2, 3, 1).contiguous().to(torch.bfloat16)
    tt = torch.from_numpy(target).to(cuda)
    tw = None if w is None else torch.from_numpy(w).to(cuda)
    dpred = torch.empty_like(pred)
    loss = torch.zeros(1, dtype=torch.float32, device=cuda)
    ws = torch.empty(l.sdb_mse_workspace_bytes(), dtype=torch.uint8, device=cuda)
    _lib.check(l.sdb_mse_fwd_bwd(pred.data_ptr(), tt.data_ptr(), _lib.ptr(tw), dpred.data_ptr(), loss.data_ptr(), B,
                                 Cc, H * W, 1.0, ws.data_ptr(), ws.numel(), _lib.stream_ptr()), "mse")
    torch.cuda.synchronize()
    pred_r = pred.float().permute(0, 3, 1, 2).cpu().numpy()
    want = O.mse_loss(pred_r, target, w)
    assert abs(loss.item() - want) <= 1e-6 * abs(want)
    wb = np.ones(B, np.float32) if w is None else w
    gref = (2.0 * (pred_r - target) * wb[:, None, None, None] / pred_r.size)
    got = dpred.float().permute(0, 3, 1, 2).cpu().numpy()
    assert np.allclose(got, gref, rtol=1 / 128, atol=1e-9)
