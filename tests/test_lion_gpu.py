"""GPU parity of the 8-bit Lion kernels (C ABI sdb_lion8_update / sdb_lion_fp32_update / sdb_global_norm) against
the NumPy oracle (oracle/lion.py, restating /root/reference/lion_quant.py). Criterion: int8 codes, reciprocal-absmax
scales AND updated fp32 parameters bit-exact given identical fp32 inputs."""
import numpy as np
import pytest
import torch

from oracle import lion as O
from stable_diffusion_training_b200 import _lib

pytestmark = pytest.mark.gpu


def _run_gpu(dev, g, p, codes, inv, gnorm, wd_flag, mode=0, shadow=False):
    l = _lib.lib()
    tg, tp = torch.from_numpy(g).to(dev), torch.from_numpy(p).to(dev)
    tc, ti = torch.from_numpy(codes.reshape(-1)).to(dev), torch.from_numpy(inv.reshape(-1)).to(dev)
    out = torch.empty_like(tp)
    sh = torch.empty(tp.shape, dtype=torch.bfloat16, device=dev) if shadow else None
    tgn = None if gnorm is None else torch.tensor([gnorm], dtype=torch.float32, device=dev)
    _lib.check(l.sdb_lion8_update(tg.data_ptr(), tp.data_ptr(), out.data_ptr(), _lib.ptr(sh), tc.data_ptr(),
                                  ti.data_ptr(), g.size, 16, 0.9, 0.99, 1e-6 / 7, 0.07, int(wd_flag), _lib.ptr(tgn),
                                  1.0, mode, 0, 0.0, _lib.stream_ptr()), "lion8")
    torch.cuda.synchronize()
    return out.cpu().numpy(), tc.cpu().numpy(), ti.cpu().numpy(), None if sh is None else sh.float().cpu().numpy()


def _oracle(g, p, codes, inv, gnorm, wd_flag):
    gc = g if gnorm is None else O.clip_leaf(g, gnorm, 1.0)
    newp, (c2, i2), upd = O.lion_leaf_update(gc, p, (codes, inv), wd_flag=wd_flag)
    return newp, c2.reshape(-1), i2.reshape(-1), upd


@pytest.mark.parametrize("n,gscale,gnorm", [(16, 1e-3, None), (4096, 3.4e-5, None), (1 << 20, 3.4e-5, 0.5),
                                            (1 << 20, 1e-2, 37.25), (16 * 1000 + 16, 1.0, 1.0), (1 << 22, 1e-8, None)])
@pytest.mark.parametrize("wd_flag", [True, False])
def test_lion8_bit_exact_multi_step(cuda, n, gscale, gnorm, wd_flag):
    rng = np.random.default_rng(n % 1000 + 7)
    p = (rng.standard_normal(n) * 0.02).astype(np.float32)
    codes, inv = O.init_state((n,), True)
    assert (codes == 3).all() and (inv == 1.0).all()
    for step in range(4):
        g = (rng.standard_normal(n) * gscale).astype(np.float32)
        if step == 2:
            g[: n // 2] = 0.0  # zero blocks -> absmax path / code 3
        want_p, want_c, want_i, _ = _oracle(g, p, codes, inv, gnorm, wd_flag)
        got_p, got_c, got_i, _ = _run_gpu(cuda, g, p, codes, inv, gnorm, wd_flag)
        assert np.array_equal(got_c, want_c), f"step {step}: {(got_c != want_c).sum()} code mismatches"
        assert np.array_equal(got_i.view(np.uint32), want_i.view(np.uint32)), f"step {step}: scale mismatch"
        assert np.array_equal(got_p.view(np.uint32), want_p.view(np.uint32)), f"step {step}: param mismatch"
        p, codes, inv = want_p, want_c.reshape(-1, 16), want_i.reshape(-1, 1)


def test_lion8_update_mode_and_shadow(cuda):
    n = 1 << 16
    rng = np.random.default_rng(3)
    p = (rng.standard_normal(n) * 0.02).astype(np.float32)
    g = (rng.standard_normal(n) * 1e-4).astype(np.float32)
    codes, inv = O.block_quantize((rng.standard_normal(n) * 1e-4).astype(np.float32))
    want_p, want_c, want_i, want_u = _oracle(g, p, codes, inv, None, True)
    got_u, got_c, _, _ = _run_gpu(cuda, g, p, codes, inv, None, True, mode=1)
    assert np.array_equal(got_u.view(np.uint32), want_u.view(np.uint32))
    assert np.array_equal(got_c, want_c)
    got_p, _, _, sh = _run_gpu(cuda, g, p, codes, inv, None, True, mode=0, shadow=True)
    assert np.array_equal(sh, torch.from_numpy(want_p).to(torch.bfloat16).float().numpy())


def test_lion8_extreme_values(cuda):
    """Edge cases: huge/tiny magnitudes, exact +-absmax (codes +-127), denormal-range momentum."""
    n = 16 * 64
    rng = np.random.default_rng(11)
    p = (rng.standard_normal(n)).astype(np.float32)
    for scale in (1e-30, 1e-20, 1e-10, 1.0, 1e10, 1e20):
        g = (rng.standard_normal(n) * scale).astype(np.float32)
        m = (rng.standard_normal(n) * scale).astype(np.float32)
        codes, inv = O.block_quantize(m)
        want_p, want_c, want_i, _ = _oracle(g, p, codes, inv, None, True)
        got_p, got_c, got_i, _ = _run_gpu(cuda, g, p, codes, inv, None, True)
        assert np.array_equal(got_c, want_c), scale
        assert np.array_equal(got_i.view(np.uint32), want_i.view(np.uint32)), scale
        assert np.array_equal(got_p.view(np.uint32), want_p.view(np.uint32)), scale


@pytest.mark.parametrize("n", [4, 1000 * 4, 1 << 20])
def test_lion_fp32_state_bit_exact(cuda, n):
    l = _lib.lib()
    rng = np.random.default_rng(5)
    p = (rng.standard_normal(n) * 0.02).astype(np.float32)
    g = (rng.standard_normal(n) * 1e-3).astype(np.float32)
    mu = (rng.standard_normal(n) * 1e-3).astype(np.float32)
    for wd_flag in (True, False):
        want_p, want_m, _ = O.lion_leaf_update(O.clip_leaf(g, 2.5, 1.0), p, mu, wd_flag=wd_flag)
        tg, tp, tm = (torch.from_numpy(a).to(cuda) for a in (g, p, mu))
        gn = torch.tensor([2.5], dtype=torch.float32, device=cuda)
        _lib.check(l.sdb_lion_fp32_update(tg.data_ptr(), tp.data_ptr(), tp.data_ptr(), 0, tm.data_ptr(), n, 0.9, 0.99,
                                          1e-6 / 7, 0.07, int(wd_flag), gn.data_ptr(), 1.0, 0, 0, 0.0, _lib.stream_ptr()), "l32")
        torch.cuda.synchronize()
        assert np.array_equal(tp.cpu().numpy().view(np.uint32), want_p.view(np.uint32))
        assert np.array_equal(tm.cpu().numpy().view(np.uint32), want_m.view(np.uint32))


def test_global_norm(cuda):
    l = _lib.lib()
    n = (1 << 22) + 4
    rng = np.random.default_rng(9)
    x = (rng.standard_normal(n) * 3.4e-4).astype(np.float32)
    tx = torch.from_numpy(x).to(cuda)
    ws = torch.empty(l.sdb_sumsq_workspace_bytes(), dtype=torch.uint8, device=cuda)
    out = torch.zeros(1, dtype=torch.float32, device=cuda)
    _lib.check(l.sdb_global_norm(tx.data_ptr(), n, ws.data_ptr(), ws.numel(), out.data_ptr(), 0, _lib.stream_ptr()), "gn")
    _lib.check(l.sdb_global_norm(tx.data_ptr(), n, ws.data_ptr(), ws.numel(), out.data_ptr(), 1, _lib.stream_ptr()), "gn")
    torch.cuda.synchronize()
    want = np.sqrt(2.0 * np.sum(x.astype(np.float64) ** 2))
    assert abs(out.item() - want) <= 2e-7 * want
