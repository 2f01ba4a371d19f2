"""GPU tests of SURVEY.md 8(f)-1: the static options of the reference's train step
(/root/reference/training_utils.py:517-521) — offset noise, perturbation noise, min-SNR-gamma weighting, EMA — the
in-step sampling from `train_rng`, and the Lion block sizes other than 16.

 * EMA (537-544, 735-739) fused into the optimizer launches: bit-exact vs oracle/lion.py:ema_update.
 * block_size 32 / 64 (lion_quant.py:66-80, default 64 at 164): codes / scales / parameters bit-exact vs the oracle.
 * min-SNR weights (546-568) computed on the device: bit-exact vs oracle/lion.py:min_snr_weight; the weighted step
   vs oracle.unet.loss_and_grads(snr_weight=...).
 * sampling (530, 590-624): a step that draws its own noise / timesteps equals a step fed the oracle's draw.
"""
import numpy as np
import pytest
import torch

from oracle import lion as OL
from oracle import sched as OS
from oracle import threefry as OT
from oracle import unet as OU
from stable_diffusion_training_b200 import _lib
from stable_diffusion_training_b200 import jax_random as JR
from stable_diffusion_training_b200 import lion_quant as LQ
from stable_diffusion_training_b200 import training_utils as TU
from stable_diffusion_training_b200 import unet as PU

pytestmark = pytest.mark.gpu
EMA_RATE = 0.99998  # model_properties_example.json:40


def _bits(a):
    return np.asarray(a, dtype=np.float32).view(np.uint32)


@pytest.mark.parametrize("bs", [16, 32, 64])
@pytest.mark.parametrize("with_ema", [False, True])
def test_lion8_block_sizes_and_fused_ema_bit_exact(cuda, bs, with_ema):
    l = _lib.lib()
    n = 64 * 3001  # not a multiple of the per-warp trip: exercises the partial last trip
    rng = np.random.default_rng(bs)
    p = (rng.standard_normal(n) * 0.02).astype(np.float32)
    ema = p.copy()
    ema0 = p.copy()
    codes, inv = OL.init_state((n,), True, block_size=bs)
    tp, te = torch.from_numpy(p).to(cuda), torch.from_numpy(ema).to(cuda)
    tc, ti = torch.from_numpy(codes.reshape(-1).copy()).to(cuda), torch.from_numpy(inv.reshape(-1).copy()).to(cuda)
    sh = torch.empty(n, dtype=torch.bfloat16, device=cuda)
    for step in range(3):
        g = (rng.standard_normal(n) * 3.4e-5).astype(np.float32)
        if step == 1:
            g[: n // 3] = 0.0
        tg = torch.from_numpy(g).to(cuda)
        p, (codes, inv), _ = OL.lion_leaf_update(g, p, (codes, inv), wd_flag=True, block_size=bs)
        ema = OL.ema_update(ema, p, EMA_RATE)
        _lib.check(l.sdb_lion8_update(tg.data_ptr(), tp.data_ptr(), tp.data_ptr(), sh.data_ptr(), tc.data_ptr(),
                                      ti.data_ptr(), n, bs, 0.9, 0.99, 1e-6 / 7, 0.07, 1, 0, 1.0, 0,
                                      te.data_ptr() if with_ema else 0, EMA_RATE, _lib.stream_ptr()), "lion8")
        torch.cuda.synchronize()
        assert np.array_equal(tc.cpu().numpy(), codes.reshape(-1)), (bs, step)
        assert np.array_equal(_bits(ti.cpu().numpy()), _bits(inv.reshape(-1))), (bs, step)
        assert np.array_equal(_bits(tp.cpu().numpy()), _bits(p)), (bs, step)
        if with_ema:
            assert np.array_equal(_bits(te.cpu().numpy()), _bits(ema)), (bs, step)
        assert torch.equal(sh, tp.to(torch.bfloat16))
    if not with_ema:
        assert np.array_equal(_bits(te.cpu().numpy()), _bits(ema0))  # untouched


def test_lion_fp32_fused_ema_and_standalone_ema_bit_exact(cuda):
    l = _lib.lib()
    n = 4 * 12345
    rng = np.random.default_rng(3)
    p = (rng.standard_normal(n) * 0.02).astype(np.float32)
    mu = (rng.standard_normal(n) * 1e-4).astype(np.float32)
    ema = (p + rng.standard_normal(n).astype(np.float32) * 1e-3).astype(np.float32)
    tp, tm, te = (torch.from_numpy(a.copy()).to(cuda) for a in (p, mu, ema))
    te2 = te.clone()
    for rate in (EMA_RATE, 0.9, 0.5):
        g = (rng.standard_normal(n) * 1e-3).astype(np.float32)
        tg = torch.from_numpy(g).to(cuda)
        p, mu, _ = OL.lion_leaf_update(g, p, mu, wd_flag=False)
        ema = OL.ema_update(ema, p, rate)
        _lib.check(l.sdb_lion_fp32_update(tg.data_ptr(), tp.data_ptr(), tp.data_ptr(), 0, tm.data_ptr(), n, 0.9, 0.99,
                                          1e-6 / 7, 0.07, 0, 0, 1.0, 0, te.data_ptr(), rate, _lib.stream_ptr()), "l32")
        _lib.check(l.sdb_ema_update(te2.data_ptr(), tp.data_ptr(), n, rate, _lib.stream_ptr()), "ema")
        torch.cuda.synchronize()
        assert np.array_equal(_bits(tp.cpu().numpy()), _bits(p))
        assert np.array_equal(_bits(tm.cpu().numpy()), _bits(mu))
        assert np.array_equal(_bits(te.cpu().numpy()), _bits(ema))
        assert torch.equal(te, te2)


@pytest.mark.parametrize("schedule", ["scaled_linear", "zero_snr_scaled_linear"])
@pytest.mark.parametrize("v_pred", [False, True])
def test_min_snr_weight_bit_exact(cuda, schedule, v_pred):
    ac = OS.common_state(schedule, 0.00085, 0.012)["alphas_cumprod"]
    t = np.array([0, 1, 17, 250, 500, 777, 998, 999], dtype=np.int32)
    for gamma in (5.0, 1.0, 0.05):
        want = OL.min_snr_weight(ac, t, gamma, v_pred)
        got = TU.min_snr_weight(torch.from_numpy(ac).to(cuda), torch.from_numpy(t).to(cuda), gamma, v_pred).cpu().numpy()
        assert np.array_equal(got, want, equal_nan=True), (schedule, v_pred, gamma, got, want)
        ok = ~np.isnan(want)
        assert np.array_equal(_bits(got[ok]), _bits(want[ok]))


def _tiny_state(cuda, block_size=16):
    cfg = PU.TINY
    params_np = PU.init_params_numpy(cfg, 0)
    tree = LQ.unflatten_dict({k: v.copy() for k, v in params_np.items()})
    tx = TU.create_lion_optimizer(tree, TU.TrainingConfig(lion_8bit_block_size=block_size))
    state = TU.TrainState.create(apply_fn=TU.UNetModel(cfg).apply, params=tree, tx=tx, device=cuda)
    return cfg, params_np, state


def _np_batch(B, H, L, D, seed=1, with_noise=True):
    rng = np.random.default_rng(seed)
    b = {"latents": rng.standard_normal((B, 4, H, H), dtype=np.float32),
         "encoder_hidden_states": rng.standard_normal((B, L, D), dtype=np.float32)}
    if with_noise:
        b["noise"] = rng.standard_normal((B, 4, H, H), dtype=np.float32)
        b["timesteps"] = rng.integers(0, 1000, B).astype(np.int32)
    return b


@pytest.mark.parametrize("ptype", ["epsilon", "v_prediction"])
def test_train_step_min_snr_matches_oracle(cuda, ptype):
    cfg, params_np, state = _tiny_state(cuda)
    sched = TU.default_scheduler(cuda, "scaled_linear", ptype)
    b = _np_batch(2, 16, 77, cfg.cross_attention_dim, seed=4)
    b["timesteps"] = np.array([40, 900], dtype=np.int32)  # one weight clamped to gamma / snr, one at 1 (epsilon)
    gamma = 5.0
    batch = {k: torch.from_numpy(v).to(cuda) for k, v in b.items()}
    state, m = TU.train_step(state, batch, None, sched, min_snr_gamma_magnitude=gamma)
    torch.cuda.synchronize()
    ac = OS.common_state("scaled_linear", 0.00085, 0.012)["alphas_cumprod"]
    w = OL.min_snr_weight(ac, b["timesteps"], gamma, ptype == "v_prediction")
    assert not np.allclose(w, 1.0)
    op = {n: torch.from_numpy(v) for n, v in params_np.items()}
    args = (torch.from_numpy(b["latents"]), torch.from_numpy(b["noise"]), torch.from_numpy(b["timesteps"]).long(),
            torch.from_numpy(b["encoder_hidden_states"]), ac)
    l64, g64, _ = OU.loss_and_grads(OU.TINY, op, *args, prediction_type=ptype, mode="fp64", snr_weight=w)
    l16, g16, _ = OU.loss_and_grads(OU.TINY, op, *args, prediction_type=ptype, mode="bf16", snr_weight=w)
    l_unw, _, _ = OU.loss_and_grads(OU.TINY, op, *args, prediction_type=ptype, mode="fp64")
    names = sorted(params_np)
    v64 = np.concatenate([g64[n].numpy().ravel() for n in names]).astype(np.float64)
    v16 = np.concatenate([g16[n].numpy().ravel() for n in names]).astype(np.float64)
    vg = np.concatenate([state.arena.g[n].detach().cpu().numpy().ravel() for n in names]).astype(np.float64)
    err = np.linalg.norm(vg - v64) / np.linalg.norm(v64)
    yard = np.linalg.norm(v16 - v64) / np.linalg.norm(v64)
    loss = float(m["loss"])
    print(f"min-SNR {ptype}: loss gpu {loss:.6f} fp64 {float(l64):.6f} (unweighted {float(l_unw):.6f}); grad err {err:.3e} "
          f"yardstick {yard:.3e}")
    assert abs(loss - float(l64)) <= max(3 * abs(float(l16) - float(l64)), 2e-3 * abs(float(l64)))
    assert abs(float(l_unw) - float(l64)) > 10 * abs(loss - float(l64))  # the weight is really applied
    assert err <= 3 * yard


def test_train_step_draws_the_reference_noise_and_advances_the_key(cuda):
    """No 'noise' / 'timesteps' in the batch: the step samples them like training_utils.py:590-624 (offset and
    perturbation noise on). Feeding the oracle's draw explicitly gives the same loss; new_train_rng is the oracle's;
    the next step draws different noise (the generator must advance)."""
    off, pert = 0.1, 0.05
    cfg, params_np, s1 = _tiny_state(cuda)
    _, _, s2 = _tiny_state(cuda)
    sched = TU.default_scheduler(cuda, "scaled_linear", "epsilon")
    b = _np_batch(2, 16, 77, cfg.cross_attention_dim, seed=6, with_noise=False)
    batch = {k: torch.from_numpy(v).to(cuda) for k, v in b.items()}
    key = JR.PRNGKey(2024)
    with pytest.raises(ValueError):
        TU.train_step(s1, batch, None, sched)
    with pytest.raises(ValueError):
        TU.train_step(s1, batch, key, sched)  # raw key without return_rng would repeat itself
    s1, m1, new_key = TU.train_step(s1, batch, key, sched, offset_noise_magnitude=off, perturbation_noise_magnitude=pert,
                                    return_rng=True)
    noise, t, want_new = OT.sample_noise_and_timesteps(OT.PRNGKey(2024), b["latents"].shape, 1000, off, pert)
    assert list(new_key) == want_new.tolist()
    fed = dict(batch, noise=torch.from_numpy(noise).to(cuda), timesteps=torch.from_numpy(t).to(cuda))
    s2, m2 = TU.train_step(s2, fed, None, sched)
    torch.cuda.synchronize()
    l1, l2 = float(m1["loss"]), float(m2["loss"])
    assert abs(l1 - l2) <= 1e-4 * abs(l2), (l1, l2)
    # second step from new_train_rng: different draw
    s1, m3, key3 = TU.train_step(s1, batch, new_key, sched, offset_noise_magnitude=off,
                                 perturbation_noise_magnitude=pert, return_rng=True)
    assert list(key3) != list(new_key)
    assert abs(float(m3["loss"]) - l1) > 1e-3 * abs(l1)
    # RngStream: same sequence, advanced in place
    _, _, s3 = _tiny_state(cuda)
    stream = JR.RngStream(2024)
    s3, m4 = TU.train_step(s3, batch, stream, sched, offset_noise_magnitude=off, perturbation_noise_magnitude=pert)
    assert stream.key == tuple(want_new.tolist())
    assert abs(float(m4["loss"]) - l1) <= 1e-6 * abs(l1)


def test_bf16_and_fp64_batches_are_promoted(cuda):
    cfg, _, s1 = _tiny_state(cuda)
    _, _, s2 = _tiny_state(cuda)
    sched = TU.default_scheduler(cuda, "scaled_linear", "v_prediction")
    b = _np_batch(2, 16, 77, cfg.cross_attention_dim, seed=8)
    lat16 = torch.from_numpy(b["latents"]).to(cuda).to(torch.bfloat16)
    nz16 = torch.from_numpy(b["noise"]).to(cuda).to(torch.bfloat16)
    batch_a = {"latents": lat16, "noise": nz16, "timesteps": torch.from_numpy(b["timesteps"]).to(cuda).long(),
               "encoder_hidden_states": torch.from_numpy(b["encoder_hidden_states"]).to(cuda)}
    batch_b = dict(batch_a, latents=lat16.double(), noise=nz16.float())
    s1, m1 = TU.train_step(s1, batch_a, None, sched)
    s2, m2 = TU.train_step(s2, batch_b, None, sched)
    torch.cuda.synchronize()
    assert abs(float(m1["loss"]) - float(m2["loss"])) <= 1e-6 * abs(float(m2["loss"]))
    # out-of-range timesteps are clamped like jnp's gather, never read out of bounds
    st = sched.params
    x = torch.randn(2, 4, 8, 8, device=cuda)
    a = sched.call.add_noise(st, x, x, torch.tensor([5000, -1], device=cuda))
    bq = sched.call.add_noise(st, x, x, torch.tensor([999, 999], device=cuda))
    assert torch.equal(a, bq)
    with pytest.raises(ValueError):
        sched.call.add_noise(st, x.cpu(), x, torch.tensor([1, 2], device=cuda))


@pytest.mark.parametrize("mode", ["eager", "graph"])
def test_train_step_ema_is_bit_exact(cuda, mode):
    cfg, params_np, state = _tiny_state(cuda)
    sched = TU.default_scheduler(cuda, "scaled_linear", "epsilon")
    ema_tree = state.create_ema()
    ema = state.arena.ema.cpu().numpy().copy()
    assert np.array_equal(_bits(ema), _bits(state.arena.master.cpu().numpy()))
    flags = dict(unet_ema_params=ema_tree, ema_rate=EMA_RATE, min_snr_gamma_magnitude=5.0)
    step = TU.compile_train_step(state, sched, **flags) if mode == "graph" else \
        (lambda s, b, r: TU.train_step(s, b, r, sched, **flags))
    for i in range(4):
        b = _np_batch(2, 16, 77, cfg.cross_attention_dim, seed=20 + i)
        state, m = step(state, {k: torch.from_numpy(v).to(cuda) for k, v in b.items()}, None)
        torch.cuda.synchronize()
        ema = OL.ema_update(ema, state.arena.master.cpu().numpy(), EMA_RATE)
        assert np.array_equal(_bits(state.arena.ema.cpu().numpy()), _bits(ema)), (mode, i)
    assert int(state.step) == 4
    leaf = LQ.flatten_dict(ema_tree)["conv_in/kernel"]
    assert leaf.data_ptr() == state.arena.e["conv_in/kernel"].data_ptr()


def test_foreign_ema_tree_goes_through_compute_model_ema(cuda):
    cfg, params_np, state = _tiny_state(cuda)
    sched = TU.default_scheduler(cuda, "scaled_linear", "epsilon")
    foreign = LQ.tree_map(lambda p: p.detach().clone(), state.params)
    before = {n: v.cpu().numpy().copy() for n, v in LQ.flatten_dict(foreign).items()}
    b = _np_batch(2, 16, 77, cfg.cross_attention_dim, seed=30)
    state, _ = TU.train_step(state, {k: torch.from_numpy(v).to(cuda) for k, v in b.items()}, None, sched,
                             unet_ema_params=foreign, ema_rate=0.9)
    torch.cuda.synchronize()
    for n, e in LQ.flatten_dict(foreign).items():
        want = OL.ema_update(before[n].reshape(-1), state.arena.p[n].cpu().numpy().reshape(-1), 0.9)
        assert np.array_equal(_bits(e.cpu().numpy().reshape(-1)), _bits(want)), n


def test_block_size_64_train_step(cuda):
    """lion_8bit's default block (lion_quant.py:164) through the arena path: leaves aligned to 64, codes bit-exact."""
    cfg, params_np, state = _tiny_state(cuda, block_size=64)
    sched = TU.default_scheduler(cuda, "scaled_linear", "epsilon")
    b = _np_batch(2, 16, 77, cfg.cross_attention_dim, seed=31)
    state, _ = TU.train_step(state, {k: torch.from_numpy(v).to(cuda) for k, v in b.items()}, None, sched)
    torch.cuda.synchronize()
    gn = np.float32(state.arena.gnorm.item())
    mu = LQ.flatten_dict(state.opt_state[1].lion.mu_quant)
    tree = LQ.unflatten_dict(params_np)
    wd = LQ.flatten_dict(TU.create_mask(tree, ["bias", "scale", "embedding"]))
    qm = LQ.flatten_dict(TU.create_mask(tree, list(TU.TrainingConfig().excluded_layer_from_quantization)))
    n_q = 0
    for n in sorted(params_np):
        g = OL.clip_leaf(state.arena.g[n].cpu().numpy(), gn, 1.0)
        st0 = OL.init_state((g.size,), bool(qm[n]), block_size=64)
        want_p, want_s, _ = OL.lion_leaf_update(g.reshape(-1), params_np[n].reshape(-1), st0, wd_flag=bool(wd[n]),
                                                block_size=64)
        assert np.array_equal(_bits(state.arena.p[n].cpu().numpy().reshape(-1)), _bits(want_p)), n
        if qm[n]:
            n_q += 1
            codes, inv = mu[n]
            assert codes.shape == (g.size // 64, 64) and inv.shape == (g.size // 64, 1)
            assert np.array_equal(codes.cpu().numpy(), want_s[0]), n
            assert np.array_equal(_bits(inv.cpu().numpy()), _bits(want_s[1])), n
    assert n_q > 0
