"""GPU parity of the full denoising train step against the CPU oracle (oracle/unet.py + oracle/lion.py + oracle/sched.py)
on the same random-init weights, synthetic latents, noise, timesteps and context (SURVEY.md 8(d) seeds).

Two criteria:
 1. loss and gradients vs. the fp64 oracle within k x the error of the oracle's own Flax-bf16 emulation
    (tolerance derived, not guessed: the reference itself computes in bf16; k = 3 for gradients, and an absolute
    floor of 2e-3 relative for the loss);
 2. given the GPU's fp32 gradients as input, the Lion int8 codes, reciprocal-absmax scales and updated fp32 masters
    are bit-exact against the NumPy restatement of lion_quant.py.
"""
import numpy as np
import pytest
import torch

from oracle import lion as OL
from oracle import sched as OS
from oracle import unet as OU
from stable_diffusion_training_b200 import lion_quant as LQ
from stable_diffusion_training_b200 import training_utils as TU
from stable_diffusion_training_b200 import unet as PU

pytestmark = pytest.mark.gpu


def _batch(B, H, W, L, D, seed=1):
    rng = np.random.default_rng(seed)
    return {
        "latents": rng.standard_normal((B, 4, H, W), dtype=np.float32),
        "noise": rng.standard_normal((B, 4, H, W), dtype=np.float32),
        "timesteps": rng.integers(0, 1000, B).astype(np.int32),
        "encoder_hidden_states": rng.standard_normal((B, L, D), dtype=np.float32),
    }


def _rel(a, b):
    a, b = np.asarray(a, np.float64).ravel(), np.asarray(b, np.float64).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


@pytest.mark.parametrize("ptype,Hh,L", [("epsilon", 32, 77), ("v_prediction", 16, 77), ("epsilon", 8, 77)])
def test_tiny_train_step_matches_oracle(cuda, ptype, Hh, L):
    cfg = PU.TINY
    model = TU.UNetModel(cfg)
    params_np = PU.init_params_numpy(cfg, 0)
    tree = LQ.unflatten_dict(params_np)
    tx = TU.create_lion_optimizer(tree)
    state = TU.TrainState.create(apply_fn=model.apply, params=tree, tx=tx, device=cuda)
    sched = TU.default_scheduler(cuda, "scaled_linear", ptype)
    b = _batch(2, Hh, Hh, L, cfg.cross_attention_dim)
    batch = {k: torch.from_numpy(v).to(cuda) for k, v in b.items()}

    p_before = {n: state.arena.p[n].detach().cpu().numpy().copy() for n in params_np}
    state, metrics = TU.train_step(state, batch, rng=0, frozen_noise_scheduler_state=sched)
    torch.cuda.synchronize()
    loss_gpu = float(metrics["loss"])
    g_gpu = {n: state.arena.g[n].detach().cpu().numpy().copy() for n in params_np}

    # ---- criterion 1: oracle in fp64 (truth) and Flax-bf16 emulation (yardstick) ----
    ac = OS.common_state("scaled_linear", 0.00085, 0.012)["alphas_cumprod"]
    op = {n: torch.from_numpy(v) for n, v in params_np.items()}
    args = (torch.from_numpy(b["latents"]), torch.from_numpy(b["noise"]), torch.from_numpy(b["timesteps"]).long(),
            torch.from_numpy(b["encoder_hidden_states"]), ac)
    loss64, g64, _ = OU.loss_and_grads(OU.TINY, op, *args, prediction_type=ptype, mode="fp64")
    loss16, g16, _ = OU.loss_and_grads(OU.TINY, op, *args, prediction_type=ptype, mode="bf16")
    names = sorted(params_np)
    v64 = np.concatenate([g64[n].numpy().ravel() for n in names])
    v16 = np.concatenate([g16[n].numpy().ravel() for n in names])
    vgpu = np.concatenate([g_gpu[n].ravel() for n in names])
    yard = _rel(v16, v64)
    err = _rel(vgpu, v64)
    print(f"loss gpu {loss_gpu:.6f} fp64 {float(loss64):.6f} bf16-emu {float(loss16):.6f}; grad rel err gpu {err:.4e} "
          f"bf16-emu yardstick {yard:.4e}")
    assert abs(loss_gpu - float(loss64)) <= max(3 * abs(float(loss16) - float(loss64)), 2e-3 * abs(float(loss64)))
    assert err <= 3 * yard, (err, yard)
    # per-leaf check on the big leaves: no leaf may be wildly off (catches a wrong-wiring bug that a global norm hides)
    # (absolute errors: some leaves have an exactly-zero true gradient, e.g. time_emb_proj in front of a GroupNorm
    #  with one channel per group, where only rounding noise remains)
    gscale = float(np.linalg.norm(v64))
    for n in names:
        if g64[n].numel() >= 1024:
            t = g64[n].numpy().astype(np.float64)
            e = float(np.linalg.norm(g_gpu[n].astype(np.float64) - t))
            y = float(np.linalg.norm(g16[n].numpy().astype(np.float64) - t))
            assert e <= max(6 * y, 0.05 * float(np.linalg.norm(t)), 1e-3 * gscale), (n, e, y)

    # ---- criterion 2: optimizer bit-exact given the GPU's own fp32 gradients ----
    gn = np.float32(state.arena.gnorm.item())
    gn64 = np.sqrt(sum(float(np.sum(g_gpu[n].astype(np.float64) ** 2)) for n in names))
    assert abs(float(gn) - gn64) <= 1e-6 * gn64
    mu = LQ.flatten_dict(state.opt_state[1].lion.mu_quant)
    wd = LQ.flatten_dict(TU.create_mask(tree, ["bias", "scale", "embedding"]))
    qm = LQ.flatten_dict(TU.create_mask(tree, list(TU.TrainingConfig().excluded_layer_from_quantization)))
    n_q = 0
    for n in names:
        g = OL.clip_leaf(g_gpu[n], gn, 1.0)
        st0 = OL.init_state((g.size,), bool(qm[n]))
        want_p, want_s, _ = OL.lion_leaf_update(g.reshape(-1), p_before[n].reshape(-1), st0, wd_flag=bool(wd[n]))
        got_p = state.arena.p[n].detach().cpu().numpy().reshape(-1)
        assert np.array_equal(got_p.view(np.uint32), want_p.view(np.uint32)), n
        if qm[n]:
            n_q += 1
            codes, inv = mu[n]
            assert codes.dtype == torch.int8 and codes.shape == (g.size // 16, 16) and inv.shape == (g.size // 16, 1)
            assert np.array_equal(codes.cpu().numpy(), want_s[0]), n
            assert np.array_equal(inv.cpu().numpy().view(np.uint32), want_s[1].view(np.uint32)), n
        else:
            assert np.array_equal(mu[n].cpu().numpy().reshape(-1).view(np.uint32), want_s.reshape(-1).view(np.uint32)), n
    assert n_q > 0 and int(state.step) == 1
    # bf16 GEMM operands follow the masters
    for n in names[:20]:
        assert torch.equal(state.arena.w[n], state.arena.p[n].to(torch.bfloat16))


def test_attention_quirk_matches_literal_chunking(cuda):
    """The additive log2(multiplicity) bias == the patched reference's clamped key chunks (oracle does them literally)."""
    from stable_diffusion_training_b200 import ops
    import math
    B, heads, Sq, Skv, d = 2, 2, 64, 77, 32
    g = torch.Generator().manual_seed(0)
    q = torch.randn(B, Sq, heads * d, generator=g).to(torch.bfloat16)
    k = torch.randn(B, Skv, heads * d, generator=g).to(torch.bfloat16)
    v = torch.randn(B, Skv, heads * d, generator=g).to(torch.bfloat16)
    m = PU.key_multiplicity(Sq, Skv)
    kb = torch.tensor([math.log2(x) for x in m], dtype=torch.float32, device=cuda)
    o, _ = ops.attention_fwd(q.to(cuda), k.to(cuda), v.to(cuda), heads, kbias2=kb)
    c = OU._Ctx({}, "fp64")

    def split(t):
        return t.double().reshape(B, -1, heads, d).permute(0, 2, 1, 3).reshape(B * heads, -1, d)
    ref = OU._chunked_attention(c, split(q), split(k), split(v))
    ref = ref.reshape(B, heads, Sq, d).permute(0, 2, 1, 3).reshape(B, Sq, heads * d)
    assert (o.cpu().double() - ref).abs().max().item() < 2e-2
    # and the quirk matters: without the bias the result differs
    o2, _ = ops.attention_fwd(q.to(cuda), k.to(cuda), v.to(cuda), heads)
    assert (o2.cpu().double() - ref).abs().max().item() > 2e-2


def test_compiled_train_step_matches_eager(cuda):
    """CompiledTrainStep (CUDA-graph replay, the counterpart of the reference's per-shape AOT executables) gives the
    same losses and parameters as eager train_step over 4 steps on rotating batches. Not bit-exact: the split-K weight
    gradients accumulate with red.add in a launch-dependent order (fp32 sums, ~1e-6 relative)."""
    cfg = PU.TINY
    params_np = PU.init_params_numpy(cfg, 0)
    sched = TU.default_scheduler(cuda, "scaled_linear", "epsilon")
    batches = [{k: torch.from_numpy(v).to(cuda) for k, v in _batch(2, 16, 16, 77, cfg.cross_attention_dim, seed=s).items()}
               for s in (1, 2, 3, 4)]
    out = {}
    for mode in ("eager", "graph"):
        tree = LQ.unflatten_dict({k: v.copy() for k, v in params_np.items()})
        model = TU.UNetModel(cfg)
        state = TU.TrainState.create(apply_fn=model.apply, params=tree, tx=TU.create_lion_optimizer(tree), device=cuda)
        step = TU.compile_train_step(state, sched) if mode == "graph" else \
            (lambda s, b, r: TU.train_step(s, b, r, frozen_noise_scheduler_state=sched))
        losses = []
        for b in batches:
            state, m = step(state, b, 0)
            losses.append(float(m["loss"]))
        torch.cuda.synchronize()
        out[mode] = (losses, state.arena.master.detach().cpu().numpy().copy(), int(state.step),
                     state.arena.codes.detach().cpu().numpy().copy())
    le, pe, se, ce = out["eager"]
    lg, pg, sg, cg = out["graph"]
    assert se == sg == 4
    np.testing.assert_allclose(lg, le, rtol=2e-3)
    # Lion moves every weight by +-lr per step: parameters may differ only where a near-zero momentum flipped sign
    assert np.mean(pe != pg) < 1e-2
    np.testing.assert_allclose(pg, pe, rtol=0, atol=4 * 4 * (1e-6 / 7) * 1.1)
    assert np.mean(ce != cg) < 1e-2


def test_sd2_style_train_step_matches_oracle(cuda):
    """SD-2.x wiring (BASELINE.json configs[3]) at a size the CPU oracle finishes in seconds: per-level head counts with
    head dim 64, Dense proj_in / proj_out (use_linear_projection), two layers per block, v-prediction, ctx dim != C."""
    kw = dict(block_out_channels=(64, 128), down_block_types=("CrossAttnDownBlock2D", "DownBlock2D"),
              up_block_types=("UpBlock2D", "CrossAttnUpBlock2D"), layers_per_block=2, attention_head_dim=(1, 2),
              cross_attention_dim=48, use_linear_projection=True)
    cfg, ocfg = PU.UNetConfig(**kw), OU.UNetConfig(**kw)
    ptype = "v_prediction"
    params_np = PU.init_params_numpy(cfg, 0)
    tree = LQ.unflatten_dict(params_np)
    state = TU.TrainState.create(apply_fn=TU.UNetModel(cfg).apply, params=tree, tx=TU.create_lion_optimizer(tree),
                                 device=cuda)
    sched = TU.default_scheduler(cuda, "scaled_linear", ptype)
    b = _batch(2, 16, 16, 77, cfg.cross_attention_dim, seed=3)
    batch = {k: torch.from_numpy(v).to(cuda) for k, v in b.items()}
    state, metrics = TU.train_step(state, batch, rng=0, frozen_noise_scheduler_state=sched)
    torch.cuda.synchronize()
    names = sorted(params_np)
    g_gpu = {n: state.arena.g[n].detach().cpu().numpy() for n in names}
    ac = OS.common_state("scaled_linear", 0.00085, 0.012)["alphas_cumprod"]
    op = {n: torch.from_numpy(v) for n, v in params_np.items()}
    args = (torch.from_numpy(b["latents"]), torch.from_numpy(b["noise"]), torch.from_numpy(b["timesteps"]).long(),
            torch.from_numpy(b["encoder_hidden_states"]), ac)
    loss64, g64, _ = OU.loss_and_grads(ocfg, op, *args, prediction_type=ptype, mode="fp64")
    loss16, g16, _ = OU.loss_and_grads(ocfg, op, *args, prediction_type=ptype, mode="bf16")
    v64 = np.concatenate([g64[n].numpy().ravel() for n in names])
    v16 = np.concatenate([g16[n].numpy().ravel() for n in names])
    vgpu = np.concatenate([g_gpu[n].ravel() for n in names])
    yard, err = _rel(v16, v64), _rel(vgpu, v64)
    print(f"loss gpu {float(metrics['loss']):.6f} fp64 {float(loss64):.6f}; grad rel err {err:.4e} yardstick {yard:.4e}")
    assert abs(float(metrics["loss"]) - float(loss64)) <= max(3 * abs(float(loss16) - float(loss64)), 2e-3 * abs(float(loss64)))
    assert err <= 3 * yard, (err, yard)
