"""GPU parity of the text-encoder path against oracle/clip.py — itself pinned against transformers'
PyTorch CLIPTextModel (tests/test_oracle_clip.py):

 * the kernels of csrc/clip.cu one by one vs torch fp32 references (embeddings with repeated tokens, 77-token causal
   attention forward / backward, quick-GELU / GELU, context gather / scatter);
 * the encoder forward + backward (tiny config and the real SD-1.5 text tower, 123M parameters) vs the fp64 oracle with
   the bf16-emulation yardstick (k = 3, as for the UNet);
 * the JOINT train step of /root/reference/training_utils.py:504-762 — text encoder inside the step, context assembly
   with BOS / EOS stripping (641-674), UNet, loss, both backward passes, both Lion updates (732-733) — loss and BOTH
   gradient sets vs the oracle; compiled (CUDA graph) == eager.
"""

import math

import numpy as np
import pytest
import torch

from oracle import clip as OC
from oracle import sched as OS
from oracle import unet as OU
from stable_diffusion_training_b200 import clip as CL
from stable_diffusion_training_b200 import lion_quant as LQ
from stable_diffusion_training_b200 import ops
from stable_diffusion_training_b200 import training_utils as TU
from stable_diffusion_training_b200 import unet as PU

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = np.asarray(a, np.float64).ravel(), np.asarray(b, np.float64).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


@pytest.mark.parametrize("N,H,S,d,causal", [(3, 12, 77, 64, True), (2, 2, 77, 16, True), (2, 4, 50, 32, False),
                                            (1, 1, 96, 64, True)])
def test_attention_small_matches_torch(cuda, N, H, S, d, causal):
    g = torch.Generator().manual_seed(S + d)
    C = H * d
    q, k, v, dO = (torch.randn(N * S, C, generator=g).to(torch.bfloat16).to(cuda) for _ in range(4))
    scale = d ** -0.5
    o, lse = ops.attention_small_fwd(q, k, v, N, H, scale, causal)
    dq, dk, dv = ops.attention_small_bwd(q, k, v, o, dO, lse, N, H, scale, causal)
    torch.cuda.synchronize()

    def heads(t):
        return t.float().cpu().reshape(N, S, H, d).permute(0, 2, 1, 3).requires_grad_(True)
    qh, kh, vh = heads(q), heads(k), heads(v)
    ref = torch.nn.functional.scaled_dot_product_attention(qh, kh, vh, is_causal=causal)
    ref.backward(heads(dO).detach())

    def back(t):
        return t.permute(0, 2, 1, 3).reshape(N * S, C)
    assert (o.float().cpu() - back(ref.detach())).abs().max().item() < 2e-2
    for got, want in ((dq, qh.grad), (dk, kh.grad), (dv, vh.grad)):
        assert _rel(got.float().cpu().numpy(), back(want).numpy()) < 2e-2


@pytest.mark.parametrize("kind", ["quick_gelu", "gelu"])
def test_activation_kernels(cuda, kind):
    x = (torch.randn(4096, 96) * 2).to(torch.bfloat16).to(cuda)
    dy = torch.randn(4096, 96).to(torch.bfloat16).to(cuda)
    xr = x.float().cpu().requires_grad_(True)
    yr = xr * torch.sigmoid(1.702 * xr) if kind == "quick_gelu" else torch.nn.functional.gelu(xr)
    yr.backward(dy.float().cpu())
    assert (ops.act_fwd(x, kind).float().cpu() - yr.detach()).abs().max().item() < 2e-2
    assert (ops.act_bwd(x, dy, kind).float().cpu() - xr.grad).abs().max().item() < 3e-2


def test_embeddings_and_context_gather(cuda):
    vocab, D, N, S = 300, 64, 4, 77
    g = torch.Generator().manual_seed(1)
    tok = (torch.randn(vocab, D, generator=g) * 0.02)
    pos = (torch.randn(S, D, generator=g) * 0.02)
    ids = torch.randint(0, vocab, (N, S), generator=g, dtype=torch.int32)
    ids[:, 0] = 7  # repeated token: atomics path
    ids[1] = ids[0]
    out = ops.embed_fwd(ids.to(cuda), tok.to(torch.bfloat16).to(cuda), pos.to(torch.bfloat16).to(cuda))
    want = (tok.to(torch.bfloat16).float()[ids.long()] + pos.to(torch.bfloat16).float()[None]).to(torch.bfloat16)
    assert torch.equal(out.cpu().view(N, S, D), want)
    dout = torch.randn(N * S, D, generator=g).to(torch.bfloat16)
    dtok = torch.full((vocab, D), 5.0, device=cuda)
    dpos = torch.full((S, D), 5.0, device=cuda)
    ops.embed_bwd(ids.to(cuda), dout.to(cuda), dtok, dpos)
    ref_tok = torch.zeros(vocab, D).index_add_(0, ids.long().reshape(-1), dout.float())
    assert (dtok.cpu() - ref_tok).abs().max().item() < 1e-4
    assert (dpos.cpu() - dout.float().view(N, S, D).sum(0)).abs().max().item() < 1e-4
    # context assembly: gather == oracle.assemble_context, scatter is its adjoint
    for chunks, strip in ((1, True), (2, True), (2, False)):
        B = N // chunks
        hid = torch.randn(B * chunks * 77, D, generator=g).to(torch.bfloat16)
        idx = torch.from_numpy(CL.context_row_index(B, chunks, strip).reshape(-1).copy()).to(cuda)
        ctx = ops.gather_rows(hid.to(cuda), idx, idx.numel())
        want = OC.assemble_context(hid.view(B * chunks, 77, D), B, strip)
        assert torch.equal(ctx.cpu().view(want.shape), want)
        back = ops.scatter_rows(ctx, idx, hid.shape[0]).cpu()
        mask = torch.zeros(hid.shape[0], dtype=torch.bool)
        mask[idx.cpu().long()] = True
        assert torch.equal(back[mask], hid[mask]) and (mask.all() or back[~mask].abs().max().item() == 0)


def _text_state(cuda, cfg, seed=1):
    params_np = CL.init_params_numpy(cfg, seed)
    tree = LQ.unflatten_dict({k: v.copy() for k, v in params_np.items()})
    model = CL.CLIPTextModel(cfg)
    state = TU.TrainState.create(apply_fn=model.__call__, params=tree, tx=TU.create_lion_optimizer(tree), device=cuda)
    return params_np, model, state


@pytest.mark.parametrize("name", ["tiny", "sd15"])
def test_text_encoder_forward_backward_matches_oracle(cuda, name):
    pcfg, ocfg = (CL.TINY_CLIP, OC.TINY_CLIP) if name == "tiny" else (CL.SD15_CLIP, OC.SD15_CLIP)
    params_np, model, state = _text_state(cuda, pcfg)
    assert CL.param_shapes(pcfg) == OC.param_shapes(ocfg)
    g = torch.Generator().manual_seed(4)
    N = 3
    ids = torch.randint(0, pcfg.vocab_size, (N, 77), generator=g, dtype=torch.int32)
    dO = torch.randn(N * 77, pcfg.hidden_size, generator=g)
    out = state.net.forward(ids.to(cuda))
    out.g = dO.to(torch.bfloat16).to(cuda)
    hid = out.v.float().cpu()
    state.net.backward()
    torch.cuda.synchronize()
    names = sorted(params_np)
    res = {}
    for mode in ("fp64", "bf16"):
        leaves = {k: torch.from_numpy(v).to(torch.float64 if mode == "fp64" else torch.float32).requires_grad_(True)
                  for k, v in params_np.items()}
        h = OC.text_encoder_forward(ocfg, leaves, ids, mode=mode)
        (h * dO.to(torch.bfloat16).to(h.dtype).view(N, 77, -1)).sum().backward()
        res[mode] = (h.detach().double().reshape(N * 77, -1),
                     np.concatenate([leaves[n].grad.double().numpy().ravel() for n in names]))
    vg = np.concatenate([state.arena.g[n].detach().cpu().double().numpy().ravel() for n in names])
    h_err, h_yard = _rel(hid.numpy(), res["fp64"][0].numpy()), _rel(res["bf16"][0].numpy(), res["fp64"][0].numpy())
    g_err, g_yard = _rel(vg, res["fp64"][1]), _rel(res["bf16"][1], res["fp64"][1])
    print(f"\nCLIP text {name}: hidden err {h_err:.3e} (bf16-emu yardstick {h_yard:.3e}); grad err {g_err:.3e} "
          f"(yardstick {g_yard:.3e})")
    assert h_err <= 3 * h_yard and g_err <= 3 * g_yard
    # the forward-only model call of the reference surface
    lhs = model(input_ids=ids.to(cuda), params=state.params)[0]
    assert lhs.shape == (N, 77, pcfg.hidden_size) and _rel(lhs.cpu().numpy(), res["fp64"][0].numpy()) <= 3 * h_yard


def _joint_batch(B, chunks, H, vocab, seed):
    rng = np.random.default_rng(seed)
    return {"latents": rng.standard_normal((B, 4, H, H), dtype=np.float32),
            "noise": rng.standard_normal((B, 4, H, H), dtype=np.float32),
            "timesteps": rng.integers(0, 1000, B).astype(np.int32),
            "input_ids": rng.integers(0, vocab, (B * chunks, 77)).astype(np.int32)}


@pytest.mark.parametrize("chunks,strip", [(3, True), (1, False)])
def test_joint_train_step_matches_oracle(cuda, chunks, strip):
    ucfg, tcfg = PU.TINY, CL.TINY_CLIP   # cross_attention_dim 32 == hidden_size 32
    up_np = PU.init_params_numpy(ucfg, 0)
    utree = LQ.unflatten_dict({k: v.copy() for k, v in up_np.items()})
    ustate = TU.TrainState.create(apply_fn=TU.UNetModel(ucfg).apply, params=utree, tx=TU.create_lion_optimizer(utree),
                                  device=cuda)
    tp_np, _, tstate = _text_state(cuda, tcfg)
    t_before = tstate.arena.master.clone()
    sched = TU.default_scheduler(cuda, "scaled_linear", "epsilon")
    B, H = 2, 16
    b = _joint_batch(B, chunks, H, tcfg.vocab_size, 5)
    batch = {k: torch.from_numpy(v).to(cuda) for k, v in b.items()}
    ustate, m = TU.train_step(ustate, batch, None, sched, text_encoder_state=tstate, strip_bos_eos_token=strip)
    torch.cuda.synchronize()
    assert int(ustate.step) == 1 and int(tstate.step) == 1
    # the text encoder was updated (token-embedding rows that do not occur in the batch have zero gradient and, being
    # exempt from weight decay, stay put: Lion's sign(0) = 0)
    assert float((tstate.arena.master != t_before).float().mean()) > 0.2

    ac = OS.common_state("scaled_linear", 0.00085, 0.012)["alphas_cumprod"]
    noisy = torch.from_numpy(OS.add_noise(ac, b["latents"], b["noise"], b["timesteps"]))
    ids = torch.from_numpy(b["input_ids"])
    out = {}
    for mode in ("fp64", "bf16"):
        dt = torch.float64 if mode == "fp64" else torch.float32
        ul = {k: torch.from_numpy(v).to(dt).requires_grad_(True) for k, v in up_np.items()}
        tl = {k: torch.from_numpy(v).to(dt).requires_grad_(True) for k, v in tp_np.items()}
        ctx = OC.assemble_context(OC.text_encoder_forward(OC.TINY_CLIP, tl, ids, mode=mode), B, strip)
        pred = OU.unet_forward(OU.TINY, ul, noisy, torch.from_numpy(b["timesteps"]).long(), ctx, mode)
        loss = ((torch.from_numpy(b["noise"]).to(pred.dtype) - pred) ** 2).mean()
        loss.backward()
        out[mode] = (float(loss), np.concatenate([ul[n].grad.double().numpy().ravel() for n in sorted(up_np)]),
                     np.concatenate([tl[n].grad.double().numpy().ravel() for n in sorted(tp_np)]))
    L = ctx.shape[1]
    assert L == (227 if (chunks, strip) == (3, True) else 77)
    ug = np.concatenate([ustate.arena.g[n].detach().cpu().double().numpy().ravel() for n in sorted(up_np)])
    tg = np.concatenate([tstate.arena.g[n].detach().cpu().double().numpy().ravel() for n in sorted(tp_np)])
    loss = float(m["loss"])
    ue, uy = _rel(ug, out["fp64"][1]), _rel(out["bf16"][1], out["fp64"][1])
    te, ty = _rel(tg, out["fp64"][2]), _rel(out["bf16"][2], out["fp64"][2])
    print(f"\njoint step (chunks {chunks}, strip {strip}, ctx {L}): loss gpu {loss:.6f} fp64 {out['fp64'][0]:.6f}; "
          f"UNet grad err {ue:.3e} (yardstick {uy:.3e}); text-encoder grad err {te:.3e} (yardstick {ty:.3e})")
    assert abs(loss - out["fp64"][0]) <= max(3 * abs(out["bf16"][0] - out["fp64"][0]), 2e-3 * abs(out["fp64"][0]))
    assert ue <= 3 * uy and te <= 3 * ty


def test_joint_compiled_step_matches_eager(cuda):
    ucfg, tcfg = PU.TINY, CL.TINY_CLIP
    up_np = PU.init_params_numpy(ucfg, 0)
    sched = TU.default_scheduler(cuda, "scaled_linear", "epsilon")
    batches = [{k: torch.from_numpy(v).to(cuda) for k, v in _joint_batch(2, 3, 16, tcfg.vocab_size, s).items()}
               for s in (1, 2, 3)]
    res = {}
    for mode in ("eager", "graph"):
        utree = LQ.unflatten_dict({k: v.copy() for k, v in up_np.items()})
        ustate = TU.TrainState.create(apply_fn=TU.UNetModel(ucfg).apply, params=utree,
                                      tx=TU.create_lion_optimizer(utree), device=cuda)
        _, _, tstate = _text_state(cuda, tcfg)
        step = TU.compile_train_step(ustate, sched, text_encoder_state=tstate) if mode == "graph" else \
            (lambda s, b, r: TU.train_step(s, b, r, sched, text_encoder_state=tstate))
        losses = []
        for b in batches:
            ustate, m = step(ustate, b, None)
            losses.append(float(m["loss"]))
        torch.cuda.synchronize()
        res[mode] = (losses, ustate.arena.master.clone(), tstate.arena.master.clone(), int(tstate.step))
    (le, ue, te, se), (lg, ug, tg, sg) = res["eager"], res["graph"]
    assert se == sg == 3
    np.testing.assert_allclose(lg, le, rtol=1e-5)
    assert float((ue != ug).float().mean()) < 1e-2 and float((te != tg).float().mean()) < 1e-2


@pytest.mark.parametrize("mode", ["eager", "graph", "foreign"])
def test_text_encoder_ema_follows_the_reference_rule(cuda, mode):
    """training_utils.py:740-746: new_text_encoder_ema = rate * ema + (1 - rate) * new_params, every step; fused into
    the text encoder's optimizer launches (eager / graph) or leaf by leaf for a tree that is not the arena's own."""
    from oracle import lion as OL
    ucfg, tcfg = PU.TINY, CL.TINY_CLIP
    rate = 0.9
    utree = LQ.unflatten_dict(PU.init_params_numpy(ucfg, 0))
    ustate = TU.TrainState.create(apply_fn=TU.UNetModel(ucfg).apply, params=utree, tx=TU.create_lion_optimizer(utree),
                                  device=cuda)
    tp_np, _, tstate = _text_state(cuda, tcfg)
    sched = TU.default_scheduler(cuda, "scaled_linear", "epsilon")
    if mode == "foreign":
        ema_tree = LQ.unflatten_dict({k: torch.from_numpy(v.copy()).to(cuda) for k, v in tp_np.items()})
    else:
        ema_tree = tstate.create_ema()
    flags = dict(text_encoder_state=tstate, text_encoder_ema_params=ema_tree, ema_rate=rate)
    step = TU.compile_train_step(ustate, sched, **flags) if mode == "graph" else \
        (lambda s, b, r: TU.train_step(s, b, r, sched, **flags))
    want = {k: v.copy() for k, v in tp_np.items()}
    for s in (1, 2, 3):
        b = {k: torch.from_numpy(v).to(cuda) for k, v in _joint_batch(2, 2, 16, tcfg.vocab_size, s).items()}
        ustate, _ = step(ustate, b, None)
        torch.cuda.synchronize()
        new = {k: v.cpu().numpy() for k, v in LQ.flatten_dict(tstate.params).items()}
        want = {k: OL.ema_update(want[k], new[k], rate) for k in want}
    got = {k: v.cpu().numpy() for k, v in LQ.flatten_dict(ema_tree).items()}
    moved = 0
    for k in want:
        assert np.array_equal(got[k].view(np.uint32), want[k].view(np.uint32)), k
        moved += int(np.any(got[k] != tp_np[k]))
    assert moved > len(want) // 2
    if mode == "graph":
        with pytest.raises(ValueError, match="text_encoder_ema_params"):
            TU.compile_train_step(ustate, sched, text_encoder_state=tstate, text_encoder_ema_params={"x": 1}, ema_rate=rate)


def test_create_lion_optimizer_states_mirrors_the_reference_factory(cuda):
    """training_utils.py:281-428: {"unet_state", "text_encoder_state"} from the models dict; the states it builds step
    exactly like states built by hand with the same hyper-parameters; without quantisation every leaf keeps an fp32
    momentum (the reference's optax.lion branch)."""
    ucfg, tcfg = PU.TINY, CL.TINY_CLIP
    cfgd = TU.TrainingConfig()

    def models():
        return {"unet": {"unet_model": TU.UNetModel(ucfg), "unet_params": LQ.unflatten_dict(PU.init_params_numpy(ucfg, 0))},
                "text_encoder": {"text_encoder_model": CL.CLIPTextModel(tcfg),
                                 "text_encoder_params": LQ.unflatten_dict(CL.init_params_numpy(tcfg, 0))}}

    kw = dict(excluded_layer_pattern_from_weight_decay=list(cfgd.excluded_layer_pattern_from_weight_decay),
              excluded_layer_from_quantization=list(cfgd.excluded_layer_from_quantization), lion_8bit_block_size=16,
              device=cuda)
    st = TU.create_lion_optimizer_states(models(), quantize_unet_state=True, quantize_text_encoder_state=True, **kw)
    assert set(st) == {"unet_state", "text_encoder_state"}
    ua, ta = st["unet_state"], st["text_encoder_state"]
    assert ua.arena.n_quant > 0 and ta.arena.n_quant > 0
    only_unet = TU.create_lion_optimizer_states(models(), train_text_encoder=False, quantize_unet_state=True, **kw)
    assert only_unet["text_encoder_state"] is None and only_unet["unet_state"] is not None
    plain = TU.create_lion_optimizer_states(models(), **kw)   # quantize_* default False
    assert plain["unet_state"].arena.n_quant == 0 and plain["text_encoder_state"].arena.n_quant == 0
    mu = LQ.flatten_dict(plain["unet_state"].opt_state[1].lion.mu_quant)
    assert all(isinstance(v, torch.Tensor) and v.dtype == torch.float32 for v in mu.values())
    # hand-built reference states
    utree = LQ.unflatten_dict(PU.init_params_numpy(ucfg, 0))
    ub = TU.TrainState.create(apply_fn=TU.UNetModel(ucfg).apply, params=utree, tx=TU.create_lion_optimizer(utree),
                              device=cuda)
    _, _, tb = _text_state(cuda, tcfg, seed=0)
    sched = TU.default_scheduler(cuda, "scaled_linear", "epsilon")
    for s in (1, 2):
        b = {k: torch.from_numpy(v).to(cuda) for k, v in _joint_batch(2, 2, 16, tcfg.vocab_size, s).items()}
        ua, ma = TU.train_step(ua, b, None, sched, text_encoder_state=ta)
        ub, mb = TU.train_step(ub, b, None, sched, text_encoder_state=tb)
    torch.cuda.synchronize()
    assert float(ma["loss"]) == float(mb["loss"])
    assert float((ua.arena.master != ub.arena.master).float().mean()) < 1e-3   # red.add weight gradients: 1e-7 noise
    assert float((ta.arena.master != tb.arena.master).float().mean()) < 1e-3
    # the unquantised states train too
    pu, pt = plain["unet_state"], plain["text_encoder_state"]
    pu, mp = TU.train_step(pu, b, None, sched, text_encoder_state=pt)
    torch.cuda.synchronize()
    assert math.isfinite(float(mp["loss"])) and int(pt.step) == 1
