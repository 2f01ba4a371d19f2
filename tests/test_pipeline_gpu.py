"""GPU parity of the validation sampler (SURVEY.md 8(f)-4;
/root/reference/models/pipeline_flax_stable_diffusion.py:160-254) against oracle/pipeline.py and oracle/vae.py: the three
elementwise kernels bit-exact, the VAE decoder at toy and REAL widths (512 / 512 / 256 / 128 channels) vs the fp64
oracle with the bf16-emulation yardstick, the whole pipeline (text encoder -> Threefry latents -> DDIM loop with
classifier-free guidance -> decode) on the tiny models, eager == CUDA-graph loop body, and executor reuse from a live
TrainState."""
import numpy as np
import pytest
import torch

from oracle import clip as OC
from oracle import pipeline as OP
from oracle import threefry as OT
from oracle import unet as OU
from oracle import vae as OV
from stable_diffusion_training_b200 import _lib
from stable_diffusion_training_b200 import clip as PC
from stable_diffusion_training_b200 import jax_random as JR
from stable_diffusion_training_b200 import lion_quant as LQ
from stable_diffusion_training_b200 import pipeline as PP
from stable_diffusion_training_b200 import schedulers as PS
from stable_diffusion_training_b200 import training_utils as TU
from stable_diffusion_training_b200 import unet as PU
from stable_diffusion_training_b200 import vae as PV

pytestmark = pytest.mark.gpu

O_TINY_UNET = OU.UNetConfig(block_out_channels=(32, 64), down_block_types=("CrossAttnDownBlock2D", "DownBlock2D"),
                            up_block_types=("UpBlock2D", "CrossAttnUpBlock2D"), layers_per_block=1, attention_head_dim=2,
                            cross_attention_dim=32)


def _rel(a, b):
    a, b = np.asarray(a, np.float64).ravel(), np.asarray(b, np.float64).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def _bf(x):
    return x.to(torch.bfloat16).float()


def test_sampler_elementwise_kernels_bit_exact(cuda):
    l = _lib.lib()
    s = _lib.stream_ptr()
    g = torch.Generator().manual_seed(0)
    B, C, h, w = 3, 4, 5, 7
    lat = torch.randn(B, C, h, w, generator=g) * 3
    lat_d = lat.to(cuda)
    out = torch.empty(2 * B, h, w, C, dtype=torch.bfloat16, device=cuda)
    sc = float(np.float32(1.0 / 0.18215))
    _lib.check(l.sdb_latents_to_nhwc_bf16(lat_d.data_ptr(), sc, 2, out.data_ptr(), B, C, h * w, s), "latents_to_nhwc")
    want = (lat * torch.tensor(sc)).to(torch.bfloat16).permute(0, 2, 3, 1)
    assert torch.equal(out.cpu()[:B], want) and torch.equal(out.cpu()[B:], want)
    # classifier-free guidance
    pred = (torch.randn(2 * B, h, w, C, generator=g) * 2).to(torch.bfloat16)
    pred_d = pred.to(cuda)
    comb = torch.empty(B, C, h, w, device=cuda)
    _lib.check(l.sdb_cfg_combine(pred_d.data_ptr(), 7.5, comb.data_ptr(), B, C, h * w, s), "cfg_combine")
    u, t = pred[:B].float(), pred[B:].float()
    want = (u + torch.tensor(7.5) * _bf(t - u)).permute(0, 3, 1, 2)
    assert torch.equal(comb.cpu(), want)
    # image post-processing (4-channel padded rows -> 3 channels)
    y = (torch.randn(B * h * w, 4, generator=g) * 1.5).to(torch.bfloat16)
    y_d = y.to(cuda)
    img = torch.empty(B * h * w, 3, device=cuda)
    _lib.check(l.sdb_image_postprocess(y_d.data_ptr(), img.data_ptr(), B * h * w, 4, 3, s), "image_postprocess")
    want = _bf(_bf(y[:, :3].float() / 2) + 0.5).clamp(0, 1)
    assert torch.equal(img.cpu(), want)
    assert float(img.min()) == 0.0 and float(img.max()) == 1.0


def _decoder(cuda, pcfg, seed=2):
    params_np = PV.init_decoder_params_numpy(pcfg, seed)
    tree = LQ.unflatten_dict(params_np)
    model = PV.AutoencoderKL(pcfg)
    return params_np, model, tree, model.decoder(tree, cuda)


@pytest.mark.parametrize("name,B,h,w", [("tiny", 2, 16, 20), ("sd", 1, 16, 16)])
def test_vae_decode_matches_oracle(cuda, name, B, h, w):
    pcfg, ocfg = (PV.TINY_VAE, OV.TINY_VAE) if name == "tiny" else (PV.SD_VAE, OV.SD_VAE)
    assert PV.decoder_param_shapes(pcfg) == OV.decoder_param_shapes(ocfg)
    params_np, model, tree, dec = _decoder(cuda, pcfg)
    g = torch.Generator().manual_seed(5)
    z = torch.randn(B, 4, h, w, generator=g)
    got = dec.decode(z.to(cuda))
    img = dec.decode_image(z.to(cuda))
    torch.cuda.synchronize()
    op = {k: torch.from_numpy(v) for k, v in params_np.items()}
    d64 = OV.decode(ocfg, op, z, mode="fp64")
    d16 = OV.decode(ocfg, op, z, mode="bf16")
    up = 2 ** (len(pcfg.block_out_channels) - 1)
    assert tuple(got.shape) == tuple(d64.shape) == (B, 3, h * up, w * up)
    err, yard = _rel(got.cpu().numpy(), d64.numpy()), _rel(d16.numpy(), d64.numpy())
    print(f"\nVAE decode {name} {h}x{w} -> {h * up}x{w * up}: sample err {err:.3e} (bf16-emu yardstick {yard:.3e})")
    assert err <= 3 * yard
    # the fused read-out is exactly the post-processing of the NCHW sample
    want = _bf(_bf(got.cpu() / 2) + 0.5).clamp(0, 1).permute(0, 2, 3, 1)
    assert torch.equal(img.cpu(), want)
    # 1 / scaling_factor folded into the entry transpose
    sc = 1.0 / 0.18215
    got_s = dec.decode(z.to(cuda), sc)
    ref_s = dec.decode((z * torch.tensor(np.float32(sc))).to(cuda))
    assert torch.equal(got_s, ref_s)
    with pytest.raises(ValueError):
        dec.decode(torch.zeros(1, 3, 8, 8, device=cuda))


class _Tok:
    """Stand-in with the CLIPTokenizer call signature (no vocabulary files on this box)."""
    model_max_length = 77

    def __init__(self, vocab):
        self.vocab = vocab

    def __call__(self, prompt, padding=None, max_length=None, truncation=None, return_tensors=None):
        prompt = [prompt] if isinstance(prompt, str) else prompt
        ids = np.full((len(prompt), max_length), self.vocab - 1, dtype=np.int32)
        ids[:, 0] = self.vocab - 2
        for i, p in enumerate(prompt):
            words = [hash(wd) % (self.vocab - 2) for wd in p.split()][: max_length - 2]
            ids[i, 1:1 + len(words)] = words
        from types import SimpleNamespace
        return SimpleNamespace(input_ids=ids)


def _tiny_pipeline(cuda, seed=3):
    up = PU.init_params_numpy(PU.TINY, seed)
    cp = PC.init_params_numpy(PC.TINY_CLIP, seed + 1)
    vp = PV.init_decoder_params_numpy(PV.TINY_VAE, seed + 2)
    params = {"unet": LQ.unflatten_dict(up), "text_encoder": LQ.unflatten_dict(cp), "vae": LQ.unflatten_dict(vp)}
    oparams = {"unet": {k: torch.from_numpy(v) for k, v in up.items()},
               "text_encoder": {k: torch.from_numpy(v) for k, v in cp.items()},
               "vae": {k: torch.from_numpy(v) for k, v in vp.items()}}
    sched = PS.FlaxDDIMScheduler(beta_start=0.00085, beta_end=0.012, beta_schedule="scaled_linear",
                                 num_train_timesteps=1000, prediction_type="epsilon", device=cuda)
    pipe = PP.FlaxStableDiffusionPipeline(vae=PV.AutoencoderKL(PV.TINY_VAE), text_encoder=PC.CLIPTextModel(PC.TINY_CLIP),
                                          tokenizer=_Tok(PC.TINY_CLIP.vocab_size), unet=TU.UNetModel(PU.TINY),
                                          scheduler=sched, device=cuda)
    return pipe, params, oparams


def test_pipeline_matches_oracle_tiny(cuda):
    pipe, params, oparams = _tiny_pipeline(cuda)
    r = np.random.default_rng(0)
    ids = r.integers(0, PC.TINY_CLIP.vocab_size, size=(2, 77)).astype(np.int32)
    neg = r.integers(0, PC.TINY_CLIP.vocab_size, size=(2, 77)).astype(np.int32)
    steps, H, W, gs = 4, 32, 48, 5.0
    out = pipe(ids, params, JR.PRNGKey(7), steps, H, W, gs, neg_prompt_ids=neg)
    img = out["images"]
    assert out["nsfw_content_detected"] is False
    assert isinstance(img, np.ndarray) and img.shape == (2, H, W, 3) and img.dtype == np.float32
    assert img.min() >= 0.0 and img.max() <= 1.0
    i64 = OP.generate(O_TINY_UNET, OC.TINY_CLIP, OV.TINY_VAE, oparams, ids, neg, OT.PRNGKey(7), steps, H, W, gs, mode="fp64")
    i16 = OP.generate(O_TINY_UNET, OC.TINY_CLIP, OV.TINY_VAE, oparams, ids, neg, OT.PRNGKey(7), steps, H, W, gs, mode="bf16")
    err, yard = _rel(img, i64.numpy()), _rel(i16.float().numpy(), i64.numpy())
    print(f"\npipeline tiny {steps} DDIM steps, guidance {gs}: image err {err:.3e} (bf16-emu yardstick {yard:.3e})")
    assert err <= 3 * yard
    assert pipe.launches_last_call > 0
    # given latents instead of a key; tuple output
    lat = JR.sample_noise(JR.PRNGKey(7), (2, 4, H // 2, W // 2), cuda).cpu().numpy()
    assert np.abs(lat - OT.normal(OT.PRNGKey(7), (2, 4, H // 2, W // 2))).max() < 1e-5
    img2, nsfw = pipe(ids, params, None, steps, H, W, gs, latents=lat, neg_prompt_ids=neg, return_dict=False)
    assert nsfw is False and np.array_equal(img2, img)
    # empty negative prompt through the tokenizer
    img3 = pipe(ids, params, JR.PRNGKey(7), 2, H, W, gs)["images"]
    uncond = pipe.tokenizer([""] * 2, padding="max_length", max_length=77, return_tensors="np").input_ids
    img4 = pipe(ids, params, JR.PRNGKey(7), 2, H, W, gs, neg_prompt_ids=uncond)["images"]
    assert np.array_equal(img3, img4)
    with pytest.raises(ValueError, match="Unexpected latents shape"):
        pipe(ids, params, None, 2, H, W, gs, latents=np.zeros((2, 4, 4, 4), np.float32), neg_prompt_ids=neg)


def test_pipeline_graph_loop_equals_eager_and_v_prediction(cuda):
    pipe, params, _ = _tiny_pipeline(cuda, seed=5)
    pipe.scheduler = PS.FlaxDDIMScheduler(beta_start=0.00085, beta_end=0.012, beta_schedule="scaled_linear",
                                          num_train_timesteps=1000, prediction_type="v_prediction", device=cuda)
    ids = pipe.prepare_inputs(["a photo of a cat", "a drawing of a dog on the moon"])
    assert ids.shape == (2, 77)
    for _ in range(2):   # the first call also pays one-time setup launches (cached key-bias rows, executors)
        eager = pipe(ids, params, JR.PRNGKey(1), 6, 32, 32, 7.5, output_type="pt")["images"]
    n_eager = pipe.launches_last_call
    graph = pipe(ids, params, JR.PRNGKey(1), 6, 32, 32, 7.5, jit=True, output_type="pt")["images"]
    assert torch.equal(eager, graph)
    assert pipe.launches_last_call == n_eager   # the replays enqueue what the eager loop launches
    again = pipe(ids, params, JR.PRNGKey(2), 6, 32, 32, 7.5, jit=True, output_type="pt")["images"]   # cached graph
    assert len(pipe._graphs) == 1 and not torch.equal(again, graph)
    want = pipe(ids, params, JR.PRNGKey(2), 6, 32, 32, 7.5, output_type="pt")["images"]
    assert torch.equal(again, want)


def test_pipeline_reuses_the_train_state_executors(cuda):
    """Validation between train steps: params that ARE a live TrainState's tree run on that state's bf16 operand copies."""
    pipe, params, _ = _tiny_pipeline(cuda, seed=8)
    ids = np.random.default_rng(3).integers(0, PC.TINY_CLIP.vocab_size, size=(1, 77)).astype(np.int32)
    neg = np.roll(ids, 5, axis=1).copy()
    base = pipe(ids, params, JR.PRNGKey(4), 3, 32, 32, 3.0, neg_prompt_ids=neg)["images"]
    ust = TU.TrainState.create(apply_fn=pipe.unet.apply, params=params["unet"],
                               tx=TU.create_lion_optimizer(params["unet"]), device=cuda)
    tst = TU.TrainState.create(apply_fn=pipe.text_encoder.__call__, params=params["text_encoder"],
                               tx=TU.create_lion_optimizer(params["text_encoder"]), device=cuda)
    live = {"unet": ust.params, "text_encoder": tst.params, "vae": params["vae"]}
    got = pipe(ids, live, JR.PRNGKey(4), 3, 32, 32, 3.0, neg_prompt_ids=neg)["images"]
    assert pipe._executor("unet", ust.params) is ust.net and pipe._executor("text_encoder", tst.params) is tst.net
    assert np.array_equal(got, base)
