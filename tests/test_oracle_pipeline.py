"""CPU tests of the validation-sampler oracle (oracle/pipeline.py, SURVEY.md 8(f)-4;
/root/reference/models/pipeline_flax_stable_diffusion.py:160-254) and of the host-side argument handling of the product
pipeline (no GPU work: the checks under test run before any kernel)."""
import numpy as np
import pytest
import torch

from oracle import clip as OC
from oracle import pipeline as OP
from oracle import sched as OS
from oracle import threefry as OT
from oracle import unet as OU
from oracle import vae as OV

TINY_UNET = OU.UNetConfig(block_out_channels=(32, 64), down_block_types=("CrossAttnDownBlock2D", "DownBlock2D"),
                          up_block_types=("UpBlock2D", "CrossAttnUpBlock2D"), layers_per_block=1, attention_head_dim=2,
                          cross_attention_dim=32)


def _params(seed=3):
    return {"unet": OU.init_params(TINY_UNET, seed), "text_encoder": OC.init_params(OC.TINY_CLIP, seed + 1),
            "vae": OV.init_decoder_params(OV.TINY_VAE, seed + 2)}


def _ids(B, seed):
    r = np.random.default_rng(seed)
    return r.integers(0, OC.TINY_CLIP.vocab_size, size=(B, 77)).astype(np.int32)


def test_ddim_timesteps_follow_set_timesteps():
    # scheduling_ddim_flax.py:177-180: (arange(n) * (T // n)).round()[::-1] + steps_offset
    assert OP.ddim_timesteps(50).tolist() == list(range(980, -1, -20))
    assert OP.ddim_timesteps(3).tolist() == [666, 333, 0]
    assert OP.ddim_timesteps(4, steps_offset=1).tolist() == [751, 501, 251, 1]


def test_ddim_timesteps_match_the_reference_run_golden():
    """tests/golden/ddim_ref.npz was produced by the UNMODIFIED reference scheduler under the jax shim
    (oracle/make_golden.py:ddim_cases: 50 steps, steps_offset 1)."""
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ddim_ref.npz"))
    for one in (0, 1):
        assert np.array_equal(OP.ddim_timesteps(50, 1000, steps_offset=1), np.asarray(g[f"timesteps/{one}"]).astype(np.int64))


def test_generate_shapes_range_and_determinism():
    p = _params()
    img, lat = OP.generate(TINY_UNET, OC.TINY_CLIP, OV.TINY_VAE, p, _ids(2, 0), _ids(2, 1), OT.PRNGKey(5), 2, 32, 48, 7.5,
                           mode="bf16", return_latents=True)
    assert tuple(img.shape) == (2, 32, 48, 3) and lat.shape == (2, 4, 16, 24) and lat.dtype == np.float32
    assert float(img.min()) >= 0.0 and float(img.max()) <= 1.0 and float(img.std()) > 1e-3
    img2 = OP.generate(TINY_UNET, OC.TINY_CLIP, OV.TINY_VAE, p, _ids(2, 0), _ids(2, 1), OT.PRNGKey(5), 2, 32, 48, 7.5,
                       mode="bf16")
    assert torch.equal(img, img2)
    with pytest.raises(ValueError, match="divisible by 8"):
        OP.generate(TINY_UNET, OC.TINY_CLIP, OV.TINY_VAE, p, _ids(1, 0), _ids(1, 1), OT.PRNGKey(5), 2, 30, 48, 7.5)
    with pytest.raises(ValueError, match="Unexpected latents shape"):
        OP.generate(TINY_UNET, OC.TINY_CLIP, OV.TINY_VAE, p, _ids(1, 0), _ids(1, 1), OT.PRNGKey(5), 2, 32, 48, 7.5,
                    latents=np.zeros((1, 4, 8, 8), np.float32))


def test_guidance_one_is_the_text_conditional_sampler_and_zero_the_unconditional_one():
    """uncond + g * (text - uncond): g = 1 must reproduce a sampler that never sees the negative prompt, g = 0 one that
    never sees the prompt (exactly, in fp64: u + (t - u) == t up to one rounding)."""
    p = _params(4)
    ids, neg = _ids(1, 2), _ids(1, 3)
    key = OT.PRNGKey(9)
    n = 3

    def single(which_ids):
        ctx = OC.text_encoder_forward(OC.TINY_CLIP, p["text_encoder"], torch.as_tensor(which_ids).long(), mode="fp64")
        lat = OT.normal(key, (1, 4, 16, 16)).astype(np.float32)
        st = OS.common_state("scaled_linear", 0.00085, 0.012, 1000)
        for t in OP.ddim_timesteps(n):
            pred = OU.unet_forward(TINY_UNET, p["unet"], torch.from_numpy(lat).double(), torch.tensor([int(t)]), ctx,
                                   mode="fp64").numpy().astype(np.float32)
            lat = OS.ddim_step(st, pred, int(t), lat, n)
        return lat

    for g, which in ((1.0, ids), (0.0, neg)):
        _, lat = OP.generate(TINY_UNET, OC.TINY_CLIP, OV.TINY_VAE, p, ids, neg, key, n, 32, 32, g, mode="fp64",
                             return_latents=True)
        want = single(which)
        assert np.abs(lat - want).max() <= 1e-5 * np.abs(want).max()


def test_bf16_emulation_stays_close_to_exact_arithmetic():
    p = _params(5)
    a, la = OP.generate(TINY_UNET, OC.TINY_CLIP, OV.TINY_VAE, p, _ids(2, 4), _ids(2, 5), OT.PRNGKey(1), 3, 32, 32, 5.0,
                        mode="fp64", return_latents=True)
    b, lb = OP.generate(TINY_UNET, OC.TINY_CLIP, OV.TINY_VAE, p, _ids(2, 4), _ids(2, 5), OT.PRNGKey(1), 3, 32, 32, 5.0,
                        mode="bf16", return_latents=True)
    assert np.linalg.norm(la - lb) / np.linalg.norm(la) < 0.1
    assert (a.double() - b.double()).abs().max().item() < 0.25


# ---- product pipeline: argument handling that needs no GPU ------------------------------------------------------
def _product_pipeline():
    from stable_diffusion_training_b200 import clip as PC, pipeline as PP, training_utils as TU, unet as PU, vae as PV
    return PP, PP.FlaxStableDiffusionPipeline(vae=PV.AutoencoderKL(PV.TINY_VAE), text_encoder=PC.CLIPTextModel(PC.TINY_CLIP),
                                              tokenizer=None, unet=TU.UNetModel(PU.TINY), scheduler=None)


def test_product_pipeline_argument_checks_cpu():
    PP, pipe = _product_pipeline()
    assert pipe.vae_scale_factor == 2
    with pytest.raises(ValueError, match="divisible by 8"):
        pipe._generate(np.zeros((1, 77), np.int32), {}, (0, 1), 2, 36, 32, 7.5)
    with pytest.raises(ValueError, match="has to be of type"):
        pipe.prepare_inputs(3)
    assert pipe._scalar_guidance(7.5) == 7.5
    assert pipe._scalar_guidance(np.array([[2.0, 2.0]])) == 2.0
    with pytest.raises(NotImplementedError):
        pipe._scalar_guidance([1.0, 2.0])
