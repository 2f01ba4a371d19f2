"""Times the validation sampler (pipeline.FlaxStableDiffusionPipeline) at SD-1.5 size on one GPU: 512x512, DDIM,
classifier-free guidance 7.5, random-init weights, eager loop body vs the CUDA-graph loop body (jit=True).
Usage: python tools/time_sampler.py [--steps 20] [--batches 1,4,8]"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stable_diffusion_training_b200 import clip as PC, jax_random as JR, lion_quant as LQ, pipeline as PP, \
    schedulers as PS, training_utils as TU, unet as PU, vae as PV  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--batches", default="1,4,8")
    ap.add_argument("--size", type=int, default=512)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    t0 = time.time()
    params = {"unet": LQ.unflatten_dict(PU.init_params_numpy(PU.SD15, 0)),
              "text_encoder": LQ.unflatten_dict(PC.init_params_numpy(PC.SD15_CLIP, 1)),
              "vae": LQ.unflatten_dict(PV.init_decoder_params_numpy(PV.SD_VAE, 2))}
    sched = PS.FlaxDDIMScheduler(beta_start=0.00085, beta_end=0.012, beta_schedule="scaled_linear",
                                 num_train_timesteps=1000, prediction_type="epsilon", device=dev)
    pipe = PP.FlaxStableDiffusionPipeline(vae=PV.AutoencoderKL(PV.SD_VAE), text_encoder=PC.CLIPTextModel(PC.SD15_CLIP),
                                          tokenizer=None, unet=TU.UNetModel(PU.SD15), scheduler=sched, device=dev)
    print(f"# weights ready in {time.time() - t0:.1f} s", flush=True)
    r = np.random.default_rng(0)
    dec = pipe.vae.decoder(params["vae"], dev)
    for B in [int(b) for b in a.batches.split(",")]:   # the two stages outside the loop, alone
        z = torch.randn(B, 4, a.size // 8, a.size // 8, device=dev)
        ms_dec = _time(lambda: dec.decode_image(z, 1.0 / 0.18215))
        text = pipe._executor("text_encoder", params["text_encoder"])
        ids2 = torch.randint(0, 49408, (2 * B, 77), device=dev, dtype=torch.int32)
        ms_txt = _time(lambda: text.forward(ids2, forward_only=True))
        print(json.dumps({"batch": B, "vae_decode_ms": round(ms_dec, 3), "text_encoder_2B_ms": round(ms_txt, 3),
                          "mem_gb": round(torch.cuda.max_memory_allocated() / 2**30, 2)}), flush=True)
    for B in [int(b) for b in a.batches.split(",")]:
        ids = r.integers(0, 49408, size=(B, 77)).astype(np.int32)
        neg = r.integers(0, 49408, size=(B, 77)).astype(np.int32)
        for jit in (False, True):
            for rep in range(3):   # first: builds executors / captures; report the last
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                w0 = time.perf_counter()
                e0.record()
                img = pipe(ids, params, JR.PRNGKey(rep), a.steps, a.size, a.size, 7.5, neg_prompt_ids=neg, jit=jit,
                           output_type="pt")["images"]
                e1.record()
                torch.cuda.synchronize()
                wall = (time.perf_counter() - w0) * 1e3
            ok = bool(torch.isfinite(img).all())
            print(json.dumps({"batch": B, "steps": a.steps, "size": a.size, "loop": "graph" if jit else "eager",
                              "ms_device": round(e0.elapsed_time(e1), 2), "ms_wall": round(wall, 2),
                              "images_per_s": round(B / wall * 1e3, 3), "launches": pipe.launches_last_call,
                              "finite": ok, "mem_gb": round(torch.cuda.max_memory_allocated() / 2**30, 2)}), flush=True)


def _time(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


if __name__ == "__main__":
    main()
