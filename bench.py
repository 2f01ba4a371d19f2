#!/usr/bin/env python
"""bench.py — UNet train samples/sec at 512^2 (64x64x4 latents), SD-1.5 UNet, batch 8 per GPU (BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config sd15|sd21|tiny|lion] [--batch B]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

One "step" = one pass of the hot path (add_noise -> UNet fwd -> MSE -> UNet bwd -> [bucketed NCCL all-reduce] ->
clip + 8-bit Lion + apply) over one synthetic batch. Prints ONE JSON line on rank 0:
  value        whole-job samples/s with the batch already resident in HBM (CUDA events, max over ranks)
  e2e          the same through the public train_step call with HOST (pinned) batches: H2D copy of latents / noise /
               timesteps / context and a D2H read of the loss inside the timed region, every step
  roofline     tensor-pipe roofline of the step (algorithmic matmul FLOPs) + the HBM roofline of the Lion
               optimizer kernels timed alone ("kernels" sub-object)
  cpu_baseline the oracle (CPU restatement of the reference: torch-CPU fp32 UNet + NumPy Lion; JAX is not installable
               here) timed on this box's host cores on a bounded sample of the same workload: 1 of the step's samples
               through the UNet + 1/B of the optimizer's elements per step — nothing is extrapolated
--impl reference times that CPU restatement as the reference arm (the reference itself is JAX/TPU code and cannot run):
               exactly --warmup + --steps bounded steps, same `config` object as our arm.
--config lion  BASELINE.json configs[4]: the 8-bit Lion step alone over the full SD-1.5 parameter set (HBM roofline).
--config sd21  BASELINE.json configs[3]: SD-2.x UNet, 96x96 latents, batch 4 per GPU, v-prediction.
Refuses to run while SDB_* experiment variables are set (--allow-experiments stamps them into the JSON instead).
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "unet_train_samples_per_sec_512px"
UNIT = "samples/s"
# algorithmic matmul-class FLOPs per trained sample (forward x3), SURVEY.md 8(d)
TRAIN_FLOP_PER_SAMPLE = {"sd15": 2.4098e12, "sd21": 6.4473e12, "tiny": 3.3166e9}
WORKLOADS = {
    "sd15": dict(name="SD-1.5 UNet (860M) train_step, 64x64x4 latents, ctx 77x768, eps-pred, Lion-8bit", H=64, L=77, batch=8),
    "sd21": dict(name="SD-2.x UNet train_step, 96x96x4 latents, ctx 77x1024, v-pred, Lion-8bit", H=96, L=77, batch=4),
    "tiny": dict(name="tiny UNet (32/64) train_step, 32x32x4 latents, ctx 77x32", H=32, L=77, batch=2),
    # the reference's COMPLETE step (training_utils.py:504-762): frozen VAE encode of 512x512 pixels, CLIP-L text encoder
    # trained inside the step on the repo's default 3 x 75 + 2 token context (227 tokens), SD-1.5 UNet, two Lion updates
    "sd15_full": dict(name="SD-1.5 complete step: VAE encode 512x512 px + CLIP-L train (227-token context) + UNet train, "
                           "eps-pred, Lion-8bit x2", H=64, L=227, batch=8),
}
TRAIN_FLOP_PER_SAMPLE["sd15_full"] = 2.4479e12  # UNet only (ctx 227); VAE / CLIP FLOPs are not counted


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sustained=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    source="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------------
def host_batches(cfg_name, cfg, B, n, seed, pin):
    import numpy as np
    import torch
    w = WORKLOADS[cfg_name]
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n):
        if cfg_name == "sd15_full":
            b = {"pixel_values": torch.from_numpy(rng.uniform(-1, 1, (B, 3, 8 * w["H"], 8 * w["H"])).astype(np.float32)),
                 "input_ids": torch.from_numpy(rng.integers(0, 49408, (B * 3, 77)).astype(np.int32))}
            out.append({k: v.pin_memory() for k, v in b.items()} if pin else b)
            continue
        b = {"latents": torch.from_numpy(rng.standard_normal((B, 4, w["H"], w["H"]), dtype=np.float32)),
             "noise": torch.from_numpy(rng.standard_normal((B, 4, w["H"], w["H"]), dtype=np.float32)),
             "timesteps": torch.from_numpy(rng.integers(0, 1000, B).astype(np.int32)),
             "encoder_hidden_states": torch.from_numpy(
                 rng.standard_normal((B, w["L"], cfg.cross_attention_dim), dtype=np.float32))}
        if pin:
            b = {k: v.pin_memory() for k, v in b.items()}
        out.append(b)
    return out


def fast_params(cfg, seed=0):
    """Synthetic random-init weights with the distributions of SURVEY.md 8(d) generated on the device (the 860M-element
    NumPy generator of the tests takes ~10 s per rank; the benchmark only needs the same shapes and statistics)."""
    import torch
    from stable_diffusion_training_b200 import unet as PU
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    out = {}
    for name, shape in PU.param_shapes(cfg).items():
        parts = name.split("/")
        leaf, parent = parts[-1], parts[-2]
        r = torch.randn(shape, generator=g, device="cuda", dtype=torch.float32)
        if leaf == "kernel":
            fan_in = 1
            for d in shape[:-1]:
                fan_in *= d
            r.mul_(1.0 / math.sqrt(fan_in))
        elif leaf == "scale":
            r.mul_(0.1).add_(1.0)
        elif parent.startswith("norm") or parent == "conv_norm_out":
            r.mul_(0.1)
        else:
            r.mul_(0.02)
        out[name] = r
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist
    from stable_diffusion_training_b200 import _lib
    from stable_diffusion_training_b200 import lion_quant as LQ
    from stable_diffusion_training_b200 import parallel
    from stable_diffusion_training_b200 import training_utils as TU
    from stable_diffusion_training_b200 import unet as PU

    os.environ.setdefault("NCCL_DEBUG", "WARN")  # keep NCCL's version banner off stdout: rank 0 prints ONE JSON line
    rank, local, world = parallel.init_process_group_from_env("nccl")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    L = _lib.lib()
    _lib.check(L.sdb_check_device(local), "sdb_check_device")
    full = args.config == "sd15_full"
    cfg = {"sd15": PU.SD15, "sd21": PU.SD21, "tiny": PU.TINY, "sd15_full": PU.SD15}[args.config]
    w = WORKLOADS[args.config]
    B = args.batch or w["batch"]
    ptype = "v_prediction" if args.config == "sd21" else "epsilon"

    params = LQ.unflatten_dict(fast_params(cfg, 0))
    model = TU.UNetModel(cfg)
    state = TU.TrainState.create(apply_fn=model.apply, params=params, tx=TU.create_lion_optimizer(params), device=dev)
    del params
    if world > 1:
        parallel.attach(state)
    sched = TU.default_scheduler(dev, "scaled_linear", ptype)
    extra = {}
    if full:
        from stable_diffusion_training_b200 import clip as CL
        from stable_diffusion_training_b200 import jax_random as JR
        from stable_diffusion_training_b200 import vae as PV
        tparams = LQ.unflatten_dict({k: torch.from_numpy(v) for k, v in CL.init_params_numpy(CL.SD15_CLIP, 1).items()})
        tstate = TU.TrainState.create(apply_fn=CL.CLIPTextModel(CL.SD15_CLIP).__call__, params=tparams,
                                      tx=TU.create_lion_optimizer(tparams), device=dev)
        vtree = LQ.unflatten_dict(PV.init_params_numpy(PV.SD_VAE, 2))
        if world > 1:
            parallel.attach(tstate)
        extra = dict(frozen_vae_state=TU.FrozenModel(call=PV.AutoencoderKL(PV.SD_VAE), params=vtree),
                     text_encoder_state=tstate)
        train_rng = JR.RngStream(1234)
    step_fn = TU.compile_train_step(state, sched, **extra) if args.graph else \
        (lambda s, b, r: TU.train_step(s, b, r, frozen_noise_scheduler_state=sched, **extra))
    compiled = step_fn if args.graph else None
    if full:  # the step draws VAE sample noise, diffusion noise and timesteps itself from train_rng
        inner = step_fn
        step_fn = lambda s, b, r: inner(s, b, train_rng)

    nb = 4
    hb = host_batches(args.config, cfg, B, nb, 1 + rank, pin=True)
    db = [{k: v.to(dev) for k, v in b.items()} for b in hb]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident leg ("value") ----
    for i in range(args.warmup):
        state, m = step_fn(state, db[i % nb], i)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    n0 = L.sdb_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        state, m = step_fn(state, db[i % nb], i)
    e1.record()
    barrier()
    launches = L.sdb_launch_count() - n0
    if compiled is not None:  # replays do not pass through the library's counter; what runs eagerly around them does
        launches += compiled.launches_per_replay * args.steps   # (the frozen VAE encode, the Threefry draws)
    ms = e0.elapsed_time(e1)
    loss_last = float(m["loss"])
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())

    # ---- end-to-end leg: host batches, H2D + D2H inside the timed region ----
    h2d = sum(v.numel() * v.element_size() for v in hb[0].values())
    for i in range(2):
        state, m = step_fn(state, {k: v.to(dev, non_blocking=True) for k, v in hb[i % nb].items()}, i)
        float(m["loss"])
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        batch = {k: v.to(dev, non_blocking=True) for k, v in hb[i % nb].items()}
        state, m = step_fn(state, batch, i)
        loss_host = float(m["loss"])  # D2H read of the step's result
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t.item())
    clocks = sampler.stop() if rank == 0 else None

    out = None
    if rank == 0:
        peaks = measured_peaks()
        sps = world * B * args.steps / (ms_total / 1e3)
        flops = TRAIN_FLOP_PER_SAMPLE[args.config] * B  # per rank per step
        step_ms = ms_total / args.steps
        ach_tf = flops / (step_ms / 1e3) / 1e12
        kernels = kernel_rooflines(state, peaks) if not args.no_kernels else {}
        out = {
            "metric": METRIC, "value": round(sps, 3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(step_ms, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": common_config(args.config, B, world),
            "run": {"cuda_graph": bool(args.graph), "loss_last": loss_last},
            "e2e": {"value": round(world * B * args.steps / e2e_s, 3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": 4, "loss_last": loss_host},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": roofline_object(kernels, peaks, ach_tf, step_ms),
        }
        if not args.no_cpu and not full:  # the CPU restatement of the complete step is not timed (UNet arm only)
            out["cpu_baseline"] = cpu_baseline(args.config, steps=2, warmup=0)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out


def ncu_traffic():
    """dram read + write bytes per launch of the dominant kernel from the committed `ncu --set full` capture of THIS
    build (profiles/ncu_traffic.json: {"build": sha1 of csrc/gemm_tc.cu, "traffic": bytes, "source": file}); None when
    there is no capture or it belongs to another build of the kernel."""
    import hashlib
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        src = os.path.join(ROOT, "stable_diffusion_training_b200", "csrc", "gemm_tc.cu")
        if d.get("build") == hashlib.sha1(open(src, "rb").read()).hexdigest():
            return int(d["traffic"]), d.get("source")
    except Exception:
        pass
    return None, None


def roofline_object(kernels, peaks, step_tf, step_ms):
    """`roofline` = the dominant kernel (the tcgen05 implicit-GEMM conv3x3, 320->320 @ 64x64, B=8: the GEMM family is
    ~50 % of the step, profiles/r3e_summary.md) timed live with CUDA events; `traffic` = dram bytes per launch from the
    committed `ncu --set full` capture of this build (ncu_traffic), else null. The whole step and the other kernels
    ride along."""
    k = kernels.get("conv3x3_fwd_320x320@64")
    step = {"bound": "tensor", "achieved": round(step_tf, 2), "peak": peaks["tf_sustained"], "unit": "TFLOP/s",
            "frac": round(step_tf / peaks["tf_sustained"], 4), "ms": round(step_ms, 3),
            "what": "whole train step: algorithmic matmul FLOPs/step (SURVEY 8(d)) / step time, sustained bf16 peak"}
    if k is None:
        return dict(step, traffic=None, of=peaks["source"], kernels=kernels)
    traffic, traffic_src = ncu_traffic()
    return {"bound": "tensor", "achieved": k["achieved"], "peak": k["peak"], "unit": "TFLOP/s", "frac": k["frac"],
            "traffic": traffic, "traffic_source": traffic_src, "kernel": "gemm_tc_kernel<FWD, 2 halves> conv3x3 320->320 @64x64 B=8",
            "flops_per_launch": k["flops"], "ms_per_launch": k["ms"],
            "of": peaks["source"] + ", burst bf16 (kernel timed alone, L2 flushed between launches)",
            "step": step, "kernels": kernels}


def kernel_rooflines(state, peaks):
    """Live CUDA-event timing of the dominant kernels, each alone on the launching stream with the L2 flushed between
    iterations: (1) the 8-bit Lion update over the whole arena (HBM bound, 14.5 B/quantised element + 20 B/fp32-state
    element); (2) the tcgen05 implicit-GEMM conv3x3 320->320 @ 64x64, batch 8 (tensor bound)."""
    import torch
    from stable_diffusion_training_b200 import ops
    res = {}
    a = state.arena
    dev = a.device
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    h = state.hyper

    def timed(fn, iters=10):
        for _ in range(3):
            fn()
        ts = []
        for _ in range(iters):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        return sum(ts) / len(ts)

    a.grad.normal_(0, 3.4e-5)
    lr = h["learning_rate"](0) if callable(h["learning_rate"]) else h["learning_rate"]
    ms = timed(lambda: a.optimizer_step(lr=float(lr), b1=h["b1"], b2=h["b2"], wd=h["weight_decay"], max_norm=None))
    nq = a.n_quant
    nf = a.n_total - a.n_quant
    by = nq * (4 + 8 + 2 + 8 / 16) + nf * 20   # shadow bf16 write excluded from the algorithmic count
    res["lion8_update"] = {"bound": "hbm", "bytes": by, "ms": round(ms, 4), "achieved": round(by / ms / 1e6, 1),
                           "peak": peaks["hbm"], "unit": "GB/s", "frac": round(by / ms / 1e6 / peaks["hbm"], 4)}
    ms = timed(lambda: a.optimizer_step(lr=float(lr), b1=h["b1"], b2=h["b2"], wd=h["weight_decay"], max_norm=1.0))
    by2 = by + a.n_total * 4
    res["clip_norm_plus_lion8"] = {"bound": "hbm", "bytes": by2, "ms": round(ms, 4), "achieved": round(by2 / ms / 1e6, 1),
                                   "peak": peaks["hbm"], "unit": "GB/s", "frac": round(by2 / ms / 1e6 / peaks["hbm"], 4)}
    for (Bn, Hh, cin, cout) in ((8, 64, 320, 320), (8, 32, 640, 640), (8, 16, 1280, 1280)):
        x = torch.randn(Bn, Hh, Hh, cin, device=dev).to(torch.bfloat16)
        wgt = (torch.randn(3, 3, cin, cout, device=dev) * 0.02).to(torch.bfloat16)
        bias = torch.zeros(cout, device=dev)
        out = torch.empty(Bn, Hh, Hh, cout, device=dev, dtype=torch.bfloat16)
        fl = 2.0 * Bn * Hh * Hh * cout * 9 * cin
        ms = timed(lambda: ops.conv3x3(x, wgt, bias=bias, out=out))
        res[f"conv3x3_fwd_{cin}x{cout}@{Hh}"] = {"bound": "tensor", "flops": fl, "ms": round(ms, 4),
                                                "achieved": round(fl / ms / 1e9, 1), "peak": peaks["tf_burst"],
                                                "unit": "TFLOP/s", "frac": round(fl / ms / 1e9 / peaks["tf_burst"], 4)}
    # GroupNorm(32)+SiLU forward on the largest common feature map (HBM bound: 2 B read + 2 B write per element) and
    # the self-attention forward / backward at 64x64 (tensor bound, head dim 40)
    Bn, HW, Cc = 8, 4096, 320
    x = torch.randn(Bn, HW, Cc, device=dev).to(torch.bfloat16)
    gam, bet = torch.ones(Cc, device=dev), torch.zeros(Cc, device=dev)
    y, _st = ops.groupnorm_fwd(x, gam, bet, silu=True)
    ms = timed(lambda: ops.groupnorm_fwd(x, gam, bet, silu=True, out=y))
    by = 4.0 * Bn * HW * Cc
    res["groupnorm_silu_fwd_320@64"] = {"bound": "hbm", "bytes": by, "ms": round(ms, 4), "achieved": round(by / ms / 1e6, 1),
                                        "peak": peaks["hbm"], "unit": "GB/s", "frac": round(by / ms / 1e6 / peaks["hbm"], 4),
                                        "note": "42 MB working set: latency-bound two-pass kernel, not bandwidth-bound"}
    Hh, S, dh = 8, 4096, 40
    q, k, v, dO = ((torch.randn(Bn, S, Hh * dh, device=dev)).to(torch.bfloat16) for _ in range(4))
    o, lse = ops.attention_fwd(q, k, v, Hh)
    fl = 4.0 * Bn * Hh * S * S * dh
    ms = timed(lambda: ops.attention_fwd(q, k, v, Hh, out=o), iters=5)
    res["attention_fwd_S4096_d40"] = {"bound": "tensor", "flops": fl, "ms": round(ms, 4), "achieved": round(fl / ms / 1e9, 1),
                                      "peak": peaks["tf_burst"], "unit": "TFLOP/s",
                                      "frac": round(fl / ms / 1e9 / peaks["tf_burst"], 4)}
    ms = timed(lambda: ops.attention_bwd(q, k, v, o, dO, lse, Hh), iters=5)
    res["attention_bwd_S4096_d40"] = {"bound": "tensor", "flops": 2.5 * fl, "ms": round(ms, 4),
                                      "achieved": round(2.5 * fl / ms / 1e9, 1), "peak": peaks["tf_burst"],
                                      "unit": "TFLOP/s", "frac": round(2.5 * fl / ms / 1e9 / peaks["tf_burst"], 4)}
    return res


# ---------------------------------------------------------------------------------------------------------------
_CPU_CACHE: dict = {}


def _cpu_setup(config: str):
    """Random-init parameters + one synthetic batch for the CPU arm (built once per process)."""
    if config in _CPU_CACHE:
        return _CPU_CACHE[config]
    import numpy as np
    import torch
    from oracle import sched as OS
    from oracle import unet as OU
    cfg = {"sd15": OU.SD15, "sd21": OU.SD21, "tiny": OU.TINY}[config]
    w = WORKLOADS[config]
    g = torch.Generator().manual_seed(0)
    params = {}
    for name, shape in OU.param_shapes(cfg).items():
        fan = 1
        for d in shape[:-1]:
            fan *= d
        params[name] = torch.randn(shape, generator=g) * (1.0 / math.sqrt(max(fan, 1)) if name.endswith("kernel") else 0.02)
        if name.endswith("scale"):
            params[name] += 1.0
    ac = OS.common_state("scaled_linear", 0.00085, 0.012)["alphas_cumprod"]
    B = w["batch"]
    batch = (torch.randn(B, 4, w["H"], w["H"], generator=g), torch.randn(B, 4, w["H"], w["H"], generator=g),
             torch.randint(0, 1000, (B,), generator=g), torch.randn(B, w["L"], cfg.cross_attention_dim, generator=g))
    quant = {n: (n.endswith("kernel") and not any(e in n.split("/") for e in ("conv_in", "conv_out", "time_embedding",
                                                                             "time_emb_proj"))
                 and params[n].numel() % 16 == 0) for n in params}
    _CPU_CACHE[config] = (cfg, params, ac, batch, quant)
    return _CPU_CACHE[config]


def cpu_step(config: str, n_samples: int, cores: int):
    """ONE bounded sample of the workload on the host cores, nothing scaled afterwards: `n_samples` of the step's B
    samples through the oracle's fp32 UNet (forward + loss + backward, torch-CPU, all threads), then the oracle's NumPy
    8-bit Lion over the SAME fraction n_samples / B of every parameter leaf's elements (the optimizer runs once per
    B-sample step, so n_samples samples carry n_samples / B of it; leaf chunks run on a thread pool — NumPy releases
    the GIL). Returns (seconds, unet seconds, lion seconds)."""
    import numpy as np
    import torch
    from concurrent.futures import ThreadPoolExecutor
    from oracle import lion as OL
    from oracle import unet as OU
    cfg, params, ac, batch, quant = _cpu_setup(config)
    w = WORKLOADS[config]
    B = w["batch"]
    ptype = "v_prediction" if config == "sd21" else "epsilon"
    lat, nz, ts, ctx = (t[:n_samples] for t in batch)
    t0 = time.perf_counter()
    loss, grads, _ = OU.loss_and_grads(cfg, params, lat, nz, ts, ctx, ac, prediction_type=ptype, mode="fp32")
    t_unet = time.perf_counter() - t0
    frac = n_samples / B

    def leaf(n):
        numel = params[n].numel()
        m = max(16, int(numel * frac) // 16 * 16) if numel >= 16 else numel
        m = min(m, numel)
        gl = grads[n].numpy().reshape(-1)[:m]
        pl = params[n].numpy().reshape(-1)[:m]
        q = quant[n] and m % 16 == 0
        OL.lion_leaf_update(gl, pl, OL.init_state((m,), q), wd_flag=n.endswith("kernel"))
        return m

    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=cores) as ex:
        done = sum(ex.map(leaf, sorted(params, key=lambda n: -params[n].numel())))
    t_lion = time.perf_counter() - t0
    del grads
    return t_unet + t_lion, t_unet, t_lion, done


def cpu_baseline(config: str, steps: int = 2, warmup: int = 0, n_samples: int = 0):
    """The oracle timed as the CPU baseline (kind = "port": the reference is JAX/Flax for TPU and cannot be imported
    or installed on this box). value = samples/s = n_samples / seconds of a bounded step (see cpu_step)."""
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    w = WORKLOADS[config]
    B = w["batch"]
    ns = n_samples or (B if config == "tiny" else 1)
    for _ in range(warmup):
        cpu_step(config, ns, cores)
    times = [cpu_step(config, ns, cores) for _ in range(max(1, steps))]
    tot = sum(t[0] for t in times)
    t_unet = sum(t[1] for t in times) / len(times)
    t_lion = sum(t[2] for t in times) / len(times)
    return {"value": round(ns * len(times) / tot, 5), "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"per step: {ns} of the step's {B} samples through the oracle's torch-CPU fp32 UNet fwd+loss+bwd "
                      f"({t_unet:.2f} s, {cores} threads) + the oracle's NumPy Lion-8bit over {ns}/{B} of every leaf "
                      f"({times[-1][3]} elements, {t_lion:.2f} s, {cores} threads); nothing extrapolated; the JAX/Flax "
                      f"reference itself is not installable here",
            "seconds_per_step": round(tot / len(times), 3), "steps_timed": len(times), "total_s": round(tot, 2)}


def common_config(config: str, B: int, world: int):
    """The `config` object — identical in both arms (ours / reference) by construction."""
    w = WORKLOADS[config]
    return {"workload": w["name"], "batch_per_gpu": B, "global_batch": B * world, "parallelism": f"dp{world}",
            "weights": "random-init (SURVEY 8(d) distributions)",
            "l2": "working set (1.7 GB bf16 weights + activations) >> 126 MB L2; 4 rotating batches"}


def run_reference(args):
    """Reference arm: the reference's own train step is JAX/Flax on TPU and cannot be imported or installed on this box
    (no jax / flax / optax / diffusers wheels, no network); the arm therefore times the CPU restatement (oracle port) on
    the host cores — exactly `--warmup` + `--steps` bounded steps (cpu_step). Rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return None
    if args.config == "lion":
        return run_lion_reference(args)
    w = WORKLOADS[args.config]
    B = args.batch or w["batch"]
    cb = cpu_baseline(args.config, steps=args.steps, warmup=args.warmup)
    return {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(1e3 * cb["seconds_per_step"], 1),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": common_config(args.config, B, args.gpus),
            "cpu_baseline": cb, "gpu_launches": 0,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


# ---------------------------------------------------------------------------------------------------------------
# BASELINE.json configs[4]: the quantised-Lion optimizer step alone over the full SD-1.5 parameter set
LION_METRIC = "lion8_optimizer_param_updates_per_sec"
LION_WORKLOAD = ("8-bit Lion step alone (clip_by_global_norm + lion_8bit + apply) over the full SD-1.5 UNet leaf set: "
                 "686 leaves, 256 quantised / 831 201 280 elements, 430 fp32-state / 28 319 684; replicated per GPU")


def run_lion(args):
    """One "step" = global-norm + clip + Lion-8bit + apply over the whole arena (3 + 2 launches), gradients resident
    in HBM (4 rotating gradient arenas would not fit the point: the 3.4 GB arena alone is 27x the L2)."""
    import torch
    import torch.distributed as dist
    from stable_diffusion_training_b200 import _lib
    from stable_diffusion_training_b200 import lion_quant as LQ
    from stable_diffusion_training_b200 import parallel
    from stable_diffusion_training_b200 import training_utils as TU
    from stable_diffusion_training_b200 import unet as PU
    os.environ.setdefault("NCCL_DEBUG", "WARN")
    rank, local, world = parallel.init_process_group_from_env("nccl")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    L = _lib.lib()
    _lib.check(L.sdb_check_device(local), "sdb_check_device")
    cfg = PU.SD15
    params = LQ.unflatten_dict(fast_params(cfg, 0))
    state = TU.TrainState.create(apply_fn=TU.UNetModel(cfg).apply, params=params, tx=TU.create_lion_optimizer(params),
                                 device=dev)
    del params
    a = state.arena
    g = torch.Generator(device=dev).manual_seed(2)
    # g ~ N(0, (3.4e-5)^2) (about unit global norm); 10 warm-up steps so that codes are not all 3
    a.grad.normal_(0, 3.4e-5, generator=g)
    for _ in range(max(args.warmup, 10)):
        state.apply_gradients(grads=None)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    n0 = L.sdb_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        state.apply_gradients(grads=None)
    e1.record()
    barrier()
    launches = L.sdb_launch_count() - n0
    t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    # e2e: a host gradient slab is copied in every step (pinned, 64 MB: the optimizer's real producer is the backward
    # pass on the device, so the copy is bounded) and the global norm is read back
    slab = torch.empty(16 << 20, dtype=torch.float32).pin_memory()
    slab.normal_(0, 3.4e-5)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        a.grad[: slab.numel()].copy_(slab, non_blocking=True)
        state.apply_gradients(grads=None)
        gn = float(a.gnorm.item())
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t.item())
    clocks = sampler.stop() if rank == 0 else None
    out = None
    if rank == 0:
        peaks = measured_peaks()
        n_params = 859_520_964
        step_ms = ms_total / args.steps
        nq, nf = a.n_quant, a.n_total - a.n_quant
        by = nq * 14.5 + nf * 20.0 + a.n_total * 4.0   # Lion + the separate global-norm read of g
        ach = by / step_ms / 1e6
        out = {"metric": LION_METRIC, "value": round(world * n_params * args.steps / (ms_total / 1e3), 1),
               "unit": "params/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 10),
               "ms_per_step": round(step_ms, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
               "dtype": "f32 math, int8 state", "data": "synthetic",
               "config": {"workload": LION_WORKLOAD, "parallelism": f"replicas x{world} (the reference replicates the "
                          "optimizer state: PartitionSpec())", "l2": "3.4 GB gradient arena + 3.4 GB masters >> 126 MB L2"},
               "e2e": {"value": round(world * n_params * args.steps / e2e_s, 1), "unit": "params/s",
                       "h2d_bytes_per_step": slab.numel() * 4, "d2h_bytes_per_step": 4, "gnorm_last": gn},
               "gpu_launches": int(launches), "clocks": clocks,
               "roofline": {"bound": "hbm", "achieved": round(ach, 1), "peak": peaks["hbm"], "unit": "GB/s",
                            "frac": round(ach / peaks["hbm"], 4), "traffic": None,
                            "kernel": "sumsq_partial + lion8_kernel x2 + lion32_kernel x2 over the arena",
                            "bytes_per_step": by, "of": peaks["source"]}}
        if not args.no_cpu:
            out["cpu_baseline"] = lion_cpu_baseline()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out


def lion_cpu_baseline(frac: float = 1.0 / 16):
    """oracle/lion.py over 1/16 of every SD-1.5 leaf (thread pool), reported as params/s of that sample."""
    import numpy as np
    from concurrent.futures import ThreadPoolExecutor
    from oracle import lion as OL
    from oracle import unet as OU
    cores = os.cpu_count() or 1
    shapes = OU.param_shapes(OU.SD15)
    rng = np.random.default_rng(2)
    excl = ("conv_in", "conv_out", "time_embedding", "time_emb_proj")
    work = []
    for n, shp in shapes.items():
        numel = int(np.prod(shp))
        m = min(numel, max(16, int(numel * frac) // 16 * 16))
        q = n.endswith("kernel") and not any(e in n.split("/") for e in excl) and m % 16 == 0
        work.append((n, m, q))
    data = {n: ((rng.standard_normal(m) * 3.4e-5).astype(np.float32), (rng.standard_normal(m) * 0.02).astype(np.float32))
            for n, m, _ in work}

    def leaf(item):
        n, m, q = item
        OL.lion_leaf_update(data[n][0], data[n][1], OL.init_state((m,), q), wd_flag=n.endswith("kernel"))
        return m

    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=cores) as ex:
        done = sum(ex.map(leaf, sorted(work, key=lambda it: -it[1])))
    dt = time.perf_counter() - t0
    return {"value": round(done / dt, 1), "unit": "params/s", "cores": cores, "kind": "port",
            "sample": f"oracle/lion.py (NumPy) over 1/16 of every SD-1.5 leaf = {done} elements in {dt:.2f} s, "
                      f"{cores} threads", "seconds": round(dt, 3)}


def run_lion_reference(args):
    cbs = [lion_cpu_baseline() for _ in range(args.warmup + args.steps)][args.warmup:]
    v = sum(c["value"] for c in cbs) / len(cbs)
    cb = dict(cbs[-1], value=round(v, 1))
    return {"impl": "reference", "metric": LION_METRIC, "value": cb["value"], "unit": "params/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(1e3 * sum(c["seconds"] for c in cbs) / len(cbs), 1),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": LION_WORKLOAD}, "cpu_baseline": cb, "gpu_launches": 0,
            "e2e": {"value": cb["value"], "unit": "params/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


EXPERIMENT_ENV = ("SDB_SKIP", "SDB_GEMM_DBG", "SDB_BN", "SDB_MT", "SDB_SPLITK", "SDB_SPLITK_FORCE", "SDB_SPLITK_RED", "SDB_DELIVERY_CLKS",
                  "SDB_SIDE_STREAM", "SDB_GN_LEGACY", "SDB_EXPERIMENTS", "SDB_LIB_PATH", "SDB_BUILD_TAG", "NVCC_EXTRA")


def experiment_env():
    return {k: v for k, v in os.environ.items() if k.startswith("SDB_") or k == "NVCC_EXTRA"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="sd15", choices=list(WORKLOADS) + ["lion"])
    ap.add_argument("--allow-experiments", action="store_true",
                    help="run although SDB_* experiment variables are set (they are stamped into the JSON line)")
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--graph", type=int, default=1, help="replay the step from a CUDA graph (0: eager launches)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-kernels", action="store_true", help="skip the per-kernel roofline leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    env = experiment_env()
    bad = sorted(k for k in env if k in EXPERIMENT_ENV)
    if bad and not args.allow_experiments:
        sys.stderr.write(f"bench.py: experiment variables are set ({', '.join(bad)}): a benchmark number under them is "
                         "meaningless. Unset them or pass --allow-experiments (they are then stamped into the JSON).\n")
        sys.exit(2)
    # stdout carries exactly ONE JSON line: libraries that print banners on fd 1 (NCCL's version line) go to stderr
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        out = run_reference(args)
    else:
        out = run_lion(args) if args.config == "lion" else run_ours(args)
    sys.stdout.flush()
    if out is not None:
        if env:
            out["experiment_env"] = env
        os.write(real_stdout, (json.dumps(out) + "\n").encode())


if __name__ == "__main__":
    main()
