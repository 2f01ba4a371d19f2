#!/usr/bin/env python
"""Data-parallel correctness on hardware (/root/reference/training_utils.py:835-840,
852-931: batch sharded on dim 0, everything else replicated, loss / gradients = global mean).

    torchrun --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/dp_check.py \
        [--config sd15|tiny] [--batch B] [--steps K] [--out profiles/rX_dp_check.json]

Checks, on N ranks (N = WORLD_SIZE):
 (a) N ranks x B  ==  1 rank on the same global batch, same weights. Strict form: the single rank runs the N shards as
     micro-batches (the kernels, tile shapes and split-K choices the ranks use) and sums the gradients in fp32 — loss,
     step-1 gradient arena, Lion codes / masters after K steps agree to fp32 summation order (a momentum sitting on a
     rounding threshold may flip a code). Loose form: the single rank runs the CONCATENATED batch (what the reference's
     unsharded program would do) — other tile shapes give other bf16 roundings, so agreement is bounded by the model's
     bf16 noise (reported beside the bf16-emulation yardstick), and every master stays within K Lion steps;
 (b) masters, Lion codes, reciprocal-absmax scales, fp32 momentum and the gradient arena are BIT-IDENTICAL across ranks
     after every step, in eager mode and in segmented-CUDA-graph mode (CompiledTrainStep);
 (c) in-step sampling: with noise / timesteps drawn from train_rng, rank r draws rows [r*B, (r+1)*B) of the global
     draw (the N-rank losses equal the 1-rank loss on the concatenated batch);
 (d) (--config tiny) the JOINT step — text encoder trained inside the step (training_utils.py:635-674, 733), both
     TrainStates attached: losses / masters of both models against one rank on the concatenated batch, both arenas
     bit-identical across ranks after every step, eager and CUDA graph.
Rank 0 prints one JSON object and (with --out) writes it.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np
import torch
import torch.distributed as dist

from stable_diffusion_training_b200 import jax_random as JR
from stable_diffusion_training_b200 import lion_quant as LQ
from stable_diffusion_training_b200 import parallel
from stable_diffusion_training_b200 import training_utils as TU
from stable_diffusion_training_b200 import unet as PU


def checksum(t: torch.Tensor) -> int:
    v = t.contiguous().view(torch.uint8) if t.dtype == torch.int8 else t.contiguous().view(torch.int32)
    # position-dependent 64-bit checksum (a plain sum would miss swapped elements)
    idx = torch.arange(v.numel(), device=v.device, dtype=torch.int64)
    return int(((v.to(torch.int64) * ((idx % 1000003) + 1)).sum()).item())


def ranks_identical(state, world) -> dict:
    a = state.arena
    out = {}
    for name, t in (("master", a.master), ("codes", a.codes), ("inv_scales", a.inv_scales), ("mu", a.mu),
                    ("grad", a.grad), ("shadow", a.shadow.view(torch.int16).to(torch.int32))):
        c = torch.tensor([checksum(t)], dtype=torch.int64, device=a.device)
        lo, hi = c.clone(), c.clone()
        if world > 1:
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        out[name] = bool(lo.item() == hi.item())
    return out


def make_state(cfg, params_np, dev):
    tree = LQ.unflatten_dict({k: torch.from_numpy(v) for k, v in params_np.items()})
    return TU.TrainState.create(apply_fn=TU.UNetModel(cfg).apply, params=tree, tx=TU.create_lion_optimizer(tree),
                                device=dev)


def joint_check(ucfg, uparams_np, sched, dev, rank, world, B, K, chunks=2):
    from stable_diffusion_training_b200 import clip as CL
    tcfg = CL.TINY_CLIP
    tparams_np = CL.init_params_numpy(tcfg, 1)
    G = B * world
    H = 16
    rng = np.random.default_rng(78)
    gb = [{"latents": rng.standard_normal((G, 4, H, H), dtype=np.float32),
           "noise": rng.standard_normal((G, 4, H, H), dtype=np.float32),
           "timesteps": rng.integers(0, 1000, G).astype(np.int32),
           "input_ids": rng.integers(0, tcfg.vocab_size, (G * chunks, 77)).astype(np.int32)} for _ in range(K)]

    def states():
        ust = make_state(ucfg, uparams_np, dev)
        ttree = LQ.unflatten_dict({k: torch.from_numpy(v) for k, v in tparams_np.items()})
        tst = TU.TrainState.create(apply_fn=CL.CLIPTextModel(tcfg).__call__, params=ttree,
                                   tx=TU.create_lion_optimizer(ttree), device=dev)
        return ust, tst

    def part(b, lo, hi):
        out = {k: torch.from_numpy(v[lo:hi]).to(dev) for k, v in b.items() if k != "input_ids"}
        out["input_ids"] = torch.from_numpy(b["input_ids"][lo * chunks:hi * chunks]).to(dev)
        return out

    ust, tst = states()
    ref_loss = []
    for i in range(K):
        ust, m = TU.train_step(ust, part(gb[i], 0, G), None, sched, text_encoder_state=tst)
        ref_loss.append(float(m["loss"]))
    ref_u, ref_t = ust.arena.master.clone(), tst.arena.master.clone()
    del ust, tst
    out = {"loss_1rank_concatenated": ref_loss, "modes": {}}
    lr_bound = 2 * K * (1e-6 / 7) * 1.1
    for mode in ("eager", "graph"):
        ust, tst = states()
        if world > 1:
            parallel.attach(ust, sync="verify")
            parallel.attach(tst, sync="verify")
        step = TU.compile_train_step(ust, sched, text_encoder_state=tst) if mode == "graph" else \
            (lambda s, b, r: TU.train_step(s, b, r, sched, text_encoder_state=tst))
        losses, ident = [], []
        for i in range(K):
            ust, m = step(ust, part(gb[i], rank * B, (rank + 1) * B), None)
            torch.cuda.synchronize()
            losses.append(float(m["loss"]))
            iu, it = ranks_identical(ust, world), ranks_identical(tst, world)
            ident.append({"unet": iu, "text_encoder": it})
        du = float((ust.arena.master - ref_u).abs().max())
        dt = float((tst.arena.master - ref_t).abs().max())
        ok = all(all(d["unet"].values()) and all(d["text_encoder"].values()) for d in ident)
        ok = ok and all(abs(a - b) <= 2e-3 * abs(b) for a, b in zip(losses, ref_loss))
        ok = ok and du <= lr_bound and dt <= lr_bound and int(tst.step) == K
        out["modes"][mode] = {"loss": losses, "ranks_bit_identical_per_step": ident,
                              "unet_masters_max_abs_diff_vs_1rank_concatenated": du,
                              "text_masters_max_abs_diff_vs_1rank_concatenated": dt,
                              "text_bytes_all_reduced": int(tst.dp.bytes_reduced) if tst.dp is not None else 0,
                              "ok": bool(ok)}
        del ust, tst, step
        torch.cuda.empty_cache()
    out["ok"] = bool(all(v["ok"] for v in out["modes"].values()))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="sd15", choices=["sd15", "tiny"])
    ap.add_argument("--batch", type=int, default=2)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    rank, local, world = parallel.init_process_group_from_env("nccl")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    cfg = PU.SD15 if args.config == "sd15" else PU.TINY
    H = 64 if args.config == "sd15" else 32
    B, K = args.batch, args.steps
    params_np = PU.init_params_numpy(cfg, 0)
    sched = TU.default_scheduler(dev, "scaled_linear", "epsilon")
    rng = np.random.default_rng(77)
    G = B * world
    gbatches = [{"latents": rng.standard_normal((G, 4, H, H), dtype=np.float32),
                 "noise": rng.standard_normal((G, 4, H, H), dtype=np.float32),
                 "timesteps": rng.integers(0, 1000, G).astype(np.int32),
                 "encoder_hidden_states": rng.standard_normal((G, 77, cfg.cross_attention_dim), dtype=np.float32)}
                for _ in range(K)]

    def shard(b, with_noise=True):
        keys = b.keys() if with_noise else ("latents", "encoder_hidden_states")
        return {k: torch.from_numpy(b[k][rank * B:(rank + 1) * B]).to(dev) for k in keys}

    def whole(b, with_noise=True):
        keys = b.keys() if with_noise else ("latents", "encoder_hidden_states")
        return {k: torch.from_numpy(b[k]).to(dev) for k in keys}

    report = {"config": args.config, "world": world, "batch_per_rank": B, "steps": K, "modes": {}}
    from stable_diffusion_training_b200 import ops

    def micro_batched_step(st, gb):
        """ONE rank computing the N shards one after the other with the kernels (tile shapes, split-K choices) the N
        ranks use, gradients summed in fp32: the all-reduce result up to fp32 summation order."""
        total, loss_sum = None, 0.0
        for r in range(world):
            sb = {k: torch.from_numpy(v[r * B:(r + 1) * B]).to(dev) for k, v in gb.items()}
            noisy = sched.call.add_noise(sched.params, sb["latents"], sb["noise"], sb["timesteps"])
            x = ops.nchw_to_nhwc_bf16(noisy)
            ctx = ops.cast_bf16(sb["encoder_hidden_states"].contiguous())
            pred = st.net.forward(x, sb["timesteps"].to(torch.int32), ctx)
            loss, dpred = TU.mse_loss_and_grad(pred.v, sb["noise"], None, 1.0 / world)
            st.net.backward(pred, dpred, None)
            total = st.arena.grad.clone() if total is None else total + st.arena.grad
            loss_sum += float(loss) / world
        st.arena.grad.copy_(total)
        st.apply_gradients(grads=None)
        return loss_sum

    # ---- reference 1: one rank, the N shards as micro-batches (strict comparison) ----
    ref = make_state(cfg, params_np, dev)
    ref_loss, ref_grad1 = [], None
    for i in range(K):
        ref_loss.append(micro_batched_step(ref, gbatches[i]))
        if i == 0:
            ref_grad1 = ref.arena.grad.clone()
    ref_master, ref_codes = ref.arena.master.clone(), ref.arena.codes.clone()
    del ref
    # ---- reference 2: one rank on the CONCATENATED batch (other tile shapes -> other fp32 summation orders -> other
    # bf16 roundings: agreement is bounded by the bf16 noise of the model, not by fp32 rounding) ----
    ref = make_state(cfg, params_np, dev)
    cat_loss, cat_grad1 = [], None
    for i in range(K):
        ref, m = TU.train_step(ref, whole(gbatches[i]), None, sched)
        cat_loss.append(float(m["loss"]))
        if i == 0:
            cat_grad1 = ref.arena.grad.clone()
    cat_master = ref.arena.master.clone()
    ref2 = make_state(cfg, params_np, dev)
    ref2, m = TU.train_step(ref2, whole(gbatches[0], False), JR.RngStream(5), sched, offset_noise_magnitude=0.1)
    ref_loss_sampled = float(m["loss"])
    del ref, ref2
    torch.cuda.empty_cache()

    for mode in ("eager", "graph"):
        st = make_state(cfg, params_np, dev)
        if world > 1:
            parallel.attach(st, sync="verify")
        step = TU.compile_train_step(st, sched) if mode == "graph" else \
            (lambda s, b, r: TU.train_step(s, b, r, sched))
        losses, ident, grad_rel, grad_rel_cat = [], [], None, None
        reps = K + (1 if mode == "graph" else 0)  # graph mode: call 1 is the eager warm-up, captured from call 2 on
        for i in range(K):
            st, m = step(st, shard(gbatches[i]), None)
            torch.cuda.synchronize()
            losses.append(float(m["loss"]))
            ident.append(ranks_identical(st, world))
            if i == 0:
                grad_rel = float((st.arena.grad - ref_grad1).double().norm() / ref_grad1.double().norm())
                grad_rel_cat = float((st.arena.grad - cat_grad1).double().norm() / cat_grad1.double().norm())
        frac_master = float((st.arena.master != ref_master).float().mean())
        frac_codes = float((st.arena.codes != ref_codes).float().mean())
        max_master = float((st.arena.master - ref_master).abs().max())
        report["modes"][mode] = {
            "loss": losses, "loss_1rank": ref_loss, "loss_1rank_concatenated": cat_loss,
            "grad_step1_rel_diff_vs_1rank": grad_rel, "grad_step1_rel_diff_vs_1rank_concatenated": grad_rel_cat,
            "masters_max_abs_diff_vs_1rank_concatenated": float((st.arena.master - cat_master).abs().max()),
            "masters_differing_frac_vs_1rank": frac_master, "codes_differing_frac_vs_1rank": frac_codes,
            "masters_max_abs_diff_vs_1rank": max_master,
            "ranks_bit_identical_per_step": ident,
            "bytes_all_reduced": int(st.dp.bytes_reduced) if st.dp is not None else 0,
            "exchange": ("nvls" if st.dp.peer.nvls else "peer") if (st.dp is not None and st.dp.peer is not None) else
                        ("nccl" if st.dp is not None else "none"),
        }
        ok = all(all(d.values()) for d in ident)
        # strict: against the micro-batched single rank (same kernels): fp32 summation order only
        ok = ok and all(abs(a - b) <= 1e-5 * abs(b) for a, b in zip(losses, ref_loss))
        ok = ok and grad_rel < 1e-5 and frac_master < 1e-3 and frac_codes < 1e-3
        ok = ok and max_master <= 2 * K * (1e-6 / 7) * 1.1
        # loose: against the concatenated batch — within half the model's own bf16 noise (gradient yardstick 1.2e-2 of
        # tests/test_train_step_full_gpu.py) and every master within K Lion steps
        ok = ok and all(abs(a - b) <= 2e-3 * abs(b) for a, b in zip(losses, cat_loss)) and grad_rel_cat < 6e-3
        ok = ok and float((st.arena.master - cat_master).abs().max()) <= 2 * K * (1e-6 / 7) * 1.1
        report["modes"][mode]["ok"] = bool(ok)
        if mode == "eager":  # (c) sharded sampling == global draw
            st2 = make_state(cfg, params_np, dev)
            if world > 1:
                parallel.attach(st2, sync="broadcast")
            st2, m = TU.train_step(st2, shard(gbatches[0], False), JR.RngStream(5), sched, offset_noise_magnitude=0.1)
            ls = float(m["loss"])
            report["sampling"] = {"loss": ls, "loss_1rank": ref_loss_sampled,
                                  "ok": bool(abs(ls - ref_loss_sampled) <= 2e-3 * abs(ref_loss_sampled))}
            del st2
        del st, step
        torch.cuda.empty_cache()
    if args.config == "tiny":
        report["joint"] = joint_check(cfg, params_np, sched, dev, rank, world, B, K)
    report["ok"] = bool(all(v["ok"] for v in report["modes"].values()) and report["sampling"]["ok"]
                        and report.get("joint", {"ok": True})["ok"])
    if rank == 0:
        print(json.dumps(report))
        if args.out:
            with open(args.out, "w") as f:
                json.dump(report, f, indent=1)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    sys.exit(0 if report["ok"] else 1)


if __name__ == "__main__":
    main()
