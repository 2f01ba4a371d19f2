"""ORACLE tooling: run the UNMODIFIED reference files under oracle/jax_shim.py and write golden vectors.

    python -m oracle.make_golden            # needs /root/reference (this container only); writes tests/golden/*.npz

  tests/golden/lion_ref.npz    /root/reference/lion_quant.py: lion_8bit(...) init + 6 update steps on a small pytree
                               (quantised + unquantised leaves, weight-decay mask), and scale_by_lion_8bit alone
  tests/golden/sched_ref.npz   /root/reference/schedulers/scheduling_{utils,ddpm}_flax.py: the four beta schedules,
                               add_noise / get_velocity, DDPM step (3 prediction types, t = 0 / 500 / 999), variances,
                               set_timesteps
The fixtures travel to the GPU box; /root/reference does not. tests/test_oracle_golden.py pins oracle/lion.py and
oracle/sched.py against them bit for bit; tests/test_golden_gpu.py pins the CUDA path against them.
"""
from __future__ import annotations

import importlib
import importlib.util
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    m = importlib.util.module_from_spec(spec)
    sys.modules[name] = m
    spec.loader.exec_module(m)
    return m


def lion_cases():
    lq = _load("ref_lion_quant", os.path.join(REF, "lion_quant.py"))
    import optax  # the shim
    rng = np.random.default_rng(7)
    A = lambda x: sys.modules["jax.numpy"].asarray(np.asarray(x, np.float32))
    shapes = {"blk/conv/kernel": (3, 3, 16, 32), "blk/conv/bias": (32,), "blk/dense/kernel": (64, 48),
              "blk/norm/scale": (48,), "time_emb_proj/kernel": (32, 16)}

    def tree(fn):
        out = {}
        for k, s in shapes.items():
            d = out
            parts = k.split("/")
            for p in parts[:-1]:
                d = d.setdefault(p, {})
            d[parts[-1]] = fn(k, s)
        return out

    params = tree(lambda k, s: A(rng.standard_normal(s) * 0.05))
    qmask = tree(lambda k, s: k.endswith("kernel") and "time_emb_proj" not in k)
    wmask = tree(lambda k, s: k.endswith("kernel"))
    lr = 1e-6 / 7
    tx = lq.lion_8bit(optax.constant_schedule(lr), b1=0.9, b2=0.99, block_size=16, weight_decay=1e-2 * 7, mask=wmask,
                      excluded_layer_mask=qmask)
    state = tx.init(params)
    save = {"lr": np.float64(lr), "names": np.array(sorted(shapes))}

    def flat(t, prefix=""):
        o = {}
        for k, v in t.items():
            n = f"{prefix}/{k}" if prefix else k
            if isinstance(v, dict):
                o.update(flat(v, n))
            else:
                o[n] = v
        return o

    def dump_state(tag, st):
        for n, m in flat(st[0].mu_quant).items():
            if isinstance(m, tuple):
                save[f"{tag}/codes/{n}"] = np.asarray(m[0])
                save[f"{tag}/inv/{n}"] = np.asarray(m[1])
            else:
                save[f"{tag}/mu/{n}"] = np.asarray(m)
        save[f"{tag}/count"] = np.asarray(st[0].count)

    dump_state("s0", state)
    for n, p in flat(params).items():
        save[f"p0/{n}"] = np.asarray(p)
    scales = [3e-5, 1.0, 1e-3, 0.0, 50.0, 1e-7]
    for step, sc in enumerate(scales, 1):
        grads = tree(lambda k, s: A(rng.standard_normal(s) * sc))
        if step == 3:  # an exactly-zero block and a block with one huge entry
            g = np.asarray(grads["blk"]["dense"]["kernel"]).copy()
            g[0, :16] = 0
            g[1, 5] = 1e4
            grads["blk"]["dense"]["kernel"] = A(g)
        updates, state = tx.update(grads, state, params)
        params = optax.apply_updates(params, updates)
        for n, g in flat(grads).items():
            save[f"g{step}/{n}"] = np.asarray(g)
        for n, u in flat(updates).items():
            save[f"u{step}/{n}"] = np.asarray(u)
        for n, p in flat(params).items():
            save[f"p{step}/{n}"] = np.asarray(p)
        dump_state(f"s{step}", state)
    save["nsteps"] = np.int32(len(scales))

    # scale_by_lion_8bit alone (update = sign(c)), block 16, everything quantised
    tx2 = lq.scale_by_lion_8bit(b1=0.9, b2=0.99, block_size=16, excluded_layer_mask={"w": True})
    st2 = tx2.init({"w": A(np.zeros(256))})
    g = A(rng.standard_normal(256) * np.repeat(10.0 ** rng.integers(-8, 3, 16), 16))
    u, st2 = tx2.update({"w": g}, st2)
    u_b, st2b = tx2.update({"w": A(-0.5 * np.asarray(g))}, st2)
    save.update({"sb/g": np.asarray(g), "sb/u1": np.asarray(u["w"]), "sb/codes1": np.asarray(st2.mu_quant["w"][0]),
                 "sb/inv1": np.asarray(st2.mu_quant["w"][1]), "sb/u2": np.asarray(u_b["w"]),
                 "sb/codes2": np.asarray(st2b.mu_quant["w"][0]), "sb/inv2": np.asarray(st2b.mu_quant["w"][1])})

    # optax.clip_by_global_norm(1) in front (training_utils.py:379-382), norm above and below the threshold
    clip = optax.clip_by_global_norm(1)
    for tag, sc in (("clip_hi", 0.5), ("clip_lo", 1e-3)):
        gr = {"a": A(rng.standard_normal(300) * sc), "b": A(rng.standard_normal((7, 9)) * sc)}
        out, _ = clip.update(gr, clip.init(gr))
        save.update({f"{tag}/a": np.asarray(gr["a"]), f"{tag}/b": np.asarray(gr["b"]),
                     f"{tag}/gn": np.asarray(optax.global_norm(gr)), f"{tag}/oa": np.asarray(out["a"]),
                     f"{tag}/ob": np.asarray(out["b"])})
    return save


def sched_cases():
    from oracle import jax_shim
    sys.path.insert(0, REF)
    try:
        ddpm = importlib.import_module("schedulers.scheduling_ddpm_flax")
    finally:
        sys.path.remove(REF)
    jnp = sys.modules["jax.numpy"]
    rng = np.random.default_rng(11)
    save = {}
    for sched_name in ("linear", "scaled_linear", "zero_snr_scaled_linear", "squaredcos_cap_v2"):
        s = ddpm.FlaxDDPMScheduler(beta_start=0.00085, beta_end=0.012, beta_schedule=sched_name,
                                   num_train_timesteps=1000, prediction_type="epsilon")
        st = s.create_state()
        save[f"{sched_name}/betas"] = np.asarray(st.common.betas)
        save[f"{sched_name}/alphas"] = np.asarray(st.common.alphas)
        save[f"{sched_name}/alphas_cumprod"] = np.asarray(st.common.alphas_cumprod)
        save[f"{sched_name}/timesteps"] = np.asarray(st.timesteps)
        save[f"{sched_name}/init_noise_sigma"] = np.asarray(st.init_noise_sigma)
    x0 = rng.standard_normal((6, 4, 8, 8)).astype(np.float32)
    nz = rng.standard_normal((6, 4, 8, 8)).astype(np.float32)
    t = np.array([0, 1, 250, 500, 998, 999], np.int32)
    save.update({"x0": x0, "noise": nz, "t": t})
    for sched_name in ("scaled_linear", "zero_snr_scaled_linear"):
        s = ddpm.FlaxDDPMScheduler(beta_start=0.00085, beta_end=0.012, beta_schedule=sched_name,
                                   num_train_timesteps=1000)
        st = s.create_state()
        save[f"{sched_name}/noisy"] = np.asarray(s.add_noise(st, jnp.asarray(x0), jnp.asarray(nz), jnp.asarray(t)))
        save[f"{sched_name}/velocity"] = np.asarray(s.get_velocity(st, jnp.asarray(x0), jnp.asarray(nz), jnp.asarray(t)))
    # reverse step
    mo = rng.standard_normal((2, 4, 8, 8)).astype(np.float32)
    smp = rng.standard_normal((2, 4, 8, 8)).astype(np.float32)
    vn = rng.standard_normal((2, 4, 8, 8)).astype(np.float32)
    save.update({"step/model_output": mo, "step/sample": smp, "step/noise": vn})
    for ptype in ("epsilon", "sample", "v_prediction"):
        for vt in ("fixed_small", "fixed_large"):
            s = ddpm.FlaxDDPMScheduler(beta_start=0.00085, beta_end=0.012, beta_schedule="scaled_linear",
                                       num_train_timesteps=1000, prediction_type=ptype, variance_type=vt,
                                       clip_sample=(ptype != "sample"))
            st = s.create_state()
            for ti in (0, 500, 999):
                jax_shim._NextNormal.queue.clear()
                jax_shim.queue_normal(vn)
                out = s.step(st, jnp.asarray(mo), ti, jnp.asarray(smp))
                save[f"step/{ptype}/{vt}/{ti}"] = np.asarray(out.prev_sample)
                save[f"var/{vt}/{ti}"] = np.asarray(s._get_variance(st, ti))
    s = ddpm.FlaxDDPMScheduler(num_train_timesteps=1000)
    st = s.set_timesteps(s.create_state(), 50)
    save["set_timesteps_50"] = np.asarray(st.timesteps)
    save["len"] = np.int32(len(s))
    return save


def ddim_cases():
    """The UNMODIFIED /root/reference/schedulers/scheduling_ddim_flax.py under the shim: the placeholder scheduler the
    reference saves (training_utils.py:997-1004: scaled_linear 0.00085-0.012, set_alpha_to_one False, steps_offset 1,
    v_prediction) and the other prediction types, 50 inference steps."""
    sys.path.insert(0, REF)
    try:
        ddim = importlib.import_module("schedulers.scheduling_ddim_flax")
    finally:
        sys.path.remove(REF)
    jnp = sys.modules["jax.numpy"]
    rng = np.random.default_rng(13)
    mo = rng.standard_normal((2, 4, 8, 8)).astype(np.float32)
    smp = rng.standard_normal((2, 4, 8, 8)).astype(np.float32)
    save = {"model_output": mo, "sample": smp}
    for ptype in ("epsilon", "sample", "v_prediction"):
        for one in (True, False):
            s = ddim.FlaxDDIMScheduler(beta_start=0.00085, beta_end=0.012, beta_schedule="scaled_linear",
                                       num_train_timesteps=1000, set_alpha_to_one=one, steps_offset=1,
                                       prediction_type=ptype)
            st = s.set_timesteps(s.create_state(), 50)
            save[f"timesteps/{int(one)}"] = np.asarray(st.timesteps)
            save[f"final/{int(one)}"] = np.asarray(st.final_alpha_cumprod)
            for ti in (981, 501, 21, 1):
                for eta in (0.0, 0.5):
                    out = s.step(st, jnp.asarray(mo), ti, jnp.asarray(smp), eta=eta)
                    save[f"step/{ptype}/{int(one)}/{ti}/{eta}"] = np.asarray(out.prev_sample)
                save[f"var/{int(one)}/{ti}"] = np.asarray(s._get_variance(st, ti, ti - 20))
    return save


def main():
    if not os.path.isdir(REF):
        raise SystemExit("make_golden needs /root/reference (build container only)")
    from oracle import jax_shim
    jax_shim.install()
    os.makedirs(OUT, exist_ok=True)
    lion = lion_cases()
    np.savez_compressed(os.path.join(OUT, "lion_ref.npz"), **lion)
    sched = sched_cases()
    np.savez_compressed(os.path.join(OUT, "sched_ref.npz"), **sched)
    ddim = ddim_cases()
    np.savez_compressed(os.path.join(OUT, "ddim_ref.npz"), **ddim)
    print(f"wrote {len(lion)} lion arrays, {len(sched)} scheduler arrays, {len(ddim)} DDIM arrays to {OUT}")


if __name__ == "__main__":
    sys.path.insert(0, ROOT)
    main()
