"""ORACLE tooling (test infrastructure only): a NumPy-backed stand-in for the few `jax` / `jax.numpy` / `optax` / `flax` /
`chex` / `diffusers` names that the reference's OWN arithmetic files import, so that the UNMODIFIED files

    /root/reference/lion_quant.py
    /root/reference/schedulers/scheduling_utils_flax.py
    /root/reference/schedulers/scheduling_ddpm_flax.py

can be executed in this container (jax, flax, optax, chex and diffusers are not installable here: no wheels, no
network). oracle/make_golden.py runs them under this shim and writes tests/golden/*.npz.

What the shim restates (published semantics of the un-vendored third-party code, not reference code):
  * jax type promotion with x64 disabled: python scalars are weakly typed, every float result is float32, int / int true
    division gives float32, int arrays stay int;
  * `x ** n` for a python int n is lax.integer_pow = repeated squaring (x**5 = x * ((x*x)*(x*x)));
    `x ** 0.5` is sqrt (XLA's algebraic simplifier rewrites pow(x, 0.5)); any other float power is the correctly
    rounded fp32 power, computed as fl32(pow_f64(x, (double)fl32(y)));
  * jnp.round = round-half-to-even; jnp.linspace = start*(1-t) + stop*t with the endpoint appended (jax's formula);
    jnp.cumprod = sequential fp32 product (XLA's scan order on TPU is not knowable here; stated in DESIGN.md);
  * jax.tree_util.tree_map over nested dicts where the FIRST tree defines the structure (so tuple leaves of later
    trees are passed whole — the reference relies on exactly that, lion_quant.py:100-113);
  * optax.chain / add_decayed_weights / scale_by_learning_rate / clip_by_global_norm / apply_updates /
    safe_int32_increment as published in optax 0.1.7; flax.struct.dataclass as a frozen dataclass with .replace;
    diffusers ConfigMixin / register_to_config as "store the constructor arguments in self.config".
Every individual arithmetic op is an IEEE fp32 NumPy op (no fused multiply-add), in the order the reference file
states it.
"""
from __future__ import annotations

import dataclasses
import inspect
import sys
import types
from typing import Any, NamedTuple

import numpy as np

F32 = np.float32
_FLOAT_ONLY = {np.true_divide, np.sqrt, np.log, np.exp, np.cos, np.sin, np.power}


class Arr(np.ndarray):
    """ndarray with jax's promotion rules (x64 disabled)."""

    def __array_ufunc__(self, ufunc, method, *inputs, out=None, **kw):
        if method != "__call__" or out is not None:
            return NotImplemented
        raw = [np.asarray(x) if isinstance(x, np.ndarray) else x for x in inputs]
        any_f = any((isinstance(x, np.ndarray) and x.dtype.kind == "f") or isinstance(x, (float, np.floating))
                    for x in raw)
        if any_f or ufunc in _FLOAT_ONLY:
            raw = [np.asarray(x, dtype=F32) if not isinstance(x, np.ndarray) else x.astype(F32, copy=False)
                   for x in raw]
        else:  # integer / bool arithmetic: python ints are weak, arrays keep their dtype
            dt = next((x.dtype for x in raw if isinstance(x, np.ndarray)), None)
            raw = [np.asarray(x, dtype=dt) if not isinstance(x, np.ndarray) else x for x in raw]
        if ufunc is np.power:
            raise TypeError("use Arr.__pow__ / jnp.power (shimmed) for powers")
        res = getattr(ufunc, method)(*raw, **kw)
        res = np.asarray(res)
        if res.dtype == np.float64:
            res = res.astype(F32)
        if res.dtype == np.int64:
            res = res.astype(np.int32)
        return res.view(Arr)

    def __pow__(self, y):
        return power(self, y)

    def __getitem__(self, idx):
        if isinstance(idx, Arr):
            idx = np.asarray(idx)
        return np.asarray(np.ndarray.__getitem__(np.asarray(self), idx)).view(Arr)

    def astype(self, dt, **kw):
        return np.asarray(self).astype(_dt(dt)).view(Arr)

    def round(self, decimals=0):
        return np.rint(np.asarray(self)).view(Arr)

    def flatten(self):
        return np.asarray(self).reshape(-1).view(Arr)


def _dt(dt):
    if dt is None:
        return F32
    if dt in (float, "float32", "fp32"):
        return F32
    return np.dtype(dt)


def _a(x, dtype=None):
    if isinstance(x, Arr) and dtype is None:
        return x
    r = np.asarray(x, dtype=_dt(dtype) if dtype is not None else None)
    if r.dtype == np.float64:
        r = r.astype(F32)
    if r.dtype == np.int64:
        r = r.astype(np.int32)
    return r.view(Arr)


def power(x, y):
    x = _a(x)
    if isinstance(y, (int, np.integer)) and not isinstance(y, bool):
        # lax.integer_pow: repeated squaring, most significant work last (x**5 -> x * ((x*x)*(x*x)))
        assert y >= 0
        if y == 0:
            return _a(np.ones_like(np.asarray(x)))
        acc = None
        base = x
        n = int(y)
        while n:
            if n & 1:
                acc = base if acc is None else acc * base
            n >>= 1
            if n:
                base = base * base
        return acc
    yf = F32(y)
    xf = np.asarray(x).astype(F32)
    if yf == F32(0.5):
        return np.sqrt(xf).view(Arr)
    return np.power(xf.astype(np.float64), np.float64(yf)).astype(F32).view(Arr)


# ---------------------------------------------------------------------------------------------------------------
# jax.numpy
def _linspace(start, stop, num, dtype=None):
    start32, stop32 = F32(start), F32(stop)
    div = num - 1
    step = np.arange(div, dtype=F32) / F32(div)
    out = start32 * (F32(1) - step) + stop32 * step
    return np.concatenate([out, np.array([stop32], F32)]).astype(_dt(dtype)).view(Arr)


def _cumprod(a, axis=0):
    a = np.asarray(a)
    assert a.ndim == 1
    out = np.empty_like(a)
    acc = a.dtype.type(1)
    for i, v in enumerate(a):
        acc = a.dtype.type(acc * v)
        out[i] = acc
    return out.view(Arr)


def _where(c, x, y):
    x, y = _a(x), _a(y)
    if x.dtype.kind == "f" or y.dtype.kind == "f":
        x, y = np.asarray(x).astype(F32), np.asarray(y).astype(F32)
    return np.where(np.asarray(c), x, y).view(Arr)


def _clip(x, a_min=None, a_max=None):
    x = np.asarray(_a(x))
    lo = None if a_min is None else x.dtype.type(a_min)
    hi = None if a_max is None else x.dtype.type(a_max)
    return np.clip(x, lo, hi).view(Arr)


def _make_jnp():
    m = types.ModuleType("jax.numpy")
    m.ndarray = Arr
    m.array = lambda x, dtype=None: _a(x, dtype)
    m.asarray = lambda x, dtype=None: _a(x, dtype)
    m.float32, m.int8, m.int32, m.bfloat16 = np.float32, np.int8, np.int32, None
    m.dtype = np.dtype
    m.sign = lambda x: np.sign(_a(x))
    m.abs = lambda x: np.abs(_a(x))
    m.power = power
    m.round = lambda x: np.rint(np.asarray(_a(x))).view(Arr)
    m.max = lambda x, axis=None, keepdims=False: np.max(np.asarray(_a(x)), axis=axis, keepdims=keepdims).view(Arr)
    m.where = _where
    m.ones_like = lambda x, dtype=None: np.ones(np.shape(x), _dt(dtype) if dtype is not None else _a(x).dtype).view(Arr)
    m.zeros_like = lambda x, dtype=None: np.zeros(np.shape(x), _dt(dtype) if dtype is not None else _a(x).dtype).view(Arr)
    m.zeros = lambda shape, dtype=None: np.zeros(shape, _dt(dtype)).view(Arr)
    m.linspace = _linspace
    m.cumprod = lambda a, axis=0: _cumprod(a, axis)
    m.sqrt = lambda x: np.sqrt(_a(x))
    m.log = lambda x: np.log(_a(x))
    m.exp = lambda x: np.exp(_a(x))
    m.concatenate = lambda xs, axis=0: np.concatenate([np.asarray(x) for x in xs], axis=axis).view(Arr)
    m.arange = lambda a, b=None: (np.arange(a, dtype=np.int32) if b is None else np.arange(a, b, dtype=np.int32)).view(Arr)
    m.clip = _clip
    m.split = lambda x, n, axis=0: [p.view(Arr) for p in np.split(np.asarray(x), n, axis=axis)]
    m.broadcast_to = lambda x, shape: np.broadcast_to(np.asarray(x), shape).view(Arr)
    return m


# ---------------------------------------------------------------------------------------------------------------
# jax.tree_util (nested dicts; the first tree defines the structure)
def _tree_map(f, tree, *rest, _path=(), _with_path=False):
    if isinstance(tree, dict):
        return {k: _tree_map(f, tree[k], *(r[k] for r in rest), _path=_path + (k,), _with_path=_with_path)
                for k in tree}
    return f(_path, tree, *rest) if _with_path else f(tree, *rest)


def _tree_leaves(tree):
    if isinstance(tree, dict):
        return [l for k in tree for l in _tree_leaves(tree[k])]
    return [tree]


class _NextNormal:
    """jax.random.normal stand-in: returns the array queued by the caller (threefry is not reproduced; the golden
    vectors carry the noise explicitly)."""
    queue: list = []


def _make_jax(jnp):
    jax = types.ModuleType("jax")
    jax.numpy = jnp
    tu = types.ModuleType("jax.tree_util")
    tu.tree_map = lambda f, t, *r: _tree_map(f, t, *r)
    tu.tree_map_with_path = lambda f, t, *r: _tree_map(f, t, *r, _with_path=True)
    tu.tree_leaves = _tree_leaves
    jax.tree_util = tu
    jax.tree_map = tu.tree_map
    jax.eval_shape = lambda f, x: types.SimpleNamespace(shape=np.shape(x), dtype=np.asarray(x).dtype)
    dt = types.ModuleType("jax.dtypes")
    dt.canonicalize_dtype = lambda d: _dt(d)
    jax.dtypes = dt
    rnd = types.ModuleType("jax.random")
    rnd.KeyArray = Any
    rnd.PRNGKey = lambda seed: np.array([0, seed], np.uint32)
    rnd.split = lambda key, num=2: key
    rnd.normal = lambda key, shape=(), dtype=F32: (_a(_NextNormal.queue.pop(0)) if _NextNormal.queue
                                                    else np.zeros(shape, F32).view(Arr))
    jax.random = rnd
    return jax, tu, dt, rnd


# ---------------------------------------------------------------------------------------------------------------
# optax 0.1.7 pieces the reference chains (lion_quant.py:7-8, 201-211; training_utils.py:355, 379-382) [upstream]
class GradientTransformation(NamedTuple):
    init: Any
    update: Any


class EmptyState(NamedTuple):
    pass


class MaskedState(NamedTuple):
    inner_state: Any


class ScaleByScheduleState(NamedTuple):
    count: Any


class ScaleState(NamedTuple):
    pass


def _safe_inc(count):
    c = np.asarray(count)
    mx = np.iinfo(np.int32).max
    return np.where(c < mx, c + np.int32(1), np.int32(mx)).astype(np.int32).view(Arr)


def _chain(*txs):
    def init(params):
        return tuple(t.init(params) for t in txs)

    def update(updates, state, params=None):
        new = []
        for t, s in zip(txs, state):
            updates, s = t.update(updates, s, params)
            new.append(s)
        return updates, tuple(new)
    return GradientTransformation(init, update)


def _add_decayed_weights(weight_decay=0.0, mask=None):
    def init(params):
        return MaskedState(EmptyState()) if mask is not None else EmptyState()

    def update(updates, state, params=None):
        if params is None:
            raise ValueError("You are using a transformation that requires the current value of parameters")
        m = mask(params) if callable(mask) else mask
        if m is None:
            out = _tree_map(lambda g, p: g + weight_decay * p, updates, params)
        else:  # optax.masked: the inner transformation touches only the leaves whose mask is True
            out = _tree_map(lambda g, p, f: (g + weight_decay * p) if f else g, updates, params, m)
        return out, state
    return GradientTransformation(init, update)


def _scale_by_learning_rate(learning_rate, flip_sign=True):
    m = -1 if flip_sign else 1
    if callable(learning_rate):
        def init(params):
            return ScaleByScheduleState(count=np.zeros([], np.int32).view(Arr))

        def update(updates, state, params=None):
            step = m * learning_rate(state.count)
            out = _tree_map(lambda g: _a(step, _a(g).dtype) * g, updates)
            return out, ScaleByScheduleState(count=_safe_inc(state.count))
        return GradientTransformation(init, update)
    return GradientTransformation(lambda p: ScaleState(),
                                  lambda u, s, p=None: (_tree_map(lambda g: (m * learning_rate) * g, u), s))


def _global_norm(updates):
    tot = F32(0)
    for l in _tree_leaves(updates):
        x = np.asarray(l).astype(F32)
        tot = F32(tot + np.sum(x * x, dtype=F32))
    return np.sqrt(tot).astype(F32).view(Arr)


def _clip_by_global_norm(max_norm):
    def update(updates, state, params=None):
        g_norm = _global_norm(updates)
        trigger = bool(np.asarray(g_norm) < F32(max_norm))
        out = _tree_map(lambda t: t if trigger else (t / g_norm.astype(_a(t).dtype)) * max_norm, updates)
        return out, state
    return GradientTransformation(lambda p: EmptyState(), update)


def _apply_updates(params, updates):
    return _tree_map(lambda p, u: _a(_a(p) + u, _a(p).dtype), params, updates)


# ---------------------------------------------------------------------------------------------------------------
def _struct_dataclass(cls):
    cls = dataclasses.dataclass(frozen=True)(cls)
    cls.replace = lambda self, **kw: dataclasses.replace(self, **kw)
    return cls


def _register_to_config(init):
    sig = inspect.signature(init)

    def wrapped(self, *a, **kw):
        ba = sig.bind(self, *a, **kw)
        ba.apply_defaults()
        cfg = {k: v for k, v in ba.arguments.items() if k != "self"}
        self.config = types.SimpleNamespace(**cfg)
        init(self, *a, **kw)
    return wrapped


def install():
    """Put the stand-in modules into sys.modules (only when the real ones are absent)."""
    jnp = _make_jnp()
    jax, tu, dt, rnd = _make_jax(jnp)
    mods = {"jax": jax, "jax.numpy": jnp, "jax.tree_util": tu, "jax.dtypes": dt, "jax.random": rnd}
    chex = types.ModuleType("chex")
    chex.Array, chex.ArrayDType = Any, Any
    mods["chex"] = chex
    optax = types.ModuleType("optax")
    src = types.ModuleType("optax._src")
    transform, combine = types.ModuleType("optax._src.transform"), types.ModuleType("optax._src.combine")
    numerics, base = types.ModuleType("optax._src.numerics"), types.ModuleType("optax._src.base")
    alias = types.ModuleType("optax._src.alias")
    transform.add_decayed_weights = _add_decayed_weights
    combine.chain = _chain
    numerics.safe_int32_increment = _safe_inc
    base.GradientTransformation, base.Params, base.EmptyState = GradientTransformation, Any, EmptyState
    alias._scale_by_learning_rate = _scale_by_learning_rate
    alias.ScalarOrSchedule = Any
    src.transform, src.combine, src.numerics, src.base, src.alias = transform, combine, numerics, base, alias
    optax._src = src
    optax.chain, optax.clip_by_global_norm, optax.apply_updates = _chain, _clip_by_global_norm, _apply_updates
    optax.constant_schedule = lambda v: (lambda count: v)
    optax.global_norm = _global_norm
    mods.update({"optax": optax, "optax._src": src, "optax._src.transform": transform, "optax._src.combine": combine,
                 "optax._src.numerics": numerics, "optax._src.base": base, "optax._src.alias": alias})
    flax = types.ModuleType("flax")
    struct = types.ModuleType("flax.struct")
    struct.dataclass = _struct_dataclass
    flax.struct = struct
    mods.update({"flax": flax, "flax.struct": struct})
    diffusers = types.ModuleType("diffusers")
    dutils = types.ModuleType("diffusers.utils")
    dutils.BaseOutput = type("BaseOutput", (), {})
    dutils.PushToHubMixin = type("PushToHubMixin", (), {})
    dcfg = types.ModuleType("diffusers.configuration_utils")
    dcfg.ConfigMixin = type("ConfigMixin", (), {})
    dcfg.register_to_config = _register_to_config
    diffusers.utils, diffusers.configuration_utils = dutils, dcfg
    mods.update({"diffusers": diffusers, "diffusers.utils": dutils, "diffusers.configuration_utils": dcfg})
    for k, v in mods.items():
        sys.modules.setdefault(k, v)
    return mods


def queue_normal(arr):
    _NextNormal.queue.append(np.asarray(arr, F32))
