"""8-bit Lion behind the reference's optimizer surface (/root/reference/lion_quant.py:12-211).

Same names, arguments, state layout and error behaviour as the reference:
  ScaleBy8bitLionState(count, mu_quant, mu_quant_flag)                      lion_quant.py:12-17
  scale_by_lion_8bit(b1, b2, mu_scale_dtype, block_size, excluded_layer_mask) -> GradientTransformation   20-156
  lion_8bit(learning_rate, b1, b2, mu_scale_dtype, block_size, weight_decay, mask, excluded_layer_mask)    159-211
plus the few optax pieces the reference chains around it (training_utils.py:355-382): clip_by_global_norm, chain,
apply_updates, constant_schedule. A quantised leaf of `mu_quant` is the tuple (int8 codes [n/bs, bs], fp32 RECIPROCAL
absmax [n/bs, 1]); an unquantised leaf is an fp32 tensor of the parameter's shape.

Pytrees are nested dicts of torch CUDA tensors. All arithmetic runs in libsdb200.so (sdb_lion8_update,
sdb_lion_fp32_update, sdb_global_norm, ...): there is no CPU or eager path.
"""
from __future__ import annotations

from typing import Any, Callable, NamedTuple, Optional, Union

import torch

from . import _lib


# ---------------------------------------------------------------------------------------------------------------
# minimal pytree helpers (nested dicts, sorted keys = jax leaf order)
def tree_map(fn, tree, *rest):
    if isinstance(tree, dict):
        return {k: tree_map(fn, tree[k], *(r[k] for r in rest)) for k in sorted(tree)}
    return fn(tree, *rest)


def tree_leaves(tree):
    if isinstance(tree, dict):
        out = []
        for k in sorted(tree):
            out.extend(tree_leaves(tree[k]))
        return out
    return [tree]


def tree_map_with_path(fn, tree, *rest, _path=()):
    if isinstance(tree, dict):
        return {k: tree_map_with_path(fn, tree[k], *(r[k] for r in rest), _path=_path + (k,)) for k in sorted(tree)}
    return fn(_path, tree, *rest)


def flatten_dict(tree, prefix=""):
    out = {}
    for k in sorted(tree):
        v = tree[k]
        name = f"{prefix}/{k}" if prefix else k
        if isinstance(v, dict):
            out.update(flatten_dict(v, name))
        else:
            out[name] = v
    return out


def unflatten_dict(flat):
    out: dict = {}
    for name, v in flat.items():
        parts = name.split("/")
        d = out
        for p in parts[:-1]:
            d = d.setdefault(p, {})
        d[parts[-1]] = v
    return out


def create_mask(pytree: dict, excluded_layer_list: list):
    """training_utils.py:116-131: True unless one of the path components equals an excluded name."""
    return tree_map_with_path(lambda path, _leaf: not any(e in path for e in excluded_layer_list), pytree)


# ---------------------------------------------------------------------------------------------------------------
class GradientTransformation:
    """optax.GradientTransformation: an (init, update) pair (unpackable like the NamedTuple). `hyper` carries the
    hyper-parameters so that TrainState can run the whole chain as fused launches over its flat arena."""

    def __init__(self, init: Callable, update: Callable, hyper: Optional[dict] = None):
        self.init = init
        self.update = update
        self.hyper = hyper or {"kind": "opaque"}

    def __iter__(self):
        yield self.init
        yield self.update


class EmptyState(NamedTuple):
    pass


class ScaleBy8bitLionState(NamedTuple):
    """State for the Lion algorithm (lion_quant.py:12-17)."""
    count: Any
    mu_quant: Any  # quantized mu
    mu_quant_flag: Any


def _expand_mask(mask, params):
    if mask is None:
        return tree_map(lambda _p: True, params)
    if callable(mask):
        mask = mask(params)
    if isinstance(mask, bool):
        return tree_map(lambda _p: mask, params)
    return mask


def _is_quantized(node):
    return isinstance(node, tuple)


def _gn_ptr(gnorm):
    return 0 if gnorm is None else gnorm.data_ptr()


def _lion_leaf(g, p, m, *, b1, b2, lr, wd, wd_flag, gnorm, max_norm, mode, out, shadow=None, block_size=16):
    l = _lib.lib()
    n = g.numel()
    if _is_quantized(m):
        codes, inv = m
        _lib.check(l.sdb_lion8_update(g.data_ptr(), p.data_ptr(), out.data_ptr(), _lib.ptr(shadow), codes.data_ptr(),
                                      inv.data_ptr(), n, block_size, b1, b2, lr, wd, int(wd_flag), _gn_ptr(gnorm),
                                      max_norm, mode, 0, 0.0, _lib.stream_ptr()), "sdb_lion8_update")
    else:
        _lib.check(l.sdb_lion_fp32_update(g.data_ptr(), p.data_ptr(), out.data_ptr(), _lib.ptr(shadow), m.data_ptr(), n,
                                          b1, b2, lr, wd, int(wd_flag), _gn_ptr(gnorm), max_norm, mode, 0, 0.0,
                                          _lib.stream_ptr()), "sdb_lion_fp32_update")


def compute_model_ema(ema_params, update_params, ema_rate: float):
    """training_utils.py:537-544: ema <- ema_rate * ema + (1 - ema_rate) * params, leaf by leaf, IN PLACE on the EMA
    tree (the reference's jit donates it). train_step itself fuses this into the optimizer launch."""
    l = _lib.lib()

    def leaf(e, p):
        _lib.check(l.sdb_ema_update(e.data_ptr(), p.data_ptr(), e.numel(), float(ema_rate), _lib.stream_ptr()),
                   "sdb_ema_update")
        return e
    return tree_map(leaf, ema_params, update_params)


def scale_by_lion_8bit(b1: float = 0.9, b2: float = 0.99, mu_scale_dtype: Optional[Any] = None,
                       block_size: Optional[int] = 16, excluded_layer_mask: Optional[Any] = None
                       ) -> GradientTransformation:
    """Rescale updates according to the Lion algorithm with block-quantised momentum (lion_quant.py:20-156).
    update() returns sign((1-b1) g + b1 m) and re-quantises m' = (1-b2) g + b2 m. The state is updated IN PLACE
    (the reference's jit donates it, training_utils.py:846-851)."""
    if mu_scale_dtype not in (None, torch.float32, "float32", "fp32"):
        raise NotImplementedError("mu_scale_dtype: only float32 (the reference's canonicalize_dtype(None)) is built")
    if block_size not in (16, 32, 64):
        raise NotImplementedError(f"block_size {block_size}: the kernels are built for 16, 32 and 64")

    def init_fn(params):
        flags = _expand_mask(excluded_layer_mask, params)

        def leaf(t, flag):
            if flag:
                if t.numel() % block_size != 0:
                    raise ValueError(f"cannot reshape array of size {t.numel()} into shape (-1, {block_size})")
                nb = t.numel() // block_size
                # _block_quantize(zeros): codes = quantize(0) = 3, scales = 1/where(0 <= 0, 1, .) = 1  (lion_quant.py:66-80)
                return (torch.full((nb, block_size), 3, dtype=torch.int8, device=t.device),
                        torch.ones((nb, 1), dtype=torch.float32, device=t.device))
            return torch.zeros(t.shape, dtype=torch.float32, device=t.device)

        return ScaleBy8bitLionState(count=torch.zeros([], dtype=torch.int32, device=tree_leaves(params)[0].device),
                                    mu_quant=tree_map(leaf, params, flags), mu_quant_flag=excluded_layer_mask)

    def update_fn(updates, state, params=None):
        del params

        def leaf(g, m):
            out = torch.empty_like(g)
            # mode 1 with lr = -1, no decay: out = sign(c) exactly
            _lion_leaf(g, g, m, b1=b1, b2=b2, lr=-1.0, wd=0.0, wd_flag=False, gnorm=None, max_norm=1.0, mode=1, out=out,
                       block_size=block_size)
            return out

        flat_u, flat_m = flatten_dict(updates), _flatten_state(state.mu_quant)
        new_updates = unflatten_dict({k: leaf(flat_u[k], flat_m[k]) for k in flat_u})
        state.count.add_(1)
        return new_updates, ScaleBy8bitLionState(count=state.count, mu_quant=state.mu_quant,
                                                 mu_quant_flag=state.mu_quant_flag)

    return GradientTransformation(init_fn, update_fn)


def _flatten_state(tree, prefix=""):
    """flatten a mu_quant tree whose leaves are tensors or (codes, scales) tuples"""
    out = {}
    for k in sorted(tree):
        v = tree[k]
        name = f"{prefix}/{k}" if prefix else k
        if isinstance(v, dict):
            out.update(_flatten_state(v, name))
        else:
            out[name] = v
    return out


ScalarOrSchedule = Union[float, Callable[[int], float]]


def constant_schedule(value: float):
    return lambda count: value


class ScaleByScheduleState(NamedTuple):
    count: Any


class MaskedState(NamedTuple):
    inner_state: Any


class LionChainState(NamedTuple):
    """(ScaleBy8bitLionState, MaskedState(add_decayed_weights), ScaleByScheduleState) — optax.chain state tuple."""
    lion: ScaleBy8bitLionState
    decay: MaskedState
    schedule: ScaleByScheduleState


def lion_8bit(learning_rate: ScalarOrSchedule, b1: float = 0.9, b2: float = 0.99, mu_scale_dtype: Optional[Any] = None,
              block_size: int = 64, weight_decay: float = 1e-3, mask: Optional[Any] = None,
              excluded_layer_mask: Optional[Any] = None) -> GradientTransformation:
    """The Lion optimizer with 8-bit state (lion_quant.py:159-211) =
    chain(scale_by_lion_8bit, add_decayed_weights(weight_decay, mask), scale_by_learning_rate(learning_rate)).
    update(updates, state, params) returns (-lr) * (sign(c) + wd * p) in ONE kernel per leaf."""
    inner = scale_by_lion_8bit(b1=b1, b2=b2, mu_scale_dtype=mu_scale_dtype, block_size=block_size,
                               excluded_layer_mask=excluded_layer_mask)

    def init_fn(params):
        s = inner.init(params)
        return LionChainState(s, MaskedState(EmptyState()),
                              ScaleByScheduleState(torch.zeros([], dtype=torch.int32, device=s.count.device)))

    def update_fn(updates, state, params=None, *, _gnorm=None, _max_norm=1.0):
        if params is None:
            raise ValueError("lion_8bit: params are required (add_decayed_weights needs them)")
        lr = learning_rate(int(state.schedule.count)) if callable(learning_rate) else float(learning_rate)
        flat_u, flat_p = flatten_dict(updates), flatten_dict(params)
        flat_m = _flatten_state(state.lion.mu_quant)
        flat_wd = flatten_dict(_expand_mask(mask, params))
        out = {}
        for k, g in flat_u.items():
            o = torch.empty_like(g)
            _lion_leaf(g, flat_p[k], flat_m[k], b1=b1, b2=b2, lr=lr, wd=weight_decay, wd_flag=bool(flat_wd[k]),
                       gnorm=_gnorm, max_norm=_max_norm, mode=1, out=o, block_size=block_size)
            out[k] = o
        state.lion.count.add_(1)
        state.schedule.count.add_(1)
        return unflatten_dict(out), state

    return GradientTransformation(init_fn, update_fn, {
        "kind": "lion8", "learning_rate": learning_rate, "b1": b1, "b2": b2, "block_size": block_size,
        "weight_decay": weight_decay, "mask": mask, "excluded_layer_mask": excluded_layer_mask})


# ---------------------------------------------------------------------------------------------------------------
def global_norm(tree) -> torch.Tensor:
    """optax.global_norm: sqrt(sum over leaves of sum(x^2)) as a 0-d... [1] fp32 device tensor."""
    l = _lib.lib()
    leaves = tree_leaves(tree)
    dev = leaves[0].device
    ws = torch.empty(l.sdb_sumsq_workspace_bytes(), dtype=torch.uint8, device=dev)
    out = torch.zeros(1, dtype=torch.float32, device=dev)
    for i, t in enumerate(leaves):
        _lib.check(l.sdb_global_norm(t.data_ptr(), t.numel(), ws.data_ptr(), ws.numel(), out.data_ptr(),
                                     1 if i else 0, _lib.stream_ptr()), "sdb_global_norm")
    return out


def clip_by_global_norm(max_norm: float) -> GradientTransformation:
    """optax.clip_by_global_norm [upstream]: g if ||g|| < max_norm else (g / ||g||) * max_norm."""
    def init_fn(params):
        del params
        return EmptyState()

    def update_fn(updates, state, params=None):
        del params
        gn = global_norm(updates)
        l = _lib.lib()

        def leaf(g):
            o = torch.empty_like(g)
            _lib.check(l.sdb_clip_by_global_norm_f32(g.data_ptr(), o.data_ptr(), g.numel(), gn.data_ptr(), max_norm,
                                                     _lib.stream_ptr()), "sdb_clip_by_global_norm_f32")
            return o
        return tree_map(leaf, updates), state

    return GradientTransformation(init_fn, update_fn, {"kind": "clip", "max_norm": float(max_norm)})


def chain(*txs: GradientTransformation) -> GradientTransformation:
    """optax.chain."""
    def init_fn(params):
        return tuple(t.init(params) for t in txs)

    def update_fn(updates, state, params=None):
        new_state = []
        for t, s in zip(txs, state):
            updates, s = t.update(updates, s, params)
            new_state.append(s)
        return updates, tuple(new_state)

    return GradientTransformation(init_fn, update_fn, {"kind": "chain", "parts": [t.hyper for t in txs]})


def apply_updates(params, updates):
    """optax.apply_updates: p + u (new tensors)."""
    l = _lib.lib()

    def leaf(p, u):
        o = torch.empty_like(p)
        _lib.check(l.sdb_add_f32(p.data_ptr(), u.data_ptr(), o.data_ptr(), p.numel(), _lib.stream_ptr()), "sdb_add_f32")
        return o
    return tree_map(leaf, params, updates)
