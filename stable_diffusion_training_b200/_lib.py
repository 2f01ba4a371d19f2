"""ctypes binding of libsdb200.so (include/sdb200.h). There is no CPU fallback: if the library cannot be loaded the
import of any operator fails loudly, and every call checks its return code and raises with the library's message."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libsdb200.so")

SDB_MAX_TAPS = 9
SDB_MAX_SRC = 4
SDB_GEMM_FWD = 0
SDB_GEMM_WGRAD = 1


class SdbError(RuntimeError):
    pass


class View4(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("C", C.c_int), ("W", C.c_int), ("H", C.c_int), ("N", C.c_int),
                ("strideW", C.c_longlong), ("strideH", C.c_longlong), ("strideN", C.c_longlong)]


class GemmDesc(C.Structure):
    _fields_ = [
        ("mode", C.c_int), ("num_src", C.c_int), ("a", View4 * SDB_MAX_SRC), ("b", View4), ("b_mn_major", C.c_int),
        ("ntaps", C.c_int), ("tap_src", C.c_int * SDB_MAX_TAPS), ("tap_dh", C.c_int * SDB_MAX_TAPS),
        ("tap_dw", C.c_int * SDB_MAX_TAPS), ("tap_brow", C.c_int * SDB_MAX_TAPS),
        ("outW", C.c_int), ("outH", C.c_int), ("outN", C.c_int), ("bw", C.c_int), ("bh", C.c_int), ("bn", C.c_int),
        ("M", C.c_int), ("N", C.c_int), ("K", C.c_int), ("block_n", C.c_int), ("splits", C.c_int),
        ("d", C.c_void_p), ("ldd", C.c_longlong), ("d_strideW", C.c_longlong), ("d_strideH", C.c_longlong),
        ("d_strideN", C.c_longlong), ("bias", C.c_void_p), ("temb", C.c_void_p), ("ldt", C.c_longlong),
        ("residual", C.c_void_p), ("ldr", C.c_longlong), ("m_halves", C.c_int), ("splitk_slab", C.c_longlong),
    ]


_P = C.c_void_p
_I = C.c_int
_LL = C.c_longlong
_D = C.c_double
_F = C.c_float
_SZ = C.c_size_t
_U32 = C.c_uint32

# name -> (restype, argtypes). Must list every symbol include/sdb200.h declares (tests/test_abi.py checks it).
SIGNATURES = {
    "sdb_last_error_string": (C.c_char_p, []),
    "sdb_abi_version": (_I, []),
    "sdb_check_device": (_I, [_I]),
    "sdb_launch_count": (_LL, []),
    "sdb_lion8_update": (_I, [_P, _P, _P, _P, _P, _P, _LL, _I, _D, _D, _D, _D, _I, _P, _D, _I, _P, _D, _P]),
    "sdb_lion_fp32_update": (_I, [_P, _P, _P, _P, _P, _LL, _D, _D, _D, _D, _I, _P, _D, _I, _P, _D, _P]),
    "sdb_ema_update": (_I, [_P, _P, _LL, _D, _P]),
    "sdb_sumsq_workspace_bytes": (_SZ, []),
    "sdb_global_norm": (_I, [_P, _LL, _P, _SZ, _P, _I, _P]),
    "sdb_clip_by_global_norm_f32": (_I, [_P, _P, _LL, _P, _D, _P]),
    "sdb_add_f32": (_I, [_P, _P, _P, _LL, _P]),
    "sdb_sum_slices_f32": (_I, [_P, _P, _I, _I, _LL, _P]),
    "sdb_nvls_allreduce_f32": (_I, [_P, _LL, _I, _P]),
    "sdb_add_noise": (_I, [_P, _P, _P, _P, _I, _P, _I, _LL, _P]),
    "sdb_get_velocity": (_I, [_P, _P, _P, _P, _I, _P, _I, _LL, _P]),
    "sdb_ddpm_step": (_I, [_P, _P, _P, _P, _LL, _F, _F, _F, _F, _F, _I, _I, _P]),
    "sdb_ddim_step": (_I, [_P, _P, _P, _LL, _F, _F, _F, _F, _I, _P]),
    "sdb_mse_workspace_bytes": (_SZ, []),
    "sdb_min_snr_weight": (_I, [_P, _I, _P, _I, _F, _I, _P, _P]),
    "sdb_threefry_bits": (_I, [_U32, _U32, _P, _LL, _LL, _LL, _P]),
    "sdb_sample_noise": (_I, [_P, _P, _P, _F, _F, _P, _I, _I, _I, _I, _I, _P]),
    "sdb_threefry_randint": (_I, [_U32, _U32, _P, _LL, _LL, _LL, _I, _I, _P]),
    "sdb_mse_fwd_bwd": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _F, _P, _SZ, _P]),
    "sdb_gemm_tc": (_I, [C.POINTER(GemmDesc), _P]),
    "sdb_gemm_simt": (_I, [C.POINTER(GemmDesc), _P]),
    "sdb_conv3x3_c4in_fwd": (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "sdb_conv3x3_c4out_fwd": (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "sdb_conv3x3_c4out_dgrad": (_I, [_P, _P, _P, _I, _I, _I, _I, _P]),
    "sdb_conv3x3_c4_wgrad_workspace_bytes": (_SZ, [_I]),
    "sdb_conv3x3_c4_wgrad": (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _P, _SZ, _P]),
    "sdb_groupnorm_workspace_bytes": (_SZ, [_I, _I, _I]),
    "sdb_groupnorm_fwd": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _F, _I, _P, _SZ, _P]),
    "sdb_groupnorm_bwd": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _P, _SZ, _P]),
    "sdb_layernorm_fwd": (_I, [_P, _P, _P, _P, _P, _LL, _I, _F, _P]),
    "sdb_layernorm_bwd_workspace_bytes": (_SZ, [_LL, _I]),
    "sdb_layernorm_bwd": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _LL, _I, _P, _SZ, _P]),
    "sdb_geglu_fwd": (_I, [_P, _P, _LL, _I, _P]),
    "sdb_geglu_bwd": (_I, [_P, _P, _P, _LL, _I, _P]),
    "sdb_attn_fwd": (_I, [_P, _LL, _P, _LL, _P, _LL, _P, _LL, _P, _P, _I, _I, _I, _I, _I, _F, _P]),
    "sdb_attn_bwd_workspace_bytes": (_SZ, [_I, _I, _I]),
    "sdb_attn_bwd": (_I, [_P, _LL, _P, _LL, _P, _LL, _P, _LL, _P, _LL, _P, _P, _P, _LL, _P, _LL, _P, _LL, _I, _I, _I,
                          _I, _I, _F, _P, _SZ, _P]),
    "sdb_embed_fwd": (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "sdb_embed_bwd": (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "sdb_attn_small_fwd": (_I, [_P, _LL, _P, _LL, _P, _LL, _P, _LL, _P, _I, _I, _I, _I, _F, _I, _P]),
    "sdb_attn_small_bwd": (_I, [_P, _LL, _P, _LL, _P, _LL, _P, _LL, _P, _LL, _P, _P, _LL, _P, _LL, _P, _LL, _I, _I, _I,
                                _I, _F, _I, _P]),
    "sdb_act_fwd": (_I, [_P, _P, _LL, _I, _P]),
    "sdb_act_bwd": (_I, [_P, _P, _P, _LL, _I, _P]),
    "sdb_gather_rows_bf16": (_I, [_P, _P, _P, _LL, _I, _LL, _P]),
    "sdb_nchw_f32_to_nhwc_bf16_pad": (_I, [_P, _P, _I, _I, _I, _I, _P]),
    "sdb_softmax_rows_bf16": (_I, [_P, _P, _LL, _I, _LL, _F, _P]),
    "sdb_vae_sample": (_I, [_P, _P, _P, _I, _I, _I, _F, _P]),
    "sdb_latents_to_nhwc_bf16": (_I, [_P, _F, _I, _P, _I, _I, _I, _P]),
    "sdb_cfg_combine": (_I, [_P, _F, _P, _I, _I, _I, _P]),
    "sdb_image_postprocess": (_I, [_P, _P, _LL, _I, _I, _P]),
    "sdb_cast_f32_bf16": (_I, [_P, _P, _LL, _P]),
    "sdb_nchw_f32_to_nhwc_bf16": (_I, [_P, _P, _I, _I, _I, _P]),
    "sdb_nhwc_bf16_to_nchw_f32": (_I, [_P, _P, _I, _I, _I, _P]),
    "sdb_bump_counters": (_I, [_P, _P, _P, _P]),
    "sdb_upsample2x_fwd": (_I, [_P, _P, _I, _I, _I, _I, _P]),
    "sdb_upsample2x_bwd": (_I, [_P, _P, _I, _I, _I, _I, _P]),
    "sdb_copy2d_bf16": (_I, [_P, _LL, _P, _LL, _LL, _I, _I, _P]),
    "sdb_add_bf16": (_I, [_P, _P, _P, _LL, _P]),
    "sdb_splitk_finish": (_I, [_P, _LL, _I, _LL, _P, _P, _LL, _LL, _P, _LL, _P, _LL, _LL, _I, _P]),
    "sdb_splitk_finish_f32": (_I, [_P, _I, _LL, _P, _LL, _P]),
    "sdb_colsum_workspace_bytes": (_SZ, [_LL, _I]),
    "sdb_colsum_bf16": (_I, [_P, _LL, _I, _LL, _I, _P, _P, _P, _SZ, _P]),
    "sdb_colsum_static_max_scratch_bytes": (_SZ, []),
    "sdb_colsum_bf16_static": (_I, [_P, _LL, _LL, _I, _P, _P]),
    "sdb_silu_fwd": (_I, [_P, _P, _LL, _P]),
    "sdb_silu_bwd": (_I, [_P, _P, _P, _LL, _P]),
    "sdb_timestep_embedding": (_I, [_P, _P, _I, _I, _I, _F, _P]),
}

_lib = None


def experiments_enabled() -> bool:
    return os.environ.get("SDB_EXPERIMENTS") == "1"


def lib() -> C.CDLL:
    """Load libsdb200.so (built in-tree by _build.py). Raises if it is missing — there is no fallback path."""
    global _lib, LIB_PATH
    if _lib is None:
        if experiments_enabled() and os.environ.get("SDB_LIB_PATH"):
            LIB_PATH = os.environ["SDB_LIB_PATH"]  # an experiment build (SDB_BUILD_TAG); never set by the product
        if not os.path.exists(LIB_PATH):
            raise SdbError(
                f"{LIB_PATH} not found: build it with `python -m stable_diffusion_training_b200._build` "
                "(the product path has no CPU or eager fallback)")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        # SDB_EXPERIMENTS=1 SDB_SKIP=sym1,sym2 (timing experiments only, tools/ablate.sh): the named entry points
        # return 0 without launching, so the marginal cost of an operator class inside a CUDA-graph replay can be
        # measured. Results are then meaningless; nothing in the product or the tests sets it, bench.py refuses to
        # run under it, and without the explicit SDB_EXPERIMENTS=1 opt-in the variable is ignored.
        skip = [x for x in os.environ.get("SDB_SKIP", "").split(",") if x] if experiments_enabled() else []
        for name in skip:
            if name in SIGNATURES:
                setattr(l, name, (lambda *a, **k: 0))
        _lib = l
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().sdb_last_error_string().decode("utf-8", "replace")
        raise SdbError(f"{what} failed ({rc}): {msg}")


def stream_ptr(device=None) -> int:
    """Raw handle of torch's current stream on `device` (default: the current device)."""
    import torch
    return torch.cuda.current_stream(device).cuda_stream


def ptr(t) -> int:
    return 0 if t is None else t.data_ptr()
