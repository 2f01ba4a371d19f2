"""CPU tests of the host logic: parameter census, layer program, masks, arena layout, attention quirk table, ABI."""
import ctypes
import os
import re

import numpy as np
import pytest

from oracle import unet as OU
from stable_diffusion_training_b200 import _lib, lion_quant as LQ, unet as PU
from stable_diffusion_training_b200.training_utils import TrainingConfig, backward_param_order, create_mask


@pytest.mark.parametrize("name,leaves,total", [("SD15", 686, 859_520_964), ("SD21", 686, 865_910_724), ("TINY", 208, 792_964)])
def test_param_census_matches_public_and_oracle(name, leaves, total):
    ps = PU.param_shapes(getattr(PU, name))
    os_ = OU.param_shapes(getattr(OU, name))
    assert ps == os_  # product enumeration == independent oracle enumeration (names, shapes, order)
    assert len(ps) == leaves
    assert sum(int(np.prod(s)) for s in ps.values()) == total


def test_masks_and_quantised_census():
    """256 quantised leaves / 831 201 280 elements; 443 844 decay-exempt elements."""
    shapes = PU.param_shapes(PU.SD15)
    tree = LQ.unflatten_dict({k: None for k in shapes})
    cfg = TrainingConfig()
    q = LQ.flatten_dict(create_mask(tree, list(cfg.excluded_layer_from_quantization)))
    wd = LQ.flatten_dict(create_mask(tree, list(cfg.excluded_layer_pattern_from_weight_decay)))
    qn = [k for k, v in q.items() if v]
    assert len(qn) == 256
    assert sum(int(np.prod(shapes[k])) for k in qn) == 831_201_280
    assert all(int(np.prod(shapes[k])) % 16 == 0 for k in qn)
    assert sum(int(np.prod(shapes[k])) for k, v in wd.items() if not v) == 443_844
    # exact path-component match (training_utils.py:120-129): "time_emb_proj" is excluded, kernels under
    # "time_embedding" too, but e.g. "to_out_0/kernel" is not caught by "conv_out"
    assert not q["down_blocks_0/resnets_0/time_emb_proj/kernel"] and not q["time_embedding/linear_1/kernel"]
    assert q["down_blocks_0/attentions_0/transformer_blocks_0/attn1/to_out_0/kernel"]


def test_backward_order_is_a_permutation():
    for cfg in (PU.SD15, PU.TINY):
        order = backward_param_order(cfg)
        assert sorted(order) == sorted(PU.param_shapes(cfg))
        assert order[0].startswith("conv_out") or order[0].startswith("conv_norm_out")
        assert order[-1].startswith("time_embedding")


def test_key_multiplicity_table():
    """SURVEY A.3 chunk table of the patched reference."""
    m = PU.key_multiplicity(64, 77)
    assert [i for i, v in enumerate(m) if v == 2] == list(range(13, 64)) and max(m) == 2
    m = PU.key_multiplicity(64, 227)
    assert [i for i, v in enumerate(m) if v == 2] == list(range(163, 192))
    m = PU.key_multiplicity(144, 227)
    assert [i for i, v in enumerate(m) if v == 2] == list(range(83, 144))
    for sq, skv in ((4096, 77), (1024, 77), (256, 77), (4096, 4096), (144, 77), (64, 64)):
        assert set(PU.key_multiplicity(sq, skv)) == {1}
    assert PU.key_multiplicity(64, 77) == OU.key_multiplicity(64, 77)


def test_layer_program_skip_channels():
    prog = PU.layer_program(PU.SD15)
    ups = [op[2] for op in prog if op[0] == "resnet" and op[1].startswith("up_blocks")]
    assert ups == [2560, 2560, 2560, 2560, 2560, 1920, 1920, 1280, 960, 960, 640, 640]
    assert sum(1 for op in prog if op[0] == "resnet") == 22 and sum(1 for op in prog if op[0] == "xformer") == 16


def test_c_abi_exports_every_declared_symbol():
    """The shared library loads and exports exactly the entry points include/sdb200.h declares (no compute calls)."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "sdb200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(sdb_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    lib = _lib.lib()
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in sdb200.h but not exported"
    assert declared == set(_lib.SIGNATURES), (declared ^ set(_lib.SIGNATURES))
    assert lib.sdb_abi_version() == 6
    assert isinstance(lib.sdb_last_error_string(), bytes)
    # struct layout agreement between ctypes and the header
    assert ctypes.sizeof(_lib.View4) == 48


def test_tile_and_splitk_pickers_return_legal_shapes():
    """The GEMM tile picker and the split-K decision are pure host logic: every answer must be a shape sdb_gemm_tc
    accepts (block_n a multiple of 16 <= 256, 1 or 2 halves, a pixel box of 128 * halves, no empty last K range)."""
    from stable_diffusion_training_b200 import ops
    shapes = [(320, (64, 64, 8), True, 45), (640, (32, 32, 8), False, 90), (1280, (16, 16, 8), True, 180),
              (1280, (8, 8, 8), False, 180), (1280, (16, 16, 8), True, 360), (2560, (8, 8, 8), False, 180),
              (10240, (2048, 1, 1), True, 20), (320, (616, 1, 1), True, 12), (200, (300, 1, 1), True, 3)]
    for N, pix, b_mn, nkb in shapes:
        mt, bn = ops._pick_tile(N, pix, b_mn, nkb)
        assert mt in (1, 2) and bn % 16 == 0 and 16 <= bn <= 256
        bw, bh, bn_ = ops._pixel_box(*pix, 128 * mt)
        assert bw * bh * bn_ == 128 * mt
        sk = ops._pick_splitk(N, pix, b_mn, nkb, mt, bn)
        if sk is not None:
            smt, sbn, sp = sk
            assert smt in (1, 2) and sbn % 16 == 0 and sbn <= 256 and 2 <= sp <= 8
            per = -(-nkb // sp)
            assert per * (sp - 1) < nkb  # the last K range is not empty
            t_split = ops._model_time(N, pix, b_mn, nkb, smt, sbn, sp)
            t_plain = ops._model_time(N, pix, b_mn, nkb, mt, bn, 1)
            assert t_split < 0.85 * t_plain
    # the 8x8 level of SD-1.5 (512 pixels, N = 1280, K = 9 * 1280) is the case split-K exists for
    assert ops._pick_splitk(1280, (8, 8, 8), True, 180, *ops._pick_tile(1280, (8, 8, 8), True, 180)) is not None
    # the 64x64 level has enough tiles: no split
    assert ops._pick_splitk(320, (64, 64, 8), True, 45, *ops._pick_tile(320, (64, 64, 8), True, 45)) is None
