"""Allocation-free entry point sdb_colsum_bf16_static (include/sdb200.h: no counterpart in the reference):
its partial sums live in a file-scope static __device__ array of 2 MiB instead of a
caller-provided workspace. Called through the C ABI (ctypes), checked against the existing entry point sdb_colsum_bf16
within the tolerance of tests/test_ops_gpu.py's column-sum checks (rtol 1e-4, atol 1e-2), on two calls in a row (the
second with other data and another shape: the array keeps the first call's contents), plus the size limit and the
"allocates nothing" claim as far as the process can see it. The file is named to be collected last."""

import pytest
import torch

from stable_diffusion_training_b200 import _lib

pytestmark = pytest.mark.gpu


def _reference_colsum(l, x, out, ws):
    rows, cols = x.shape
    _lib.check(l.sdb_colsum_bf16(x.data_ptr(), x.stride(0), 1, rows, cols, out.data_ptr(), None, ws.data_ptr(),
                                 ws.numel(), _lib.stream_ptr()), "sdb_colsum_bf16")


def test_static_scratch_colsum_matches_the_workspace_entry_point(cuda):
    l = _lib.lib()
    limit = l.sdb_colsum_static_max_scratch_bytes()
    assert limit == 2 << 20
    g = torch.Generator().manual_seed(11)
    # two different shapes / contents, back to back on one stream (bias-gradient shapes of the UNet: conv 320 at 64x64,
    # B = 8; FF 2560 at 32x32, B = 2)
    cases = [(32768, 320), (2048, 2560)]
    xs = [(torch.randn(r, c, generator=g) * (i + 1)).to(torch.bfloat16).to(cuda) for i, (r, c) in enumerate(cases)]
    outs = [torch.full((c,), float("nan"), device=cuda) for _, c in cases]
    refs = [torch.empty(c, device=cuda) for _, c in cases]
    ws = torch.empty(l.sdb_colsum_workspace_bytes(32768, 2560), dtype=torch.uint8, device=cuda)
    for x, ref in zip(xs, refs):
        _reference_colsum(l, x, ref, ws)
    # one untimed call first: CUDA loads a kernel's code lazily at its first launch, which may take device memory that
    # has nothing to do with the entry point (the second-stage kernel is first launched here in a fresh process)
    warm = torch.empty(cases[0][1], device=cuda)
    assert l.sdb_colsum_bf16_static(xs[0].data_ptr(), xs[0].stride(0), cases[0][0], cases[0][1], warm.data_ptr(),
                                    _lib.stream_ptr()) == 0, l.sdb_last_error_string().decode()
    torch.cuda.synchronize()
    alloc0, reserved0 = torch.cuda.memory_allocated(), torch.cuda.memory_reserved()
    free0 = torch.cuda.mem_get_info()[0]
    for x, out in zip(xs, outs):   # two calls in a row, nothing in between
        rc = l.sdb_colsum_bf16_static(x.data_ptr(), x.stride(0), x.shape[0], x.shape[1], out.data_ptr(),
                                      _lib.stream_ptr())
        assert rc == 0, l.sdb_last_error_string().decode()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == alloc0 and torch.cuda.memory_reserved() == reserved0
    assert torch.cuda.mem_get_info()[0] == free0   # no device memory taken behind torch's back either
    for x, out, ref in zip(xs, outs, refs):
        assert torch.isfinite(out).all()
        assert torch.allclose(out, ref, rtol=1e-4, atol=1e-2)
        assert torch.allclose(out, x.float().sum(0), rtol=1e-4, atol=1e-2)
    # the first shape again after the second: stale partial sums of the other shape must not leak into the result
    again = torch.full((cases[0][1],), float("nan"), device=cuda)
    rc = l.sdb_colsum_bf16_static(xs[0].data_ptr(), xs[0].stride(0), cases[0][0], cases[0][1], again.data_ptr(),
                                  _lib.stream_ptr())
    assert rc == 0, l.sdb_last_error_string().decode()
    torch.cuda.synchronize()
    assert torch.equal(again, outs[0])


def test_static_scratch_colsum_rejects_what_does_not_fit(cuda):
    l = _lib.lib()
    rows, cols = 262144, 2560   # 1024 row blocks x 2560 columns x 4 B = 10 MiB of partial sums > 2 MiB
    x = torch.zeros((8, cols), dtype=torch.bfloat16, device=cuda)   # never read: the call must fail before any launch
    out = torch.zeros(cols, device=cuda)
    n0 = l.sdb_launch_count()
    rc = l.sdb_colsum_bf16_static(x.data_ptr(), cols, rows, cols, out.data_ptr(), _lib.stream_ptr())
    assert rc == -3 and l.sdb_launch_count() == n0
    msg = l.sdb_last_error_string().decode()
    assert "sdb_colsum_bf16_static" in msg and "static scratch" in msg and str(2 << 20) in msg
    # argument checks
    assert l.sdb_colsum_bf16_static(x.data_ptr(), cols, 8, 12, out.data_ptr(), _lib.stream_ptr()) == -1
    assert l.sdb_colsum_bf16_static(None, cols, 8, cols, out.data_ptr(), _lib.stream_ptr()) == -1
    # and the library still works afterwards
    rc = l.sdb_colsum_bf16_static(x.data_ptr(), cols, 8, cols, out.data_ptr(), _lib.stream_ptr())
    torch.cuda.synchronize()
    assert rc == 0 and float(out.abs().max()) == 0.0
