"""The C-ABI library loads and exports exactly the symbols include/sdb200.h declares (no compute calls: CPU-only)."""
import ctypes
import os
import re

from stable_diffusion_training_b200 import _build, _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "sdb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sdb_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported_and_bound():
    path = _build.build()
    names = _declared()
    assert len(names) >= 35
    l = ctypes.CDLL(path)
    for n in names:
        assert hasattr(l, n), f"{n} declared in sdb200.h but not exported"
    assert set(names) == set(_lib.SIGNATURES), set(names) ^ set(_lib.SIGNATURES)


def test_library_level_calls_without_gpu():
    l = _lib.lib()
    assert l.sdb_abi_version() == 6
    assert l.sdb_sumsq_workspace_bytes() > 0
    assert l.sdb_launch_count() >= 0
    # argument validation happens before any CUDA call and reports through the error string
    rc = l.sdb_gemm_tc(None, None)
    assert rc < 0 and b"null descriptor" in l.sdb_last_error_string()


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    try:
        _lib.lib()
    except _lib.SdbError as e:
        assert "no CPU or eager fallback" in str(e)
    else:
        raise AssertionError("expected SdbError")
