"""In-tree build of the C-ABI CUDA library (libsdb200.so) for sm_100a with plain nvcc.

The .so is git-ignored but travels to the GPU box with the gpurun snapshot. No torch headers are involved: the
library's signatures are plain pointers and sizes (include/sdb200.h).
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")
# SDB_BUILD_TAG=<tag> (with NVCC_EXTRA=-D...) builds an experiment variant next to the product library
# (libsdb200_<tag>.so, objects under _build_<tag>/); _lib loads it only with SDB_EXPERIMENTS=1 SDB_LIB_PATH=<path>.
_TAG = os.environ.get("SDB_BUILD_TAG", "")
BUILD_DIR = os.path.join(HERE, "_build" + ("_" + _TAG if _TAG else ""))
LIB_PATH = os.path.join(HERE, "libsdb200" + ("_" + _TAG if _TAG else "") + ".so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-I", INCLUDE,
] + [f for f in os.environ.get("NVCC_EXTRA", "").split() if f]  # e.g. -DSDB_GEMM_EXPERIMENTS, -DSDB_ATTN_TIMELINE


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def sources() -> list[str]:
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _digest(path: str) -> str:
    h = hashlib.sha1()
    h.update(" ".join(NVCC_FLAGS).encode())
    with open(path, "rb") as f:
        h.update(f.read())
    # headers shared by every translation unit
    for hdr in sorted(os.listdir(CSRC)):
        if hdr.endswith((".cuh", ".inc", ".h")):
            with open(os.path.join(CSRC, hdr), "rb") as f:
                h.update(f.read())
    with open(os.path.join(INCLUDE, "sdb200.h"), "rb") as f:
        h.update(f.read())
    return h.hexdigest()


def _compile_one(src: str, verbose: bool) -> str:
    os.makedirs(BUILD_DIR, exist_ok=True)
    obj = os.path.join(BUILD_DIR, os.path.basename(src)[:-3] + ".o")
    stamp = obj + ".sha1"
    dig = _digest(src)
    if os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == dig:
        return obj
    cmd = [_nvcc(), *NVCC_FLAGS, "-c", src, "-o", obj]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    if verbose:
        sys.stderr.write(r.stderr)
    with open(stamp, "w") as f:
        f.write(dig)
    return obj


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every csrc/*.cu for sm_100a and link libsdb200.so. Returns the library path."""
    if force:
        for f in os.listdir(BUILD_DIR) if os.path.isdir(BUILD_DIR) else []:
            os.remove(os.path.join(BUILD_DIR, f))
    srcs = sources()
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile_one(s, verbose), srcs))
    newest = max(os.path.getmtime(o) for o in objs)
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < newest:
        cmd = [_nvcc(), "-shared", "-o", LIB_PATH, *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-lcudart"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
