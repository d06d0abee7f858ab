"""In-tree build of libcnx_b200.so (hand-written sm_100a kernels + C ABI).

`python -m image_classification_b200.build` or `build_library()`; nvcc cross-compiles without a
GPU.  The .so lands next to this file so it travels to the GPU box with the repo snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ_DIR = os.path.join(CSRC, "_obj")
LIB_PATH = os.path.join(HERE, "libcnx_b200.so")

SOURCES = [
    "cnx_util.cu",
    "cnx_api.cu",
    "dwconv7.cu",
    "dwconvk.cu",
    "bnse.cu",
    "ln_rows.cu",
    "gemm2.cu",
    "elementwise.cu",
    "optim.cu",
    "prep.cu",
    "layout.cu",
]
HEADERS = ["cnx_common.cuh", "cnx_kernels.h", os.path.join("..", "..", "include", "cnx_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libcnx_b200.so cannot be built (no fallback path exists)")


def _newest_header_mtime() -> float:
    return max(os.path.getmtime(os.path.join(CSRC, h)) for h in HEADERS)


def _compile_one(nvcc: str, src: str, force: bool) -> tuple[str, str]:
    src_path = os.path.join(CSRC, src)
    obj_path = os.path.join(OBJ_DIR, src.replace(".cu", ".o"))
    if (not force and os.path.exists(obj_path)
            and os.path.getmtime(obj_path) > max(os.path.getmtime(src_path), _newest_header_mtime())):
        return obj_path, ""
    cmd = [nvcc, *NVCC_FLAGS, "-c", src_path, "-o", obj_path]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{res.stdout}\n{res.stderr}")
    return obj_path, res.stderr


def build_library(force: bool = False, verbose: bool = False) -> str:
    """Compile every CUDA source for sm_100a and link libcnx_b200.so. Returns its path."""
    nvcc = _nvcc()
    os.makedirs(OBJ_DIR, exist_ok=True)
    with ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as ex:
        results = list(ex.map(lambda s: _compile_one(nvcc, s, force), SOURCES))
    objs = [r[0] for r in results]
    log = "\n".join(r[1] for r in results if r[1])
    if log:
        with open(os.path.join(OBJ_DIR, "ptxas.log"), "a") as f:
            f.write(log)
        if verbose:
            print(log)
    need_link = force or not os.path.exists(LIB_PATH) or any(
        os.path.getmtime(o) > os.path.getmtime(LIB_PATH) for o in objs)
    if need_link:
        cmd = [nvcc, "-shared", "-o", LIB_PATH, *objs, "-lcudart_static", "-ldl", "-lpthread", "-lrt"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    return LIB_PATH


if __name__ == "__main__":
    path = build_library(force="--force" in sys.argv, verbose=True)
    print(path)
