"""Build libssl4p.so (the C-ABI library of sm_100a kernels) in-tree with nvcc.

    python -m ssl4polyp_b200.csrc.build [--force]

nvcc cross-compiles for sm_100a without a GPU; the resulting `ssl4polyp_b200/libssl4p.so` is git-ignored
but travels to the GPU box with the repo snapshot.  Objects are rebuilt only when a source/header is newer.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
ROOT = os.path.dirname(PKG)
OBJ_DIR = os.path.join(HERE, "build")
LIB_PATH = os.path.join(PKG, "libssl4p.so")
SOURCES = ["api.cu", "gemm_tcgen05.cu", "tokens.cu", "layernorm.cu", "attention_tc.cu", "loss.cu", "optim.cu", "input_pipeline.cu", "augment.cu", "host_pack.cu"]
HEADERS = [os.path.join(HERE, "common.cuh"), os.path.join(HERE, "attn_common.cuh"), os.path.join(ROOT, "include", "ssl4p.h")]
# test infrastructure: an independent on-device implementation the product kernels are cross-checked against (tests only)
CHECK_DIR = os.path.join(ROOT, "tests", "csrc")
CHECK_SOURCES = ["attention_mma.cu"]
CHECK_LIB_PATH = os.path.join(ROOT, "tests", "libssl4p_check.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC", "-I" + os.path.join(ROOT, "include"), "-I" + HERE]


def _stale(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _compile(src: str, force: bool, src_dir: str = HERE) -> str:
    obj = os.path.join(OBJ_DIR, src.replace(".cu", ".o"))
    path = os.path.join(src_dir, src)
    if force or _stale(obj, [path] + HEADERS):
        cmd = [NVCC] + FLAGS + ["-I" + src_dir, "-c", path, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    return obj


def build(force: bool = False, verbose: bool = True) -> str:
    os.makedirs(OBJ_DIR, exist_ok=True)
    srcs = [s for s in SOURCES if os.path.exists(os.path.join(HERE, s))]
    missing = sorted(set(SOURCES) - set(srcs))
    if missing:
        raise RuntimeError(f"missing kernel sources: {missing}")
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile(s, force), srcs))
    if force or _stale(LIB_PATH, objs):
        cmd = [NVCC, "-shared", "-o", LIB_PATH] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
        if verbose:
            print(f"built {LIB_PATH}")
    elif verbose:
        print(f"up to date: {LIB_PATH}")
    if os.path.isdir(CHECK_DIR):                              # tests/libssl4p_check.so (skipped in a source tree without tests/)
        cobjs = [_compile(s, force, CHECK_DIR) for s in CHECK_SOURCES]
        if force or _stale(CHECK_LIB_PATH, cobjs):
            r = subprocess.run([NVCC, "-shared", "-o", CHECK_LIB_PATH] + cobjs + ["-gencode", "arch=compute_100a,code=sm_100a"],
                               capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
            if verbose:
                print(f"built {CHECK_LIB_PATH}")
    return LIB_PATH


if __name__ == "__main__":
    build(force="--force" in sys.argv)
