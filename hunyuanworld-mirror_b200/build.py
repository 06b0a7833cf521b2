"""In-tree build of libwm_b200.so (and the native self-test) with nvcc for sm_100a.

    python hunyuanworld-mirror_b200/build.py [--force]

nvcc cross-compiles without a GPU; the resulting .so lives in hunyuanworld-mirror_b200/lib/ (git-ignored,
but shipped to the GPU box by gpurun).  No torch headers are involved: the library has a plain C ABI
(include/wm_b200.h) and is bound with ctypes.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
LIBDIR = os.path.join(PKG, "lib")
LIB = os.path.join(LIBDIR, "libwm_b200.so")
SOURCES = ["gemm.cu", "attn.cu", "attn_split.cu", "attn_pipe.cu", "rowops.cu", "conv.cu", "headops.cu", "camhead.cu", "priors.cu", "splat.cu", "raster.cu", "loader.cu", "capi.cu"]
HEADERS = ["common.cuh", "ptx.cuh", "ops.cuh", "attn_common.cuh", os.path.join(ROOT, "include", "wm_b200.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--shared",
              "-Xcompiler", "-fPIC"]
SELFTEST_SRC = os.path.join(ROOT, "tests", "native", "selftest.cu")
SELFTEST_BIN = os.path.join(ROOT, "tests", "native", "selftest")


def _nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: libwm_b200 cannot be built (there is no prebuilt or CPU fallback)")
    return exe


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force: bool = False, verbose: bool = False) -> str:
    """One object file per translation unit (compiled in parallel, rebuilt only when its source or a header changed),
    then one link.  No relocatable device code: no kernel calls across translation units."""
    from concurrent.futures import ThreadPoolExecutor

    hdrs = [h if os.path.isabs(h) else os.path.join(CSRC, h) for h in HEADERS]
    objdir = os.path.join(PKG, "build")
    os.makedirs(objdir, exist_ok=True)
    os.makedirs(LIBDIR, exist_ok=True)
    flags = [f for f in NVCC_FLAGS if f != "--shared"]

    def compile_one(name):
        src, obj = os.path.join(CSRC, name), os.path.join(objdir, name[:-3] + ".o")
        if force or _stale(obj, [src] + hdrs):
            cmd = [_nvcc()] + flags + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", obj, src]
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError(f"nvcc failed on {name}:\n" + r.stdout + r.stderr)
            if verbose:
                print(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(len(SOURCES), os.cpu_count() or 4)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    if force or _stale(LIB, objs):
        cmd = [_nvcc(), "--shared", "-Xcompiler", "-fPIC", "-o", LIB] + objs + ["-lcudart", "-ldl"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc (link) failed:\n" + r.stdout + r.stderr)
    return LIB


def build_selftest(force: bool = False) -> str:
    if force or _stale(SELFTEST_BIN, [SELFTEST_SRC, LIB, os.path.join(ROOT, "include", "wm_b200.h")]):
        cmd = [_nvcc(), "-O2", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-o", SELFTEST_BIN, SELFTEST_SRC, "-L" + LIBDIR,
               "-lwm_b200", "-Xlinker", "-rpath", "-Xlinker", "$ORIGIN/../../hunyuanworld-mirror_b200/lib"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc (selftest) failed:\n" + r.stdout + r.stderr)
    return SELFTEST_BIN


if __name__ == "__main__":
    force = "--force" in sys.argv
    print(build_library(force=force, verbose="-v" in sys.argv))
    print(build_selftest(force=force))
