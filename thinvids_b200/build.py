"""Builds thinvids_b200/libtvscale.so (the C-ABI library of include/tvscale.h) in-tree with nvcc for sm_100a.

    python -m thinvids_b200.build [--force]

The library is git-ignored but travels with the tree (it is what the GPU box loads).  nvcc cross-compiles
without a GPU, so this also runs on CPU-only machines.
"""
from __future__ import annotations

import os
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libtvscale.so")
SOURCES = ["filter_tables.cpp", "kernels_generic.cu", "kernels_fused_w256.cu", "kernels_fused_w128.cu", "kernels_stream.cu", "kernels_slab.cu", "kernels_bwdif.cu", "tvscale_abi.cu"]
HEADERS = ["swscale_shim.cpp", "filter_tables.h", "tvs_internal.h", "kernels_fused_impl.cuh", "scale_device.cuh", os.path.join("..", "..", "include", "tvscale.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC,-ffp-contract=off,-fvisibility=hidden", "--fmad=true"]
NVCC_FLAGS += os.environ.get("TVS_NVCC_EXTRA", "").split()   # experiments, e.g. -DTVS_HALVES=1


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.sep not in cand or os.path.exists(cand)):
            return cand
    raise RuntimeError("nvcc not found")


def needs_build() -> bool:
    if not os.path.exists(LIB) or not os.path.exists(os.path.join(PKG, "libswscale_tvs.so")):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    build_dir = os.path.join(PKG, "build")
    os.makedirs(build_dir, exist_ok=True)

    def compile_one(src):
        obj = os.path.join(build_dir, os.path.splitext(src)[0] + ".o")
        cmd = [_nvcc(), *NVCC_FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose and src.endswith(".cu"):
            cmd += ["-Xptxas", "-v"]
        proc = subprocess.run(cmd, capture_output=True, text=True)
        return src, obj, cmd, proc

    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(max_workers=min(len(SOURCES), os.cpu_count() or 1)) as pool:
        results = list(pool.map(compile_one, SOURCES))
    objs = []
    for src, obj, cmd, proc in results:
        if verbose or proc.returncode:
            print(" ".join(cmd))
            sys.stdout.write(proc.stdout)
            sys.stdout.write(proc.stderr)
        if proc.returncode:
            raise subprocess.CalledProcessError(proc.returncode, cmd)
        objs.append(obj)
    tmp = LIB + ".tmp.%d" % os.getpid()
    subprocess.check_call([_nvcc(), "-shared", "-o", tmp, *objs, "-gencode", "arch=compute_100a,code=sm_100a",
                           "-cudart", "static"])
    os.replace(tmp, LIB)
    build_shim()
    return LIB


SHIM_LIB = os.path.join(PKG, "libswscale_tvs.so")


def build_shim() -> str:
    """libswscale's sws_getContext / sws_scale / sws_freeContext on top of libtvscale.so (csrc/swscale_shim.cpp): a small
    host-only library that finds libtvscale.so next to itself."""
    src = os.path.join(CSRC, "swscale_shim.cpp")
    tmp = SHIM_LIB + ".tmp.%d" % os.getpid()
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-fvisibility=hidden", "-o", tmp, src,
                           "-L" + PKG, "-l:libtvscale.so", "-Wl,-rpath,$ORIGIN"])
    os.replace(tmp, SHIM_LIB)
    return SHIM_LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
