"""In-tree nvcc build of libqpolicy.so for sm_100a (no JIT cache: the .so travels with the repo)."""
from __future__ import annotations

import os
import shutil
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libqpolicy.so")
SOURCES = [os.path.join(_HERE, "csrc", "qp_api.cu")]
HEADERS = [os.path.join(_HERE, "csrc", "qp_kernels.cuh"), os.path.join(_HERE, "csrc", "qp_wide.cuh"),
           os.path.join(_HERE, "csrc", "qp_tc.cuh"), os.path.join(_HERE, "csrc", "qp_ep_tc.cuh"),
           os.path.join(_HERE, "csrc", "qp_probe.cuh"), os.path.join(_HERE, "csrc", "qp_fanout.cuh"), os.path.join(_HERE, "csrc", "qp_ep_c2.cuh"),
           os.path.join(_HERE, "csrc", "qp_ep_c2_kernel.cuh"), os.path.join(_HERE, "csrc", "qp_ep_c2_phaseA.inl"),
           os.path.join(_HERE, "csrc", "qp_ep_c2_backward.inl"), os.path.join(_HERE, "csrc", "qp_ep_c2_phaseB.inl"),
           os.path.join(os.path.dirname(_HERE), "include", "qpolicy.h")]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared", "-cudart", "static",
]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: cannot build libqpolicy.so")


def stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.exists(p) and os.path.getmtime(p) > t for p in SOURCES + HEADERS)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every CUDA source of the package into libqpolicy.so (sm_100a, -lineinfo)."""
    if not force and not stale():
        return LIB_PATH
    cmd = [_nvcc(), *NVCC_FLAGS, "-o", LIB_PATH, *SOURCES, "-ldl"]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return LIB_PATH
