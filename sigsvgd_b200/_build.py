"""In-tree nvcc build of libsigsvgd_b200.so (sm_100a only; cross-compiles without a GPU)."""
from __future__ import annotations

import glob
import os
import shutil
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "csrc")
_OUT = os.path.join(_HERE, "libsigsvgd_b200.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "-shared",
    "-cudart", "static",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def sources():
    return sorted(glob.glob(os.path.join(_SRC, "*.cu")))


def is_stale() -> bool:
    if not os.path.exists(_OUT):
        return True
    t = os.path.getmtime(_OUT)
    deps = sources() + glob.glob(os.path.join(_SRC, "*.cuh")) + glob.glob(
        os.path.join(_HERE, "..", "include", "*.h"))
    return any(os.path.getmtime(p) > t for p in deps)


def build(verbose: bool = False, force: bool = False) -> str:
    if not force and not is_stale():
        return _OUT
    # no -lcuda: the one driver entry point used (cuTensorMapEncodeTiled) is resolved at run time
    # through cudaGetDriverEntryPoint, so the library loads on a box without a GPU driver
    cmd = [_nvcc(), *NVCC_FLAGS, "-o", _OUT, *sources()]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
        print(" ".join(cmd))
    env = dict(os.environ)
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)
    if verbose or res.returncode != 0:
        print(res.stdout)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed ({res.returncode}):\n{res.stdout[-4000:]}")
    return _OUT
