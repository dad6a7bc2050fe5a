"""In-tree nvcc build of libsigsvgd_b200.so (sm_100a only; cross-compiles without a GPU).

Every ``csrc/*.cu`` is compiled to an object file under ``sigsvgd_b200/build/`` (in parallel, only when it or
a header changed) and the objects are linked into the shared library."""
from __future__ import annotations

import glob
import os
import shutil
import subprocess
from concurrent.futures import ThreadPoolExecutor

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "csrc")
_OBJ = os.path.join(_HERE, "build")
_OUT = os.path.join(_HERE, "libsigsvgd_b200.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
]
LINK_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-cudart", "static"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def sources():
    return sorted(glob.glob(os.path.join(_SRC, "*.cu")))


def _headers():
    return glob.glob(os.path.join(_SRC, "*.cuh")) + glob.glob(os.path.join(_HERE, "..", "include", "*.h"))


def _obj(src: str) -> str:
    return os.path.join(_OBJ, os.path.basename(src)[:-3] + ".o")


def is_stale() -> bool:
    if not os.path.exists(_OUT):
        return True
    t = os.path.getmtime(_OUT)
    return any(os.path.getmtime(p) > t for p in sources() + _headers())


def _run(cmd, verbose):
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or res.returncode != 0:
        print(" ".join(cmd))
        print(res.stdout)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed ({res.returncode}):\n{res.stdout[-6000:]}")
    return res.stdout


def build(verbose: bool = False, force: bool = False) -> str:
    if not force and not is_stale():
        return _OUT
    os.makedirs(_OBJ, exist_ok=True)
    nvcc = _nvcc()
    hdr_t = max(os.path.getmtime(p) for p in _headers())
    todo = []
    for src in sources():
        o = _obj(src)
        if force or not os.path.exists(o) or os.path.getmtime(o) < max(os.path.getmtime(src), hdr_t):
            todo.append(src)
    extra = ["-Xptxas", "-v"] if verbose else []
    with ThreadPoolExecutor(max_workers=min(8, max(1, len(todo)))) as pool:
        list(pool.map(lambda s: _run([nvcc, *NVCC_FLAGS, *extra, "-c", s, "-o", _obj(s)], verbose), todo))
    # stale objects of deleted sources must not be linked
    keep = {_obj(s) for s in sources()}
    for o in glob.glob(os.path.join(_OBJ, "*.o")):
        if o not in keep:
            os.remove(o)
    # no -lcuda: the one driver entry point used (cuTensorMapEncodeTiled) is resolved at run time
    # through cudaGetDriverEntryPoint, so the library loads on a box without a GPU driver
    _run([nvcc, *LINK_FLAGS, "-o", _OUT, *sorted(keep)], verbose)
    return _OUT
