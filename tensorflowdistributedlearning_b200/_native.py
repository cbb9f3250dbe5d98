"""Build + load the sm_100a native library (``lib/libtdl_b200.so``) and expose its C ABI.

The library is plain CUDA C++ with an ``extern "C"`` surface (declared in ``include/tdl_b200.h``)
so it can be driven from Python (ctypes, here), C, or any runtime that can pass device pointers
and a ``cudaStream_t``.  It is built in-tree with nvcc for ``sm_100a`` only -- there is no
multi-arch dispatch and no JIT cache.

The reference has no native code of its own; these kernels stand in for
the cuDNN / cuBLAS / Eigen / NCCL kernels its TensorFlow runtime dispatches.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
import sys
import threading
from pathlib import Path

_PKG_DIR = Path(__file__).resolve().parent
_REPO = _PKG_DIR.parent
CSRC = _REPO / "csrc"
INCLUDE = _REPO / "include"
LIB_DIR = _PKG_DIR / "lib"
LIB_PATH = LIB_DIR / "libtdl_b200.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--use_fast_math",
    "-Xptxas", "-v",
]

_lock = threading.Lock()
_lib = None


def sources():
    return sorted(CSRC.glob("*.cu"))


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def needs_build() -> bool:
    if not LIB_PATH.exists():
        return True
    t = LIB_PATH.stat().st_mtime
    deps = list(sources()) + list(CSRC.glob("*.cuh")) + list(INCLUDE.glob("*.h"))
    return any(p.stat().st_mtime > t for p in deps)


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every ``csrc/*.cu`` for sm_100a and link ``libtdl_b200.so`` (parallel nvcc -c)."""
    import fcntl

    LIB_DIR.mkdir(parents=True, exist_ok=True)
    # one builder at a time across PROCESSES too: the towers of a torchrun job import the package at the same moment,
    # and a stale library would otherwise be rebuilt by all of them into the same object files
    with _lock, open(LIB_DIR / ".build.lock", "w") as lock_file:
        fcntl.flock(lock_file, fcntl.LOCK_EX)
        if not force and not needs_build():  # (a tower that waited for the lock finds the library fresh)
            return LIB_PATH
        obj_dir = LIB_DIR / "obj"
        obj_dir.mkdir(exist_ok=True)
        nvcc = _nvcc()
        procs = []
        for src in sources():
            obj = obj_dir / (src.stem + ".o")
            cmd = [nvcc, *NVCC_FLAGS, "-I", str(CSRC), "-I", str(INCLUDE), "-c", str(src), "-o", str(obj)]
            procs.append((src, obj, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs, log = [], []
        for src, obj, p in procs:
            out, _ = p.communicate()
            log.append(f"==== {src.name} ====\n{out}")
            if p.returncode != 0:
                raise RuntimeError(f"nvcc failed for {src}:\n{out}")
            objs.append(str(obj))
        (LIB_DIR / "build.log").write_text("\n".join(log))
        if verbose:
            print("\n".join(log))
        tmp = LIB_PATH.with_suffix(".so.tmp")
        cmd = [nvcc, "-shared", "-o", str(tmp), *objs, "-lcudart"]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}")
        os.replace(tmp, LIB_PATH)
        return LIB_PATH


class NativeError(RuntimeError):
    pass


def available() -> bool:
    """True when the library exists on disk (it may still need a GPU to *run*)."""
    return LIB_PATH.exists()


def lib() -> ctypes.CDLL:
    """Load (building if needed) the native library.  Fails loudly -- there is no silent fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if needs_build():
        build()
    # cudart must be resolvable: torch ships one; load torch first so its libs are mapped.
    import torch  # noqa: F401

    with _lock:
        if _lib is None:
            _lib = ctypes.CDLL(str(LIB_PATH), mode=ctypes.RTLD_GLOBAL)
            _lib.tdl_last_error.restype = ctypes.c_char_p
            _lib.tdl_bn_scratch_floats.restype = ctypes.c_int64
    return _lib


# kernels launched per C-ABI entry point (default 1); used for the bench's "gpu_launches" count
_KERNELS_PER_CALL = {"tdl_bn_fwd_stats": 2, "tdl_bn_bwd_reduce": 2, "tdl_depthwise_bwd": 2, "tdl_gap_conv1x1_bwd": 2}
kernel_launches = 0


def check(rc: int, what: str = "") -> None:
    global kernel_launches
    kernel_launches += _KERNELS_PER_CALL.get(what, 1)
    if rc != 0:
        msg = lib().tdl_last_error()
        raise NativeError(f"{what} failed (rc={rc}): {msg.decode() if msg else ''}")


# ---- argument helpers -------------------------------------------------------------------------
c_void_p = ctypes.c_void_p
c_int = ctypes.c_int
c_int64 = ctypes.c_int64
c_float = ctypes.c_float


def ptr(t) -> ctypes.c_void_p:
    """Device pointer of a torch tensor (or NULL for None)."""
    if t is None:
        return ctypes.c_void_p(0)
    return ctypes.c_void_p(t.data_ptr())


def stream_ptr() -> ctypes.c_void_p:
    import torch

    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
    print(LIB_PATH)
