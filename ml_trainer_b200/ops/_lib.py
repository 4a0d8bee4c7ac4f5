"""ctypes binding to the in-tree sm_100a kernel library (``lib/libmlt_b200.so``).

The library exposes a plain C ABI (``include/mlt_b200.h``): every entry point takes raw
device pointers, sizes and a ``cudaStream_t`` and returns a ``cudaError_t``-style int
(0 = success).  That keeps the build to a few seconds of ``nvcc`` per file, has no
link-time dependency on libtorch or libcuda (the driver API is resolved at run time), and
lets the same ``.so`` be driven from C++ or Python.

Policy: on a machine with a CUDA device the native library is REQUIRED: a missing or
stale ``.so`` raises instead of silently falling back to eager PyTorch.  On a CPU-only
machine the ops layer uses its PyTorch statements of the same math.
"""
from __future__ import annotations

import ctypes
import os
import threading
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_DIR = os.path.join(os.path.dirname(_HERE), "lib")
LIB_PATH = os.path.join(LIB_DIR, "libmlt_b200.so")

_lock = threading.Lock()
_lib: Optional[ctypes.CDLL] = None
_load_error: Optional[BaseException] = None


class NativeLibraryError(RuntimeError):
    pass


def _load() -> Optional[ctypes.CDLL]:
    global _lib, _load_error
    with _lock:
        if _lib is not None or _load_error is not None:
            return _lib
        try:
            lib = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
            lib.mlt_last_error.restype = ctypes.c_char_p
            lib.mlt_version.restype = ctypes.c_int
            _lib = lib
        except OSError as e:      # not built (or wrong platform)
            _load_error = e
        return _lib


def lib_or_none() -> Optional[ctypes.CDLL]:
    return _load()


def lib() -> ctypes.CDLL:
    l = _load()
    if l is None:
        raise NativeLibraryError(
            f"native kernel library not loadable from {LIB_PATH}: {_load_error}. "
            "Build it with `python -c 'import __graft_entry__ as g; g.build()'` (or `make -C csrc`).")
    return l


def native_available() -> bool:
    """True when CUDA is present and the kernel library is loaded."""
    if not torch.cuda.is_available():
        return False
    return lib() is not None        # raises loudly on a GPU box without the library


def check(status: int, what: str = "") -> None:
    if status != 0:
        msg = lib().mlt_last_error()
        raise NativeLibraryError(f"{what or 'native call'} failed (status {status}): "
                                 f"{msg.decode() if msg else 'unknown error'}")


def ptr(t: Optional[torch.Tensor]) -> ctypes.c_void_p:
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def stream_handle(stream=None) -> ctypes.c_void_p:
    s = stream if stream is not None else torch.cuda.current_stream()
    return ctypes.c_void_p(s.cuda_stream)


# launch counter: every successful native launch bumps this (bench.py reports it as gpu_launches)
class _Counter:
    def __init__(self):
        self.n = 0
        self.by_name = {}

    def bump(self, name: str, k: int = 1):
        self.n += k
        self.by_name[name] = self.by_name.get(name, 0) + k

    def reset(self):
        self.n = 0
        self.by_name.clear()


launches = _Counter()
