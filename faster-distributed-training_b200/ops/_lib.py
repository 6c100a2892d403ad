"""ctypes loader for ``lib/libfdt_b200.so`` (the hand-written sm_100a kernels).

Every entry point is ``extern "C" int fdt_xxx(..., cudaStream_t)``, takes raw device pointers and
returns 0 or a CUDA error code.  ``call`` converts tensors to pointers, appends the current torch
CUDA stream and raises on a non-zero return — so a missing library or a failed launch is loud,
never a silent fallback.  ``LAUNCHES`` counts kernel launches issued through this module (the
``gpu_launches`` figure ``bench.py`` reports).
"""
from __future__ import annotations

import ctypes
import os
import pathlib
from typing import Optional

import torch

PKG = pathlib.Path(__file__).resolve().parent.parent
LIB_PATH = PKG / "lib" / "libfdt_b200.so"

_lib: Optional[ctypes.CDLL] = None


class LaunchCounter:
    def __init__(self):
        self.count = 0
        self.by_name = {}

    def add(self, name: str, n: int = 1):
        self.count += n
        self.by_name[name] = self.by_name.get(name, 0) + n


LAUNCHES = LaunchCounter()


def available() -> bool:
    return LIB_PATH.exists()


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            if os.environ.get("FDT_AUTOBUILD", "1") == "1":
                from ..build import build_native
                build_native(verbose=False)
            else:
                raise RuntimeError(f"native library {LIB_PATH} is missing; run `python -m fdt.build`")
        _lib = ctypes.CDLL(str(LIB_PATH))
    return _lib


def _conv(a):
    if a is None:
        return ctypes.c_void_p(0)
    if torch.is_tensor(a):
        return ctypes.c_void_p(a.data_ptr())
    if isinstance(a, bool):
        return ctypes.c_int(int(a))
    if isinstance(a, int):
        return ctypes.c_int64(a) if abs(a) > 0x7FFFFFFF else ctypes.c_int(a)
    if isinstance(a, float):
        return ctypes.c_float(a)
    return a      # already a ctypes object (c_int64, byref(struct), ...)


def i64(v: int):
    return ctypes.c_int64(int(v))


def stream_ptr() -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def call(name: str, *args, launches: int = 1):
    fn = getattr(lib(), name)
    fn.restype = ctypes.c_int
    rc = fn(*[_conv(a) for a in args], stream_ptr())
    if rc != 0:
        raise RuntimeError(f"{name} failed with code {rc}")
    LAUNCHES.add(name, launches)
