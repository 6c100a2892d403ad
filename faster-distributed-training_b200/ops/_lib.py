"""ctypes loader for ``lib/libfdt_b200.so`` (the hand-written sm_100a kernels).

Every entry point is ``extern "C" int fdt_xxx(..., cudaStream_t)``, takes raw device pointers and
returns 0 or a CUDA error code.  The argument types of each entry point are read from the C
declarations in ``include/fdt_b200.h`` (the one ABI description, shared with C callers) and
installed as ctypes ``argtypes``, so a Python ``int`` reaches an ``int64_t`` parameter as 64 bits
and a wrong argument count fails in Python rather than in the kernel.  ``call`` converts tensors
to pointers, appends the current torch CUDA stream and raises on a non-zero return — so a missing
library or a failed launch is loud, never a silent fallback.

``LAUNCHES`` counts kernel launches issued through this module (the ``gpu_launches`` figure
``bench.py`` reports).  ``FALLBACKS`` counts every call that a CUDA tensor took through a
*library* kernel (cuDNN / cuBLAS / ATen) because the native kernels do not cover its shape:
:func:`note_fallback` is called at each such site, the totals are printed at interpreter exit
(``FDT_FALLBACK_REPORT=0`` silences it) and ``FDT_STRICT=1`` turns the first one into an error.
"""
from __future__ import annotations

import atexit
import ctypes
import os
import pathlib
import re
import sys
from typing import Dict, List, Optional

import torch

PKG = pathlib.Path(__file__).resolve().parent.parent
LIB_PATH = PKG / "lib" / "libfdt_b200.so"
HEADER_PATH = PKG.parent / "include" / "fdt_b200.h"

_lib: Optional[ctypes.CDLL] = None


class LaunchCounter:
    def __init__(self):
        self.count = 0
        self.by_name = {}

    def add(self, name: str, n: int = 1):
        self.count += n
        self.by_name[name] = self.by_name.get(name, 0) + n


LAUNCHES = LaunchCounter()
FALLBACKS = LaunchCounter()
STRICT = os.environ.get("FDT_STRICT", "0") == "1"


class NativeCoverageError(RuntimeError):
    """Raised under ``FDT_STRICT=1`` when a CUDA call would run on a library kernel."""


def set_strict(flag: bool) -> None:
    global STRICT
    STRICT = bool(flag)


def note_fallback(op: str, why: str = "") -> None:
    """Record that ``op`` is about to run on a library (cuDNN / cuBLAS / ATen) kernel instead of a native one."""
    key = f"{op}: {why}" if why else op
    if STRICT:
        raise NativeCoverageError(f"FDT_STRICT=1: {key} has no native sm_100a kernel for this call")
    FALLBACKS.add(key)


def _report_fallbacks():
    if FALLBACKS.count and os.environ.get("FDT_FALLBACK_REPORT", "1") == "1":
        rows = sorted(FALLBACKS.by_name.items(), key=lambda kv: -kv[1])
        print(f"[fdt] {FALLBACKS.count} call(s) ran on library kernels instead of the native sm_100a path "
              f"(FDT_STRICT=1 makes this an error):", file=sys.stderr)
        for k, v in rows[:20]:
            print(f"[fdt]   {v:8d} x {k}", file=sys.stderr)


atexit.register(_report_fallbacks)


def available() -> bool:
    return LIB_PATH.exists()


# ----------------------------------------------------------------------------- ABI from the header
_CTYPE = {"int": ctypes.c_int, "int64_t": ctypes.c_int64, "float": ctypes.c_float, "double": ctypes.c_double,
          "cudaStream_t": ctypes.c_void_p, "uint32_t": ctypes.c_uint32, "uint64_t": ctypes.c_uint64}
_SIGS: Optional[Dict[str, List]] = None


def _parse_header() -> Dict[str, List]:
    sigs: Dict[str, List] = {}
    if not HEADER_PATH.exists():
        return sigs
    text = re.sub(r"/\*.*?\*/", " ", HEADER_PATH.read_text(), flags=re.S)
    for m in re.finditer(r"\bint\s+(fdt_\w+)\s*\(([^)]*)\)\s*;", text):
        types = []
        for arg in m.group(2).split(","):
            arg = arg.strip()
            if "*" in arg:
                types.append(ctypes.c_void_p)
                continue
            words = [w for w in arg.split() if w != "const"]
            types.append(_CTYPE[words[0]])
        sigs[m.group(1)] = types
    return sigs


def signatures() -> Dict[str, List]:
    global _SIGS
    if _SIGS is None:
        _SIGS = _parse_header()
    return _SIGS


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
        for name, types in signatures().items():
            fn = getattr(_lib, name, None)
            if fn is not None:
                fn.argtypes = types
                fn.restype = ctypes.c_int
    return _lib


def _conv(a):
    if a is None:
        return None
    if torch.is_tensor(a):
        return ctypes.c_void_p(a.data_ptr())
    if isinstance(a, bool):
        return int(a)
    if isinstance(a, (ctypes.c_int64, ctypes.c_int, ctypes.c_float)):
        return a.value
    return a      # python int / float (converted per argtypes), byref(struct), ctypes arrays, c_void_p


def i64(v: int):
    return ctypes.c_int64(int(v))


def stream_ptr() -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def call(name: str, *args, launches: int = 1):
    fn = getattr(lib(), name)
    if fn.argtypes is None:
        raise RuntimeError(f"{name} is not declared in {HEADER_PATH.name}: every entry point needs its C declaration there")
    if len(args) + 1 != len(fn.argtypes):
        raise TypeError(f"{name} takes {len(fn.argtypes) - 1} arguments + stream, got {len(args)}")
    rc = fn(*[_conv(a) for a in args], stream_ptr())
    if rc != 0:
        raise RuntimeError(f"{name} failed with code {rc}")
    LAUNCHES.add(name, launches)
