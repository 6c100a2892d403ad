"""Build the native sm_100a library in-tree: ``csrc/*.cu`` → ``lib/libfdt_b200.so``.

Plain ``nvcc`` (no torch headers: the library exposes a C ABI declared in
``include/fdt_b200.h`` and is driven through ctypes with raw device pointers and the caller's
CUDA stream), so a full rebuild takes seconds and cross-compiles on a machine without a GPU.
The ``.so`` stays in the tree so it travels to the GPU box and is visible to the driver's
"which native code was loaded" check.
"""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import pathlib
import shutil
import subprocess
import sys

PKG = pathlib.Path(__file__).resolve().parent
CSRC = PKG / "csrc"
LIBDIR = PKG / "lib"
OBJDIR = PKG / "build"
LIB = LIBDIR / "libfdt_b200.so"
GENCODE = "arch=compute_100a,code=sm_100a"


def nvcc_path() -> str:
    for c in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found")


def _flags():
    return ["-gencode", GENCODE, "-lineinfo", "-O3", "-std=c++17",
            "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "-I", str(CSRC)]


def _digest(src: pathlib.Path) -> str:
    h = hashlib.sha1()
    h.update(src.read_bytes())
    for hdr in sorted(CSRC.glob("*.cuh")):
        h.update(hdr.read_bytes())
    h.update(" ".join(_flags()).encode())
    return h.hexdigest()


def _compile(src: pathlib.Path) -> pathlib.Path:
    obj = OBJDIR / (src.stem + ".o")
    stamp = OBJDIR / (src.stem + ".sha1")
    dg = _digest(src)
    if obj.exists() and stamp.exists() and stamp.read_text() == dg:
        return obj
    cmd = [nvcc_path(), *_flags(), "-c", str(src), "-o", str(obj)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
    stamp.write_text(dg)
    return obj


def build_native(verbose: bool = True, force: bool = False) -> pathlib.Path:
    OBJDIR.mkdir(exist_ok=True)
    LIBDIR.mkdir(exist_ok=True)
    if force:
        for f in OBJDIR.glob("*.sha1"):
            f.unlink()
    srcs = sorted(CSRC.glob("*.cu"))
    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(_compile, srcs))
    newest = max(o.stat().st_mtime for o in objs)
    if force or not LIB.exists() or LIB.stat().st_mtime < newest:
        cmd = [nvcc_path(), "-shared", "-gencode", GENCODE, "-o", str(LIB), *map(str, objs)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    if verbose:
        print(f"[fdt.build] {LIB} ({LIB.stat().st_size // 1024} KiB, {len(srcs)} sources)")
    return LIB


if __name__ == "__main__":
    build_native(force="--force" in sys.argv)
