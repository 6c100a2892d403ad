"""Locate and import the reference's dependency-free modules (the numerical oracle).

Looks in ``baseline/_ref`` (the copy that travels to the GPU box) and then ``/root/reference``.
Nothing is copied into this repository's history; tests that need the oracle skip if neither
location exists.
"""
import importlib.util
import pathlib
import sys

import pytest

ROOT = pathlib.Path(__file__).resolve().parent.parent
_CANDIDATES = [ROOT / "baseline" / "_ref", pathlib.Path("/root/reference")]
_cache = {}


def ref_dir():
    for c in _CANDIDATES:
        if (c / "ngd_optimizer.py").exists():
            return c
    return None


def load(name: str):
    """Import reference module ``name`` (``resnet`` | ``transformer`` | ``ngd_optimizer``)."""
    if name in _cache:
        return _cache[name]
    d = ref_dir()
    if d is None:
        pytest.skip("reference tree not available")
    spec = importlib.util.spec_from_file_location(f"_ref_{name}", d / f"{name}.py")
    mod = importlib.util.module_from_spec(spec)
    sys.modules[spec.name] = mod
    spec.loader.exec_module(mod)
    _cache[name] = mod
    return mod
