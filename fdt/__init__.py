"""Import alias for the framework package.

The product package lives in ``faster-distributed-training_b200/`` (the directory
name the project brief asks for).  A hyphen is not importable, so ``import fdt``
re-points its ``__path__`` at that directory; ``fdt.models``, ``fdt.ops`` ... resolve
to the real sub-packages.
"""
import pathlib as _pl

_REAL = _pl.Path(__file__).resolve().parent.parent / "faster-distributed-training_b200"
__path__ = [str(_REAL)]  # sub-module lookups go to the real package directory
PACKAGE_DIR = _REAL

from fdt._pkginit import *  # noqa: E402,F401,F403
from fdt._pkginit import __version__  # noqa: E402,F401
