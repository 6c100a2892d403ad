"""Materialise the unmodified reference under ``baseline/_ref`` (git-ignored, travels with gpurun).

The reference is a flat directory of scripts with no ``setup.py`` / ``pyproject.toml``, so
``pip install --no-index --target baseline/_ref /root/reference`` fails with "not installable"
(outcome recorded in DESIGN.md).  The fallback "install" is a byte-for-byte copy of its tree;
``bench.py --impl reference`` and the oracle tests import ``resnet``, ``transformer`` and
``ngd_optimizer`` from there, unmodified.
"""
import pathlib
import shutil
import sys

SRC = pathlib.Path("/root/reference")
DST = pathlib.Path(__file__).resolve().parent / "_ref"


def main() -> int:
    if not SRC.is_dir():
        print(f"{SRC} not found; nothing installed", file=sys.stderr)
        return 1
    if DST.exists():
        shutil.rmtree(DST)
    shutil.copytree(SRC, DST, ignore=shutil.ignore_patterns("figures", ".git", "__pycache__"))
    print(f"copied {SRC} -> {DST}")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
