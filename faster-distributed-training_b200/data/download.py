"""Dataset download / integrity / extraction helpers.

The reference vendors 512 lines of ``torchvision.datasets.utils`` (``torchvision_utils.py``) and
uses two functions from it: ``check_integrity`` and ``download_and_extract_archive``
(``resnet50_test.py:36``).  Those two, plus ``verify_str_arg`` and the archive extractors they
need, are provided here on top of the standard library.  The GPU box has no network, so on it
only the integrity / extraction halves are exercised.
"""
from __future__ import annotations

import bz2
import gzip
import hashlib
import lzma
import os
import shutil
import tarfile
import urllib.request
import zipfile
from typing import Iterable, Optional

USER_AGENT = "fdt-b200/0.1"


def calculate_md5(fpath: str, chunk_size: int = 1 << 20) -> str:
    md5 = hashlib.md5()
    with open(fpath, "rb") as f:
        for chunk in iter(lambda: f.read(chunk_size), b""):
            md5.update(chunk)
    return md5.hexdigest()


def check_md5(fpath: str, md5: str) -> bool:
    return md5 == calculate_md5(fpath)


def check_integrity(fpath: str, md5: Optional[str] = None) -> bool:
    if not os.path.isfile(fpath):
        return False
    return True if md5 is None else check_md5(fpath, md5)


def download_url(url: str, root: str, filename: Optional[str] = None, md5: Optional[str] = None,
                 max_redirect_hops: int = 3) -> str:
    root = os.path.expanduser(root)
    filename = filename or os.path.basename(url)
    fpath = os.path.join(root, filename)
    os.makedirs(root, exist_ok=True)
    if check_integrity(fpath, md5):
        return fpath
    req = urllib.request.Request(url, headers={"User-Agent": USER_AGENT})
    try:
        with urllib.request.urlopen(req, timeout=60) as r, open(fpath, "wb") as f:
            shutil.copyfileobj(r, f)
    except Exception as e:
        if url.startswith("https:"):
            with urllib.request.urlopen(urllib.request.Request("http:" + url[6:], headers={"User-Agent": USER_AGENT}),
                                        timeout=60) as r, open(fpath, "wb") as f:
                shutil.copyfileobj(r, f)
        else:
            raise RuntimeError(f"download of {url} failed: {e}") from e
    if not check_integrity(fpath, md5):
        raise RuntimeError("File not found or corrupted.")
    return fpath


_COMPRESSED = {".gz": gzip.open, ".bz2": bz2.open, ".xz": lzma.open}


def extract_archive(from_path: str, to_path: Optional[str] = None, remove_finished: bool = False) -> str:
    to_path = to_path or os.path.dirname(from_path)
    lower = from_path.lower()
    if lower.endswith((".tar", ".tar.gz", ".tgz", ".tar.bz2", ".tbz2", ".tar.xz", ".txz")):
        with tarfile.open(from_path, "r:*") as tar:
            tar.extractall(to_path, filter="data")
    elif lower.endswith(".zip"):
        with zipfile.ZipFile(from_path) as z:
            z.extractall(to_path)
    else:
        ext = os.path.splitext(lower)[1]
        if ext not in _COMPRESSED:
            raise RuntimeError(f"Unknown archive type: {from_path}")
        out = os.path.join(to_path, os.path.basename(from_path)[: -len(ext)])
        with _COMPRESSED[ext](from_path, "rb") as src, open(out, "wb") as dst:
            shutil.copyfileobj(src, dst)
        to_path = out
    if remove_finished:
        os.remove(from_path)
    return to_path


def download_and_extract_archive(url: str, download_root: str, extract_root: Optional[str] = None,
                                 filename: Optional[str] = None, md5: Optional[str] = None,
                                 remove_finished: bool = False) -> None:
    download_root = os.path.expanduser(download_root)
    extract_root = extract_root or download_root
    filename = filename or os.path.basename(url)
    download_url(url, download_root, filename, md5)
    extract_archive(os.path.join(download_root, filename), extract_root, remove_finished)


def verify_str_arg(value, arg: Optional[str] = None, valid_values: Optional[Iterable[str]] = None):
    if not isinstance(value, str):
        raise ValueError(f"Expected type str for argument {arg}, but got type {type(value)}.")
    if valid_values is not None and value not in valid_values:
        raise ValueError(f"Unknown value '{value}' for argument {arg}. Valid values are {set(valid_values)}.")
    return value
