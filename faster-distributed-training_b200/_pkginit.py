"""Package-level constants shared by every sub-package."""
__version__ = "0.1.0"

#: the only architecture the native code is written for
CUDA_ARCH = "sm_100a"
GENCODE = "arch=compute_100a,code=sm_100a"

__all__ = ["CUDA_ARCH", "GENCODE"]
