"""`pccl.cuda.is_available()` (reference: python/framework/pccl/_cuda.py)."""
from ._pccl import get_build_info

# the reference keeps the answer in a module attribute of this name (set when the package is imported)
is_cuda_available = bool(get_build_info()["has_cuda_support"])


def is_available() -> bool:
    return is_cuda_available
