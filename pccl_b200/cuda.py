"""`pccl.cuda.is_available()` (reference: python/framework/pccl/_cuda.py)."""
from ._pccl import get_build_info


def is_available() -> bool:
    return get_build_info()["has_cuda_support"]
