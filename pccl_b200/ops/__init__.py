"""Torch-facing wrappers around the sm_100a kernels in libpccl.so."""
from .kernels import (EmulatedWorld, LocalArena, block_dequantize_reduce, block_quantize, crc32c, local_all_reduce, local_outer_step,
                      mset64, mset64_finalize, reduce_, scale_inv_, simplehash)

__all__ = ["simplehash", "crc32c", "mset64", "mset64_finalize", "reduce_", "scale_inv_", "block_quantize", "block_dequantize_reduce",
           "local_all_reduce", "local_outer_step", "LocalArena", "EmulatedWorld"]
