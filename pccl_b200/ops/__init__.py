"""Torch-facing wrappers around the sm_100a kernels in libpccl.so."""
from .kernels import (LocalArena, block_dequantize_reduce, block_quantize, crc32c, local_all_reduce, local_outer_step, reduce_,
                      scale_inv_, simplehash)

__all__ = ["simplehash", "crc32c", "reduce_", "scale_inv_", "block_quantize", "block_dequantize_reduce", "local_all_reduce",
           "local_outer_step", "LocalArena"]
