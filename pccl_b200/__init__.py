"""B200-native PCCL: fault-tolerant collective communication with a CCoIP control plane and
NVLink data plane.  `import pccl_b200 as pccl` gives the surface of the reference package."""
from ._pccl import (AsyncReduceHandle, Attribute, Communicator, DataType, DeviceType, DistributionHint, MasterNode,
                    PCCLError, QuantizationAlgorithm, QuantizationOptions, ReduceDescriptor, ReduceInfo, ReduceOp,
                    ReduceOpDescriptor, ReduceOperandDescriptor, Result, SharedState, SharedStateSyncInfo,
                    SharedStateSyncStrategy, TensorInfo, get_build_info)
from . import cuda

__version__ = "0.1.0"

__all__ = ["cuda", "Communicator", "MasterNode", "ReduceOp", "Attribute", "SharedStateSyncStrategy", "DataType",
           "DeviceType", "DistributionHint", "QuantizationAlgorithm", "ReduceOperandDescriptor", "QuantizationOptions",
           "ReduceDescriptor", "TensorInfo", "SharedState", "SharedStateSyncInfo", "ReduceInfo", "ReduceOpDescriptor",
           "AsyncReduceHandle", "PCCLError", "Result", "get_build_info"]
