"""Python surface of the B200-native PCCL.

Class and method names follow the reference package so that scripts written for it run unchanged
after `import pccl_b200 as pccl` (reference: python/framework/pccl/_pccl.py:42-834).  Differences:

* tensors may live on a CUDA device; peers of one NVLink domain then reduce on the GPUs
  (the reference asserts `device.type == 'cpu'`, _pccl.py:617);
* extra quantisation algorithms (`BLOCK_MIN_MAX`, `BLOCK_FP8`) and a few `ext_*` helpers.
"""
from __future__ import annotations

import ctypes as C
import sys
import time
from enum import Enum
from ipaddress import IPv4Address, IPv6Address, ip_address
from typing import Dict, List, Optional, Tuple, Union

from . import _lib
from ._lib import lib

try:  # both are optional, exactly like in the reference package
    import torch
except ImportError:  # pragma: no cover
    torch = None
try:
    import numpy as np
except ImportError:  # pragma: no cover
    np = None


class Result(Enum):
    SUCCESS = 0
    NOT_INITIALIZED = 1
    INTERNAL_ERROR = 2
    INVALID_ARGUMENT = 3
    INVALID_USAGE = 4
    TOO_FEW_PEERS = 5
    MASTER_CONNECTION_FAILED = 6
    RANK_CONNECTION_FAILED = 7
    RANK_CONNECTION_LOST = 8
    NO_SHARED_STATE_AVAILABLE = 9
    PENDING_ASYNC_OPS = 10
    UPDATE_TOPOLOGY_FAILED = 11
    TOPOLOGY_OPTIMIZATION_FAILED = 12


class PCCLError(Exception):
    def __init__(self, result: Result, func_name: str):
        self.result = result
        self.func_name = func_name
        super().__init__(f"{func_name} failed with {result.name}")

    @staticmethod
    def check(result: int, func_name: str) -> None:
        if result != 0:
            raise PCCLError(Result(result), func_name)


class ReduceOp(Enum):
    SUM = 0
    AVG = 1
    PROD = 2
    MAX = 3
    MIN = 4


class Attribute(Enum):
    GLOBAL_WORLD_SIZE = 1
    PEER_GROUP_WORLD_SIZE = 2
    LOCAL_WORLD_SIZE = 2  # the name the reference's Python package uses for the same attribute (_pccl.py:106)
    NUM_DISTINCT_PEER_GROUPS = 3
    LARGEST_PEER_GROUP_WORLD_SIZE = 4


class SharedStateSyncStrategy(Enum):
    ENFORCE_POPULAR = 0
    RECEIVE_ONLY = 1
    SEND_ONLY = 2


class DataType(Enum):
    UINT8 = 0
    INT8 = 1
    INT16 = 2
    UINT16 = 3
    UINT32 = 4
    INT32 = 5
    UINT64 = 6
    INT64 = 7
    FLOAT16 = 8
    BFLOAT16 = 9
    FLOAT = 10
    DOUBLE = 11
    FLOAT8_E4M3 = 12  # extension: only as quantized_datatype of BLOCK_FP8

    def to_torch_dtype(self):
        return _torch_table()[self]

    def to_numpy_dtype(self):
        return _numpy_table()[self]

    @classmethod
    def from_torch_dtype(cls, dtype) -> "DataType":
        try:
            return _torch_reverse()[dtype]
        except KeyError:
            raise ValueError(f"unsupported torch dtype {dtype}") from None

    @classmethod
    def from_numpy_dtype(cls, dtype) -> "DataType":
        dtype = np.dtype(dtype)
        for k, v in _numpy_table().items():
            if np.dtype(v) == dtype:
                return k
        raise ValueError(f"unsupported numpy dtype {dtype}")

    @property
    def itemsize(self) -> int:
        return int(lib().pcclExtDataTypeSize(self.value))


_TORCH_TABLE = None
_TORCH_REVERSE = None


def _torch_reverse():
    global _TORCH_REVERSE
    if _TORCH_REVERSE is None:
        _TORCH_REVERSE = {v: k for k, v in _torch_table().items()}
    return _TORCH_REVERSE


def _torch_table() -> Dict[DataType, "torch.dtype"]:
    global _TORCH_TABLE
    if _TORCH_TABLE is not None:
        return _TORCH_TABLE
    t = {
        DataType.UINT8: torch.uint8, DataType.INT8: torch.int8, DataType.INT16: torch.int16,
        DataType.INT32: torch.int32, DataType.INT64: torch.int64, DataType.FLOAT16: torch.float16,
        DataType.BFLOAT16: torch.bfloat16, DataType.FLOAT: torch.float32, DataType.DOUBLE: torch.float64,
    }
    for name, dt in (("uint16", DataType.UINT16), ("uint32", DataType.UINT32), ("uint64", DataType.UINT64)):
        if hasattr(torch, name):
            t[dt] = getattr(torch, name)
    _TORCH_TABLE = t
    return t


def _numpy_table() -> Dict[DataType, type]:
    return {
        DataType.UINT8: np.uint8, DataType.INT8: np.int8, DataType.INT16: np.int16, DataType.UINT16: np.uint16,
        DataType.UINT32: np.uint32, DataType.INT32: np.int32, DataType.UINT64: np.uint64, DataType.INT64: np.int64,
        DataType.FLOAT16: np.float16, DataType.FLOAT: np.float32, DataType.DOUBLE: np.float64,
    }


class DeviceType(Enum):
    CPU = 0
    CUDA = 1

    @classmethod
    def from_torch_device_type(cls, device_type: str) -> "DeviceType":
        if device_type == "cpu":
            return cls.CPU
        if device_type == "cuda":
            return cls.CUDA
        raise ValueError(f"unsupported device type {device_type}")


class DistributionHint(Enum):
    NONE = 0
    NORMAL = 1
    UNIFORM = 2


class QuantizationAlgorithm(Enum):
    NONE = 0
    MIN_MAX = 1
    ZERO_POINT_SCALE = 2
    BLOCK_MIN_MAX = 3  # extension
    BLOCK_FP8 = 4      # extension


class ReduceOperandDescriptor:
    def __init__(self, datatype: DataType, distribution_hint: DistributionHint = DistributionHint.NONE):
        self.datatype = datatype
        self.distribution_hint = distribution_hint


class QuantizationOptions:
    def __init__(self, quantized_datatype: DataType = DataType.UINT8,
                 algorithm: QuantizationAlgorithm = QuantizationAlgorithm.NONE):
        self.quantized_datatype = quantized_datatype
        self.algorithm = algorithm


class ReduceDescriptor:
    def __init__(self, count: int, op: ReduceOp, tag: int, operand_descriptor: ReduceOperandDescriptor,
                 quantization_options: Optional[QuantizationOptions] = None):
        self.count = count
        self.op = op
        self.tag = tag
        self.operand_descriptor = operand_descriptor
        self.quantization_options = quantization_options or QuantizationOptions()

    def to_c(self) -> _lib.ReduceDescriptorC:
        d = _lib.ReduceDescriptorC()
        d.count = self.count
        d.op = self.op.value
        d.tag = self.tag
        d.src_descriptor.datatype = self.operand_descriptor.datatype.value
        d.src_descriptor.distribution_hint = self.operand_descriptor.distribution_hint.value
        d.quantization_options.quantized_datatype = self.quantization_options.quantized_datatype.value
        d.quantization_options.algorithm = self.quantization_options.algorithm.value
        return d


def _tensor_view(t) -> Tuple[int, int, DataType, DeviceType, object]:
    """(data pointer, element count, dtype, device type, keep-alive object)"""
    if torch is not None and isinstance(t, torch.Tensor):
        if type(t).__name__ == "DTensor":
            raise ValueError("DTensor is not supported; pass the local shard")
        if not t.is_contiguous():
            raise ValueError("tensor must be contiguous")
        return t.data_ptr(), t.numel(), DataType.from_torch_dtype(t.dtype), DeviceType.from_torch_device_type(t.device.type), t
    if np is not None and isinstance(t, np.ndarray):
        if not t.flags["C_CONTIGUOUS"]:
            raise ValueError("array must be C-contiguous")
        return t.ctypes.data, t.size, DataType.from_numpy_dtype(t.dtype), DeviceType.CPU, t
    raise ValueError(f"unsupported buffer type {type(t)}; expected torch.Tensor or numpy.ndarray")


class ReduceOpDescriptor:
    def __init__(self, sendbuf_ptr: int, recvbuf_ptr: int, reduce_descriptor: ReduceDescriptor, _keep=None):
        self.sendbuf_ptr = sendbuf_ptr
        self.recvbuf_ptr = recvbuf_ptr
        self.reduce_descriptor = reduce_descriptor
        self._keep = _keep

    @staticmethod
    def from_torch(send, recv, reduce_descriptor: ReduceDescriptor) -> "ReduceOpDescriptor":
        sp, sn, sdt, _, _ = _tensor_view(send)
        rp, rn, rdt, _, _ = _tensor_view(recv)
        assert sn == rn and sdt == rdt, "send and recv must match in size and dtype"
        return ReduceOpDescriptor(sp, rp, reduce_descriptor, (send, recv))

    from_numpy = from_torch

    def to_c(self) -> _lib.ReduceOpDescriptorC:
        c_desc = _lib.ReduceOpDescriptorC()
        self.to_c_inpl(c_desc)
        return c_desc

    def to_c_inpl(self, c_desc: _lib.ReduceOpDescriptorC) -> None:
        c_desc.sendbuf = self.sendbuf_ptr
        c_desc.recvbuf = self.recvbuf_ptr
        c_desc.descriptor = self.reduce_descriptor.to_c()


class TensorInfo:
    def __init__(self, name: str, data_ptr: int, *, numel: int, dtype: DataType, device_type: DeviceType,
                 allow_content_inequality: bool = False, _keep=None):
        self.name = name
        self.data_ptr = data_ptr
        self.numel = numel
        self.dtype = dtype
        self.device_type = device_type
        self.allow_content_inequality = allow_content_inequality
        self._keep = _keep

    @classmethod
    def from_torch(cls, tensor, name: str, *, allow_content_inequality: bool = False) -> "TensorInfo":
        ptr, n, dt, dev, keep = _tensor_view(tensor)
        return cls(name, ptr, numel=n, dtype=dt, device_type=dev, allow_content_inequality=allow_content_inequality, _keep=keep)

    from_numpy = from_torch


class SharedState:
    def __init__(self, tensor_infos: List[TensorInfo]):
        self._infos = list(tensor_infos)
        self._names = [ti.name.encode("utf-8") for ti in self._infos]  # kept alive for the C side
        self._array = (_lib.TensorInfoC * len(self._infos))()
        for i, ti in enumerate(self._infos):
            c = self._array[i]
            c.name = self._names[i]
            c.data = ti.data_ptr
            c.count = ti.numel
            c.datatype = ti.dtype.value
            c.device_type = ti.device_type.value
            c.allow_content_inequality = ti.allow_content_inequality
        self._state = _lib.SharedStateC()
        self._state.revision = 0
        self._state.count = len(self._infos)
        self._state.infos = self._array

    @property
    def revision(self) -> int:
        return int(self._state.revision)

    @revision.setter
    def revision(self, value: int) -> None:
        self._state.revision = value

    def push_revision(self) -> None:
        self._state.revision += 1


class SharedStateSyncInfo:
    def __init__(self, tx_bytes: int, rx_bytes: int):
        self.tx_bytes = tx_bytes
        self.rx_bytes = rx_bytes


class ReduceInfo:
    def __init__(self, local_world_size: int, tx_bytes: int, rx_bytes: int):
        self.world_size = local_world_size
        self.local_world_size = local_world_size
        self.tx_bytes = tx_bytes
        self.rx_bytes = rx_bytes


class AsyncReduceHandle:
    def __init__(self, handle: _lib.AsyncReduceOpC, keep=None):
        self._handle = handle
        self._keep = keep
        self._result: Optional[Tuple[bool, int, ReduceInfo]] = None

    def wait(self) -> Tuple[bool, int, ReduceInfo]:
        """(success, status code, info); may be called more than once."""
        if self._result is None:
            info = _lib.ReduceInfoC()
            status = lib().pcclAwaitAsyncReduce(C.byref(self._handle), C.byref(info))
            self._result = (status == 0, status, ReduceInfo(info.local_world_size, info.tx_bytes, info.rx_bytes))
            self._keep = None
        return self._result


def _socket_address(addr: Union[IPv4Address, IPv6Address], port: int) -> _lib.SocketAddress:
    sa = _lib.SocketAddress()
    if isinstance(addr, IPv4Address):
        sa.inet.protocol = 0
        for i, b in enumerate(addr.packed):
            sa.inet.ipv4.data[i] = b
    else:
        sa.inet.protocol = 1
        for i, b in enumerate(addr.packed):
            sa.inet.ipv6.data[i] = b
    sa.port = port
    return sa


def _parse_host_port(text: str) -> Tuple[Union[IPv4Address, IPv6Address], int]:
    assert ":" in text, f"Invalid address: {text}, expected format: ip:port"
    host, port = text.rsplit(":", 1)
    return ip_address(host.strip("[]")), int(port)


class Communicator:
    """One peer of a run."""

    def __init__(self, address: str, peer_group: int = 0, p2p_connection_pool_size: int = 0,
                 public_advertise_ip: Optional[str] = None, p2p_listen_port: int = 48149,
                 shared_state_listen_port: int = 48150, benchmark_listen_port: int = 48151):
        host, port = _parse_host_port(address)
        params = _lib.CommCreateParams()
        params.master_address = _socket_address(host, port)
        params.peer_group = peer_group
        params.p2p_connection_pool_size = p2p_connection_pool_size
        params.internal_p2p_listen_port = p2p_listen_port
        params.internal_shared_state_listen_port = shared_state_listen_port
        params.internal_benchmark_listen_port = benchmark_listen_port
        if public_advertise_ip is not None:
            pub = ip_address(public_advertise_ip)
            params.use_explicit_p2p_addresses = True
            params.advertised_p2p_address = _socket_address(pub, p2p_listen_port)
            params.advertised_shared_state_address = _socket_address(pub, shared_state_listen_port)
            params.advertised_benchmark_address = _socket_address(pub, benchmark_listen_port)
        self._comm = C.c_void_p()
        PCCLError.check(lib().pcclCreateCommunicator(C.byref(params), C.byref(self._comm)), "pcclCreateCommunicator")

    def __del__(self):
        self.destroy()

    def destroy(self) -> None:
        comm, self._comm = getattr(self, "_comm", None), None
        if comm:
            try:
                lib().pcclDestroyCommunicator(comm)
            except Exception:  # interpreter shutdown: module globals may already be gone
                pass

    def get_attribute(self, attribute: Attribute) -> int:
        out = C.c_int()
        PCCLError.check(lib().pcclGetAttribute(self._comm, attribute.value, C.byref(out)), "pcclGetAttribute")
        return out.value

    def connect(self, n_attempts: int = 5) -> None:
        """Joins the run; retries once per second while the master is unreachable."""
        assert n_attempts > 0
        for attempt in range(n_attempts):
            status = lib().pcclConnect(self._comm)
            if status == 0:
                return
            if attempt + 1 == n_attempts:
                PCCLError.check(status, "pcclConnect")
            time.sleep(1)

    def update_topology(self) -> None:
        PCCLError.check(lib().pcclUpdateTopology(self._comm), "pcclUpdateTopology")

    def are_peers_pending(self) -> bool:
        out = C.c_bool()
        PCCLError.check(lib().pcclArePeersPending(self._comm, C.byref(out)), "pcclArePeersPending")
        return bool(out.value)

    def optimize_topology(self) -> None:
        PCCLError.check(lib().pcclOptimizeTopology(self._comm), "pcclOptimizeTopology")

    def sync_shared_state(self, shared_state: SharedState,
                          strategy: SharedStateSyncStrategy = SharedStateSyncStrategy.ENFORCE_POPULAR) -> SharedStateSyncInfo:
        _sync_current_stream(shared_state._infos)
        info = _lib.SharedStateSyncInfoC()
        PCCLError.check(lib().pcclSynchronizeSharedState(self._comm, C.byref(shared_state._state), strategy.value, C.byref(info)),
                        "pcclSynchronizeSharedState")
        return SharedStateSyncInfo(info.tx_bytes, info.rx_bytes)

    def _descriptor(self, send, recv, op: ReduceOp, tag: int, operand_descriptor, quantization_options, sync_stream: bool = True):
        sp, sn, sdt, sdev, _ = _tensor_view(send)
        rp, rn, rdt, rdev, _ = _tensor_view(recv)
        assert sn == rn, "send and recv must have the same number of elements"
        assert sdt == rdt, "send and recv must have the same dtype"
        assert sdev == rdev, "send and recv must live on the same kind of device"
        operand_descriptor = operand_descriptor or ReduceOperandDescriptor(sdt)
        desc = ReduceDescriptor(sn, op, tag, operand_descriptor, quantization_options)
        if sdev == DeviceType.CUDA and sync_stream:
            _sync_current_stream([send, recv])
        return sp, rp, desc.to_c()

    def all_reduce(self, send, recv, *, op: ReduceOp, tag: int = 0,
                   operand_descriptor: Optional[ReduceOperandDescriptor] = None,
                   quantization_options: Optional[QuantizationOptions] = None,
                   stream=None, enqueue: bool = False) -> ReduceInfo:
        """Reduces `send` over all peers of the group into `recv`.

        CUDA tensors are reduced in stream order: the operation starts after the work queued on `stream` (default: torch's
        current stream of the tensor's device) and that stream waits for the result -- no stream or device synchronisation
        happens here, and kernels on other streams overlap with the reduction.  By default the call still returns only
        when the reduction has finished (the reference's contract).  enqueue=True returns after the launch whenever the
        operation needs no vote at the master (every operation on a tag after the first one, until membership changes);
        a failure then surfaces at the next stream-ordered call or at ext_wait_stream_ops()."""
        on_device = torch is not None and isinstance(send, torch.Tensor) and send.is_cuda
        if on_device and operand_descriptor is None and type(recv) is type(send):
            # hot path (small stream-ordered reductions are host-bound): no intermediate descriptor objects
            if not (send.is_contiguous() and recv.is_contiguous()):
                raise ValueError("tensor must be contiguous")
            assert send.numel() == recv.numel() and send.dtype == recv.dtype and recv.is_cuda, "send and recv must match"
            q = quantization_options
            key = (send.dtype, op, tag, None if q is None else (q.quantized_datatype, q.algorithm))
            cache = self.__dict__.setdefault("_desc_cache", {})
            desc = cache.get(key)
            if desc is None:
                desc = ReduceDescriptor(0, op, tag, ReduceOperandDescriptor(DataType.from_torch_dtype(send.dtype)), q).to_c()
                cache[key] = desc
            desc.count = send.numel()
            sp, rp = send.data_ptr(), recv.data_ptr()
        else:
            sp, rp, desc = self._descriptor(send, recv, op, tag, operand_descriptor, quantization_options, sync_stream=not on_device)
        info = _lib.ReduceInfoC()
        if on_device:
            handle = (stream if stream is not None else torch.cuda.current_stream(send.device)).cuda_stream
            PCCLError.check(lib().pcclExtAllReduceOnStream(sp, rp, C.byref(desc), self._comm, C.c_void_p(handle), 1 if enqueue else 0, C.byref(info)),
                            "pcclExtAllReduceOnStream")
        else:
            PCCLError.check(lib().pcclAllReduce(sp, rp, C.byref(desc), self._comm, C.byref(info)), "pcclAllReduce")
        return ReduceInfo(info.local_world_size, info.tx_bytes, info.rx_bytes)

    def ext_wait_stream_ops(self, tag: Optional[int] = None) -> ReduceInfo:
        """Takes the verdict of the enqueued (all_reduce(..., enqueue=True)) operations on `tag`, or of all of them.
        Raises PCCLError(RANK_CONNECTION_LOST) when one of them failed."""
        info = _lib.ReduceInfoC()
        PCCLError.check(lib().pcclExtWaitStreamOps(self._comm, 0 if tag is None else tag, 1 if tag is None else 0, C.byref(info)),
                        "pcclExtWaitStreamOps")
        return ReduceInfo(info.local_world_size, info.tx_bytes, info.rx_bytes)

    def all_reduce_async(self, send, recv, *, op: ReduceOp, tag: int = 0,
                         operand_descriptor: Optional[ReduceOperandDescriptor] = None,
                         quantization_options: Optional[QuantizationOptions] = None) -> AsyncReduceHandle:
        sp, rp, desc = self._descriptor(send, recv, op, tag, operand_descriptor, quantization_options)
        handle = _lib.AsyncReduceOpC()
        PCCLError.check(lib().pcclAllReduceAsync(sp, rp, C.byref(desc), self._comm, C.byref(handle)), "pcclAllReduceAsync")
        return AsyncReduceHandle(handle, keep=(send, recv))

    def all_reduce_multiple_with_retry(self, descriptors: List[ReduceOpDescriptor], *, max_in_flight: int = 4) -> ReduceInfo:
        arr = (_lib.ReduceOpDescriptorC * len(descriptors))()
        for i, d in enumerate(descriptors):
            d.to_c_inpl(arr[i])
        _sync_current_stream([d._keep for d in descriptors])
        info = _lib.ReduceInfoC()
        PCCLError.check(lib().pcclAllReduceMultipleWithRetry(arr, len(descriptors), self._comm, C.byref(info), max_in_flight),
                        "pcclAllReduceMultipleWithRetry")
        return ReduceInfo(info.local_world_size, info.tx_bytes, info.rx_bytes)

    # ---- extensions -------------------------------------------------------------------------
    def ext_last_reduce_stats(self, tag: int = 0) -> Dict[str, Union[int, float, str]]:
        st = _lib.ExtReduceStatsC()
        PCCLError.check(lib().pcclExtGetLastReduceStats(self._comm, tag, C.byref(st)), "pcclExtGetLastReduceStats")
        names = {0: "none", 1: "tcp_ring", 2: "p2p_one_shot", 3: "p2p_two_shot", 4: "nvls", 5: "tcp_hierarchical"}
        return {"plane": names.get(st.plane, str(st.plane)), "device_ms": float(st.device_ms), "seq_nr": int(st.seq_nr),
                "kernel_launches": int(st.kernel_launches), "leased": bool(st.leased),
                "phase_us": [round((int(st.phase_ns[i]) - int(st.phase_ns[0])) / 1e3, 2) if st.phase_ns[i] else None for i in range(7)],
                "content_hash": [int(st.content_hash[0]), int(st.content_hash[1])] if st.has_content_hash else None}

    def ext_trust_content_hashes(self, on: bool = True) -> None:
        """Lets shared-state synchronisation reuse the content hashes the collective kernels produce while they write a
        tensor (hash type mset64: set PCCL_SHARED_STATE_HASH=mset on every peer).  You promise not to modify tensors the
        library wrote last (outer-step parameters and momentum) except through the library."""
        PCCLError.check(lib().pcclExtTrustContentHashes(self._comm, 1 if on else 0), "pcclExtTrustContentHashes")

    def ext_hash_kernels_skipped(self) -> int:
        out = C.c_uint64()
        PCCLError.check(lib().pcclExtHashKernelsSkipped(self._comm, C.byref(out)), "pcclExtHashKernelsSkipped")
        return int(out.value)

    def ext_diloco_outer_step(self, params, inner, momentum, *, lr: float = 0.7, mu: float = 0.9, first: bool = False,
                              tag: int = 0) -> ReduceInfo:
        """DiLoCo outer step as ONE collective: g = mean over peers of (params - inner); momentum = g if `first` else
        mu*momentum + g; params -= lr*(g + mu*momentum); inner = params  (torch.optim.SGD with nesterov=True).
        float32, contiguous, all on the host or all on one GPU.  Between GPUs of one NVLink domain this is a single
        kernel per peer; see pcclExtDiLoCoOuterStep in pccl.h.  Raises like `all_reduce`; tensors are untouched on failure."""
        views = [_tensor_view(t) for t in (params, inner, momentum)]
        n = views[0][1]
        for _, cnt, dt, dev, _ in views:
            assert cnt == n and dt == DataType.FLOAT and dev == views[0][3], "three float32 tensors of equal size on one device"
        if views[0][3] == DeviceType.CUDA:
            _sync_current_stream([params, inner, momentum])
        d = _lib.ExtOuterStepDescriptorC(views[0][0], views[1][0], views[2][0], n, lr, mu, 1 if first else 0, tag)
        info = _lib.ReduceInfoC()
        PCCLError.check(lib().pcclExtDiLoCoOuterStep(C.byref(d), self._comm, C.byref(info)), "pcclExtDiLoCoOuterStep")
        return ReduceInfo(info.local_world_size, info.tx_bytes, info.rx_bytes)

    def ext_symmetric_empty(self, numel: int, dtype) -> "torch.Tensor":
        """A CUDA tensor inside the communicator's peer-visible arena (needs PCCL_SYMMETRIC_MB).
        In-place reductions on such tensors skip the staging copy.  Every peer must allocate
        the same sizes in the same order."""
        dt = DataType.from_torch_dtype(dtype)
        ptr = C.c_void_p()
        nbytes = numel * dt.itemsize
        PCCLError.check(lib().pcclExtSymmetricAlloc(self._comm, nbytes, C.byref(ptr)), "pcclExtSymmetricAlloc")

        class _Holder:
            pass

        holder = _Holder()
        holder.__cuda_array_interface__ = {
            "shape": (numel,), "typestr": _typestr(dt), "data": (ptr.value, False), "version": 3, "strides": None,
        }
        t = torch.as_tensor(holder, device="cuda")
        return t.view(dtype) if t.dtype != dtype else t

    def ext_symmetric_free(self, tensor: "torch.Tensor") -> None:
        """Gives a tensor from ext_symmetric_empty back to the arena (same order on every peer, like the allocation).
        The tensor must not be used afterwards."""
        PCCLError.check(lib().pcclExtSymmetricFree(self._comm, tensor.data_ptr()), "pcclExtSymmetricFree")


def _typestr(dt: DataType) -> str:
    return {
        DataType.UINT8: "|u1", DataType.INT8: "|i1", DataType.INT16: "<i2", DataType.UINT16: "<u2", DataType.UINT32: "<u4",
        DataType.INT32: "<i4", DataType.UINT64: "<u8", DataType.INT64: "<i8", DataType.FLOAT16: "<f2", DataType.BFLOAT16: "<u2",
        DataType.FLOAT: "<f4", DataType.DOUBLE: "<f8",
    }[dt]


def _sync_current_stream(objs) -> None:
    """The library orders its device work after the legacy default stream.  Work queued on any
    other torch stream has to be finished before the buffers are handed over."""
    if torch is None or not torch.cuda.is_available():
        return
    def walk(o):
        if isinstance(o, (list, tuple)):
            for x in o:
                yield from walk(x)
        elif isinstance(o, TensorInfo):
            yield from walk(o._keep)
        elif isinstance(o, torch.Tensor):
            yield o
    for t in walk(objs):
        if t.is_cuda:
            s = torch.cuda.current_stream(t.device)
            if s.cuda_stream != 0:  # the legacy default stream is ordered by the library itself
                s.synchronize()
            return


class MasterNode:
    """A CCoIP master.  `run()` starts it on its own thread and returns; `interrupt()` stops it.

    Like the reference's class (python/framework/pccl/_pccl.py:814-834), dropping the last reference to a master that runs
    and was not interrupted BLOCKS until somebody interrupts it: the reference's master scripts consist of
    `MasterNode(addr).run()` and rely on exactly that to stay alive.  Unlike the reference, the wait is interruptible
    (Ctrl-C ends it and stops the master), and the object is a context manager whose exit interrupts:
    `with MasterNode(addr) as m: m.run(); ...`."""

    def __init__(self, listen_address: str):
        host, port = _parse_host_port(listen_address)
        self._master = C.c_void_p()
        self._running = False
        self._interrupted = False
        PCCLError.check(lib().pcclCreateMaster(_socket_address(host, port), C.byref(self._master)), "pcclCreateMaster")

    def run(self) -> None:
        """Non-blocking: the master runs on its own thread."""
        PCCLError.check(lib().pcclRunMaster(self._master), "pcclRunMaster")
        self._running = True

    @property
    def port(self) -> int:
        out = C.c_uint16()
        PCCLError.check(lib().pcclExtMasterGetPort(self._master, C.byref(out)), "pcclExtMasterGetPort")
        return out.value

    def interrupt(self) -> None:
        PCCLError.check(lib().pcclInterruptMaster(self._master), "pcclInterruptMaster")
        self._interrupted = True

    def __enter__(self) -> "MasterNode":
        return self

    def __exit__(self, *exc) -> None:
        if self._running and not self._interrupted and self._master:
            self.interrupt()

    def __del__(self):
        m, self._master = getattr(self, "_master", None), None
        if not m:
            return
        try:
            if self._running:
                try:
                    while not self._interrupted and not sys.is_finalizing():
                        time.sleep(0.1)  # stays alive until interrupt(); a KeyboardInterrupt lands here
                finally:
                    if not self._interrupted:
                        lib().pcclInterruptMaster(m)
                    lib().pcclMasterAwaitTermination(m)
            lib().pcclDestroyMaster(m)
        except Exception:  # interpreter shutdown
            pass


def get_build_info() -> Dict[str, bool]:
    info = _lib.BuildInfoC()
    PCCLError.check(lib().pcclGetBuildInfo(C.byref(info)), "pcclGetBuildInfo")
    return {"has_cuda_support": bool(info.has_cuda_support)}
