from .numa import bind_to_device, device_numa_node
from .profiler import Profiler

__all__ = ["Profiler", "bind_to_device", "device_numa_node"]
