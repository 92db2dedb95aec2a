from .numa import bind_to_device, device_numa_node
from .profiler import Profiler, ProfilerCollection

__all__ = ["Profiler", "ProfilerCollection", "bind_to_device", "device_numa_node"]
