from .profiler import Profiler

__all__ = ["Profiler"]
