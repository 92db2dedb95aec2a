"""Training-loop level building blocks on top of the communicator: DiLoCo outer steps, DDP-style
gradient averaging."""
from .diloco import DiLoCo, LocalOuterStep

__all__ = ["DiLoCo", "LocalOuterStep"]
