"""Training-loop level building blocks on top of the communicator: DiLoCo outer steps, DDP-style
gradient averaging."""
from .ddp import GradientAverager
from .diloco import AsyncDiLoCo, DiLoCo, LocalOuterStep

__all__ = ["DiLoCo", "AsyncDiLoCo", "LocalOuterStep", "GradientAverager"]
