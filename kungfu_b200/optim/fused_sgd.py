"""Kept for import compatibility: the fused optimizers live in ``kungfu_b200.optim.fused``."""
from .fused import FusedOptimizer, FusedSGD  # noqa: F401
