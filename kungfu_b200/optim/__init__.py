from .fused import FusedAdam, FusedAdamW, FusedOptimizer, FusedSGD  # noqa: F401
