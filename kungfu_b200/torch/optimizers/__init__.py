"""Distributed optimizer wrappers.

The reference's torch binding only had SynchronousSGDOptimizer
(/root/reference/srcs/python/kungfu/torch/optimizers/sync_sgd.py:6-32); SMA,
PairAveraging and the monitors existed for TensorFlow only
(/root/reference/srcs/python/kungfu/tensorflow/optimizers/). All of them are
provided here for torch with the TF semantics, and each has a fused single-kernel
fast path when the wrapped optimizer is ``kungfu_b200.optim.FusedSGD``.
"""
from .sync_sgd import SynchronousSGDOptimizer  # noqa: F401
from .sma_sgd import SynchronousAveragingOptimizer  # noqa: F401
from .async_sgd import PairAveragingOptimizer  # noqa: F401
from .grad_noise_scale import MonitorGradientNoiseScaleOptimizer  # noqa: F401
from .grad_variance import MonitorGradientVarianceOptimizer  # noqa: F401
from .ada_sgd import AdaptiveSGDOptimizer  # noqa: F401
