"""Small helpers shared by examples, benchmarks and policies."""


class ExponentialMovingAverage(object):
    """v <- alpha v + (1 - alpha) x, first sample passes through
    (/root/reference/srcs/cpp/include/kungfu/utils/ema.hpp:19-28,
    /root/reference/srcs/python/kungfu/tensorflow/v1/utils/ema.py)."""

    def __init__(self, alpha):
        self.alpha = alpha
        self.value = None

    def update(self, x):
        self.value = x if self.value is None else self.alpha * self.value + (1 - self.alpha) * x
        return self.value


def step_based_schedule(config, step, default=1, strict=True):
    """Cluster size for `step` from a schedule string "size:steps,size:steps,...".
    Past the end: `default` (strict=False) or the last size is kept... the
    reference's op (/root/reference/srcs/cpp/src/tensorflow/ops/cpu/elastic.cpp:16-81)
    returns the default when the schedule is exhausted and errors when strict."""
    off = 0
    for part in config.split(','):
        part = part.strip()
        if not part:
            continue
        size, n = part.split(':')
        off += int(n)
        if step < off:
            return int(size)
    if strict:
        raise ValueError('step %d is beyond the schedule %r' % (step, config))
    return default
