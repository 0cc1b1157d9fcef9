"""Stateful helpers: step counter and exponential moving average.
Reference: KungfuCounter / KungfuExponentialMovingAverage
(/root/reference/srcs/cpp/src/tensorflow/ops/cpu/state.cpp:6-77,
/root/reference/srcs/python/kungfu/tensorflow/ops/state.py)."""


class Counter(object):
    """Returns init, init+incr, init+2*incr, ... on successive calls."""

    def __init__(self, init=0, incr=1):
        self._next = init
        self._incr = incr

    def __call__(self):
        v = self._next
        self._next += self._incr
        return v


def counter(init=0, incr=1):
    """A callable that returns init, init + incr, init + 2 incr, ... (the reference's Counter op)."""
    return Counter(init, incr)


class _EMA(object):
    def __init__(self, alpha):
        self.alpha = alpha
        self.value = None

    def __call__(self, x):
        x = float(x)
        self.value = x if self.value is None else self.alpha * self.value + (1 - self.alpha) * x
        return self.value


def exponential_moving_average(alpha=0.9):
    """v <- alpha v + (1-alpha) x; the first sample passes through."""
    return _EMA(alpha)
