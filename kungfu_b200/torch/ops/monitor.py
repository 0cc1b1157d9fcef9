"""Monitoring ops. Reference: /root/reference/srcs/python/kungfu/tensorflow/ops/monitor.py:6-27."""
import torch

from ... import python as kfpy


class NoiseScale(object):
    """G / S estimators + EMA of each, ratio S_ema / G_ema
    (/root/reference/srcs/cpp/src/tensorflow/ops/cpu/collective.cpp:223-258)."""

    def __init__(self, alpha=0.6):
        self.alpha = alpha
        self.g = self.s = None

    def __call__(self, batch_small, batch_big, tensor, avg_tensor):
        gs = float(torch.as_tensor(tensor).float().pow(2).sum())
        gb = float(torch.as_tensor(avg_tensor).float().pow(2).sum())
        G = (batch_big * gb - batch_small * gs) / (batch_big - batch_small)
        S = (gs - gb) / (1.0 / batch_small - 1.0 / batch_big)
        self.g = G if self.g is None else self.alpha * self.g + (1 - self.alpha) * G
        self.s = S if self.s is None else self.alpha * self.s + (1 - self.alpha) * S
        return self.s / self.g if self.g != 0 else float('inf')


def global_noise_scale(batch_small, batch_big, tensor, avg_tensor, alpha=0.6, _state={}):
    ns = _state.setdefault(alpha, NoiseScale(alpha))
    return ns(batch_small, batch_big, tensor, avg_tensor)


def egress_rates():
    """float32[np] bytes/s sent to each peer: control-plane sockets (with
    KUNGFU_CONFIG_ENABLE_MONITORING=true) plus the NVLink bytes of the device kernels since the previous call
    (reference: /root/reference/srcs/python/kungfu/tensorflow/ops/monitor.py:20-27)."""
    return torch.tensor(kfpy.egress_rates(), dtype=torch.float32)
