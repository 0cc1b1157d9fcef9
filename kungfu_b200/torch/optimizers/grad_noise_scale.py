"""S-SGD + gradient-noise-scale monitor.

Reference: /root/reference/srcs/python/kungfu/tensorflow/optimizers/grad_noise_scale.py:11-88,
/root/reference/srcs/python/kungfu/tensorflow/ops/monitor.py:6-17 and the EMA op
/root/reference/srcs/cpp/src/tensorflow/ops/cpu/collective.cpp:223-258.
  G = (B_big |g_avg|^2 - B_small |g_local|^2) / (B_big - B_small)
  S = (|g_local|^2 - |g_avg|^2) / (1/B_small - 1/B_big)
  noise scale = EMA(S) / EMA(G)
The two norms cost nothing extra here: |g_local|^2 falls out of the bucket
gather kernel, |g_avg|^2 out of the all-reduce (or fused step) kernel.
"""
import torch

from ... import python as kfpy
from . import core
from .sync_sgd import _SynchronousSGD


class _GradientNoiseScale(_SynchronousSGD):
    pass


def _gns_setup(self):
    _SynchronousSGD._kf_setup(self)
    self._kf_gns_state = None
    self._kf_gns_step = 0
    self.noise_scale = None
    if core.is_fused(self):
        self.monitor_norms = True
    else:
        self._kf_sync.monitor_norms = True


def _gns_step(self, closure=None):
    from ...ops import _cuda as K
    loss = _SynchronousSGD.step(self, closure)
    src = self if core.is_fused(self) else self._kf_sync
    if src.sumsq_local is None or not src.sumsq_local.is_cuda:
        return loss
    if self._kf_gns_step % self._kf_interval == 0:
        if self._kf_gns_state is None:
            self._kf_gns_state = torch.zeros(4, dtype=torch.float32, device=src.sumsq_local.device)
        np_ = kfpy.current_cluster_size()
        b_small = float(self._kf_device_batch)
        b_big = b_small * max(np_, 1)
        if np_ > 1:
            K.call('kfb_gns_finalize', src.sumsq_local.data_ptr(), src.sumsq_global.data_ptr(),
                   b_small, b_big, float(self._kf_gns_alpha), self._kf_gns_state.data_ptr(), K.stream_ptr())
            self.noise_scale = self._kf_gns_state[3]
    self._kf_gns_step += 1
    return loss


def MonitorGradientNoiseScaleOptimizer(optimizer, named_parameters=None, device_batch_size=1,
                                       monitor_interval=1, alpha=0.9, op=None, overlap=False):
    """S-SGD whose ``noise_scale`` attribute (a device scalar) tracks the
    gradient noise scale; read it with ``float(opt.noise_scale)``. ``overlap``
    as in SynchronousSGDOptimizer: the two norms then come out of the
    per-bucket kernels (fixed slots, summed in order)."""
    if named_parameters is None:
        named = [('p%d' % i, p) for i, p in enumerate(q for g in optimizer.param_groups for q in g['params'])]
    else:
        named = list(named_parameters)
    mixin = type('_GNS', (object, ), {k: v for k, v in _SynchronousSGD.__dict__.items() if k not in ('__dict__', '__weakref__')})
    mixin._kf_setup = _gns_setup
    mixin.step = torch.no_grad()(_gns_step)
    return core.wrap(optimizer, mixin, _kf_named=named, _kf_average=True, _kf_overlap=bool(overlap),
                     _kf_device_batch=device_batch_size, _kf_interval=max(1, int(monitor_interval)),
                     _kf_gns_alpha=float(alpha))
