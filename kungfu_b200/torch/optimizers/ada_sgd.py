"""AdaptiveSGD: SMA until `change_step`, then S-SGD (with a one-off broadcast
so every worker starts S-SGD from rank 0's model).
Reference: /root/reference/srcs/python/kungfu/tensorflow/optimizers/ada_sgd.py:12-83.
"""
import torch

from ..ops import collective as tops
from . import core
from .sma_sgd import _SynchronousAveraging
from .sync_sgd import _SynchronousSGD


def _ada_setup(self):
    _SynchronousAveraging._kf_setup(self)
    self._kf_sync = None
    self._kf_global_step = 0


def _ada_step(self, closure=None):
    if self._kf_global_step < self._kf_change_step:
        loss = _SynchronousAveraging.step(self, closure)
    else:
        if self._kf_global_step == self._kf_change_step:
            # no more model averages from here on (zero_grad() used to start one per step)
            self.overlap_average = False
            if self._kf_avg_pending:
                torch.cuda.current_stream().wait_stream(self._kf_side)
                self._kf_avg_pending = False
            tops.broadcast_parameters({n: p.data for n, p in self._kf_named})
            if core.is_fused(self):
                for _, arenas in self.arenas():
                    for a in arenas:
                        a.ensure_master(slots=self._n_slots())
                        a.master.copy_(a.flat_params.float())
        loss = _SynchronousSGD.step(self, closure)
    self._kf_global_step += 1
    return loss


def AdaptiveSGDOptimizer(optimizer, named_parameters=None, change_step=0, alpha=0.1):
    """Synchronous model averaging (SMA, blend factor ``alpha``) until ``change_step``, then S-SGD; the switch
    broadcasts rank 0's model once so that every replica starts the synchronous phase from the same point
    (reference: tensorflow/optimizers/ada_sgd.py:12-83)."""
    if named_parameters is None:
        named = [('p%d' % i, p) for i, p in enumerate(q for g in optimizer.param_groups for q in g['params'])]
    else:
        named = [(n, p) for n, p in named_parameters if p.requires_grad]
    d = {k: v for k, v in _SynchronousAveraging.__dict__.items() if k not in ('__dict__', '__weakref__')}
    d.update({k: v for k, v in _SynchronousSGD.__dict__.items() if k in ('sync_gradients', )})
    mixin = type('_AdaptiveSGD', (object, ), d)
    mixin._kf_setup = _ada_setup
    mixin.step = torch.no_grad()(_ada_step)
    return core.wrap(optimizer, mixin, _kf_named=named, _kf_alpha=float(alpha), _kf_average=True,
                     _kf_overlap=False, _kf_change_step=int(change_step))
