"""Synchronous SGD (S-SGD): every step all-reduces the gradients.

Reference: /root/reference/srcs/python/kungfu/torch/optimizers/sync_sgd.py:6-32
(torch stub: sum, no division) and
/root/reference/srcs/python/kungfu/tensorflow/optimizers/sync_sgd.py:78-109
(TF: sum then /np). Default here is the TF / documented behaviour (average);
``op='sum'`` reproduces the torch stub.
"""
import torch

from . import core


class _SynchronousSGD(object):
    def _kf_setup(self):
        if core.is_fused(self):
            self._kf_sync = None
            if self._kf_overlap:
                # bucketed fused steps launched from autograd hooks while backward runs
                self.enable_overlap(sync=True, average=self._kf_average)
            self.rebind_overlap()
        else:
            self._kf_sync = core.GradSync(self._kf_named, overlap=self._kf_overlap, average=self._kf_average)

    def sync_gradients(self):
        """All-reduce (average) every gradient in place; no parameter update."""
        if self._kf_sync is None:
            if core.is_fused(self) and self._ov is not None:
                raise RuntimeError('sync_gradients() is not available while the fused step overlaps backward '
                                   '(the exchange and the update are one kernel); build the optimizer with '
                                   'overlap=False')
            self._kf_sync = core.GradSync(self._kf_named, overlap=False, average=self._kf_average)
        self._kf_sync.sync(average=self._kf_average)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if core.is_fused(self) and self._kf_sync is None:
            # one kernel: reduce-scatter + scale + SGD on the owned shard +
            # all-gather of the new parameters over peer memory
            self.sync_step(average=self._kf_average)
            return loss
        self.sync_gradients()
        super(self.__class__, self).step()
        return loss


def SynchronousSGDOptimizer(optimizer, named_parameters=None, op=None, overlap=False):
    """Wrap `optimizer` so that step() first synchronises gradients.

    op: None|'mean'|'average' -> average (default), 'sum' -> plain sum.
    overlap: launch the buckets from autograd hooks while backward runs: for
             the fused optimizers each bucket is one gather + one fused
             exchange-and-update kernel, for other optimizers one gather + one
             all-reduce (one backward pass per step() is assumed).
    """
    if named_parameters is None:
        named = [('p%d' % i, p) for i, p in enumerate(q for g in optimizer.param_groups for q in g['params'])]
    else:
        named = list(named_parameters)
    average = op in (None, 'mean', 'average', 'avg')
    if not average and op != 'sum':
        raise ValueError("op must be None, 'mean' or 'sum'")
    return core.wrap(optimizer, _SynchronousSGD, _kf_named=named, _kf_average=average, _kf_overlap=overlap)
