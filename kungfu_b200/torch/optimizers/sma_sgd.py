"""Synchronous model averaging (SMA / EA-SGD).

Each step: avg = mean over workers of the model, v <- (1-alpha) v + alpha avg,
then the LOCAL gradients are applied. Reference:
/root/reference/srcs/python/kungfu/tensorflow/optimizers/sma_sgd.py:50-74.
With FusedSGD the model all-reduce, the blend and the SGD update are ONE kernel
over peer memory (csrc/kernels/optim.cu:k_sma_fused).
"""
import ctypes as C

import torch

from ... import python as kfpy
from ..ops import collective as tops
from . import core


class _SynchronousAveraging(object):
    def _kf_setup(self):
        self._kf_stage = {}

    def _fused_step(self):
        from ...ops import _cuda as K
        comm = self.comm
        blocks = self._blocks or comm.default_blocks
        threads = self._threads or comm.default_threads
        for g, arenas in self.arenas():
            h = self._hyper(g)
            for a in arenas:
                a.ensure_master()
                a.fuse_grads()
                stage = self._kf_stage.get(id(a))
                if stage is None and comm.np > 1:
                    stage = comm.alloc(a.count * 4)
                    self._kf_stage[id(a)] = stage
                pa, _k1 = a.params_buf.ptr_array()
                sa, _k2 = stage.ptr_array() if stage is not None else (None, None)
                K.call('kfb_sma_fused_step', comm.cstruct, pa, a.code, sa,
                       a.grads_buf.ptrs[max(comm.rank, 0)], a.code, a.master.data_ptr(),
                       a.mom.data_ptr(), a.count, float(self._kf_alpha), 1.0, C.byref(h),
                       blocks, threads, K.stream_ptr())
        self._steps += 1

    def _generic_average(self):
        np_ = kfpy.current_cluster_size()
        if np_ <= 1:
            return
        for n, p in self._kf_named:
            avg = tops.all_reduce_fn(p.data, 'sum', name='sma::' + n)
            p.data.mul_(1 - self._kf_alpha).add_(avg, alpha=self._kf_alpha / np_)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if core.is_fused(self):
            self._fused_step()
            return loss
        self._generic_average()
        super(self.__class__, self).step()
        return loss


def SynchronousAveragingOptimizer(optimizer, named_parameters=None, alpha=0.1):
    """Synchronous model averaging (EA-SGD): ``v <- (1 - alpha) v + alpha * mean_over_workers(v)``, then the LOCAL
    gradients are applied (reference: optimizers/sma_sgd.py:50-74). With FusedSGD: average + blend + update in one kernel."""
    if named_parameters is None:
        named = [('p%d' % i, p) for i, p in enumerate(q for g in optimizer.param_groups for q in g['params'])]
    else:
        named = [(n, p) for n, p in named_parameters if p.requires_grad]
    return core.wrap(optimizer, _SynchronousAveraging, _kf_named=named, _kf_alpha=float(alpha))
