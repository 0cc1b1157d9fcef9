"""S-SGD + gradient-variance monitor.

Reference: /root/reference/srcs/python/kungfu/tensorflow/optimizers/grad_variance.py:46-59:
a second all-reduce of g^2, then sum over tensors of || E[g^2] - E[g]^2 ||_2.
Here one kernel reads every rank's gradients once and keeps both accumulators
(csrc/kernels/optim.cu:k_grad_variance); the reduced per-tensor partial sums
are a few hundred floats.
"""
import torch

from ... import python as kfpy
from ..ops import collective as tops
from . import core
from .sync_sgd import _SynchronousSGD


def _var_setup(self):
    self._kf_sync = core.GradSync(self._kf_named, overlap=False)
    self._kf_var_step = 0
    self._kf_var_bufs = {}
    self.variance = None


def _variance_of(self):
    from ...ops import _cuda as K
    from ...ops import collective as coll
    sync = self._kf_sync
    total = None
    np_ = kfpy.current_cluster_size()
    # CUDA: gradients are already gathered in the flat buckets (pre-reduction)
    if sync.cuda_named:
        comm = sync._ensure()
        for a in sync._arenas:
            if a.fuse_grads() == 0:
                continue
            ends = a.segment_ends()
            nseg = ends.numel()
            bufs = self._kf_var_bufs.get((id(comm), id(a)))
            if bufs is None:
                # symmetric allocations are collective and never shrink: made once, reused every step
                bufs = (comm.alloc(max(nseg * 4, 16)), comm.alloc(max(nseg * 4, 16)) if comm.np > 1 else None,
                        torch.zeros(K.VAR_WORK_WORDS * comm.n_local_ranks, dtype=torch.float32, device=comm.device))
                self._kf_var_bufs[(id(comm), id(a))] = bufs
            seg, out, work = bufs
            sv = seg.view(torch.float32)
            sv[:nseg].zero_()
            ga, _k = a.grads_buf.ptr_array()
            K.call('kfb_grad_variance', comm.cstruct, ga, a.code, a.count, ends.data_ptr(), nseg,
                   sv.data_ptr(), work.data_ptr(), comm.default_blocks, comm.default_threads, K.stream_ptr())
            if comm.np > 1:
                coll.all_reduce(comm, seg, (nseg + 3) // 4 * 4, K.F32, 'sum', algo=K.ALGO_ONESHOT, out=out)
                sv = out.view(torch.float32)
            v = sv[:nseg].sqrt().sum()
            total = v if total is None else total + v
    for n, p in sync.cpu_named:
        if p.grad is None:
            continue
        g = p.grad
        s1 = tops.all_reduce_fn(g, 'sum', name='var1::' + n) / np_
        s2 = tops.all_reduce_fn(g * g, 'sum', name='var2::' + n) / np_
        v = torch.norm(s2 - s1 * s1)
        total = v if total is None else total + v.to(total.device)
    return total


def _var_step(self, closure=None):
    loss = None
    if closure is not None:
        with torch.enable_grad():
            loss = closure()
    if self._kf_var_step % self._kf_interval == 0:
        self.variance = _variance_of(self)
        if self._kf_verbose and self.variance is not None and kfpy.current_rank() == 0:
            print('Variance:', float(self.variance))
    self._kf_var_step += 1
    self._kf_sync.sync(average=True)
    super(self.__class__, self).step()
    return loss


def MonitorGradientVarianceOptimizer(optimizer, named_parameters=None, monitor_interval=1, verbose=False):
    """S-SGD that also measures the gradient variance across workers every ``monitor_interval`` steps:
    ``opt.variance`` = sum over tensors of ``|| E[g^2] - E[g]^2 ||`` (reference: optimizers/grad_variance.py:46-59);
    with FusedSGD both moments come out of one pass over peer memory."""
    if named_parameters is None:
        named = [('p%d' % i, p) for i, p in enumerate(q for g in optimizer.param_groups for q in g['params'])]
    else:
        named = list(named_parameters)
    mixin = type('_GradVariance', (object, ), {})
    mixin._kf_setup = _var_setup
    mixin.step = torch.no_grad()(_var_step)
    mixin.sync_gradients = _SynchronousSGD.sync_gradients
    return core.wrap(optimizer, mixin, _kf_named=named, _kf_average=True, _kf_overlap=False,
                     _kf_interval=max(1, int(monitor_interval)), _kf_verbose=verbose)
