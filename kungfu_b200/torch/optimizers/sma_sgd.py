"""Synchronous model averaging (SMA / EA-SGD).

Each step: avg = mean over workers of the model, v <- (1-alpha) v + alpha avg,
then the LOCAL gradients are applied. Reference:
/root/reference/srcs/python/kungfu/tensorflow/optimizers/sma_sgd.py:50-74.

With a fused optimizer (FusedSGD / FusedAdamW) the step is two kernels of
csrc/kernels/optim.cu:k_sma_fused over peer memory:

* phase 1, the model average (through the NVSwitch with multimem.ld_reduce /
  multimem.st when the arenas have a multicast alias). It only READS
  parameters, so ``zero_grad()`` -- the first thing a training step does --
  launches it on a side stream and it runs under forward and backward;
* phase 2, blend + update with the local gradient, purely local.

Any other optimizer on CUDA parameters takes the same route with library
pieces: the parameters are moved into a flat symmetric arena once, the average
is one all-reduce kernel into a staging buffer, the blend one element-wise
pass, then the wrapped optimizer's own step. No tensor goes through the host.
"""
import ctypes as C

import torch

from ... import python as kfpy
from ..ops import collective as tops
from . import core

SMA_CHANNEL = 2


class _SynchronousAveraging(object):
    def _kf_setup(self):
        self._kf_stage = {}
        self._kf_side = None
        self._kf_avg_pending = False
        self._kf_garenas = None
        self._kf_gcomm = None

    # -- fused path ----------------------------------------------------------------------
    def _sma_stage(self, a):
        comm = self.comm
        stage = self._kf_stage.get(id(a))
        if stage is None and (comm.np > 1 or comm.cross):
            stage = comm.alloc(a.count * 4, multicast=True)
            self._kf_stage[id(a)] = stage
        return stage

    def _sma_launch(self, phase, stream=None, channel=0):
        from ...ops import _cuda as K
        comm = self.comm
        threads = self._threads or comm.default_threads
        for g, arenas in self.arenas():
            for a in arenas:
                a.ensure_master(slots=self._n_slots())
                h = self._hyper(g, a)
                stage = self._sma_stage(a)
                np1 = comm.np <= 1 or comm.cross
                if phase == 1 and np1:
                    continue
                if phase != 1:
                    a.fuse_grads()
                pa, _k1 = a.params_buf.ptr_array()
                sa, _k2 = stage.ptr_array() if stage is not None else (None, None)
                use_mc = (not np1 and a.params_buf.mc_ptr and stage is not None and stage.mc_ptr
                          and getattr(self, 'use_nvls', True))
                if phase == 2 and not comm.emulated:
                    # purely local and HBM-bound: all the CTAs the launch check allows
                    blocks, cs = self.local_blocks, comm.cstruct
                else:
                    blocks, cs = comm.channel_blocks(channel, self._blocks), comm.channel(channel)
                grad = a.grads_buf.ptrs[max(comm.rank, 0)]
                if comm.emulated:
                    garr, _k3 = a.grads_buf.ptr_array()
                    grad = C.cast(garr, C.c_void_p)
                if comm.cross and phase != 1:
                    # several hosts: NVLink inside the host, control plane between the host masters
                    from ...ops import collective as coll
                    sv = stage.view(torch.float32)[:a.count]
                    sv.copy_(a.flat_params)
                    coll.hier_all_reduce(comm, stage, a.count, K.F32, 'sum', scale=1.0 / comm.global_np)
                    self._sma_blend_cross(a, sv)
                    pa, _k1 = K.ptr_array([a.params_buf.ptrs[max(comm.rank, 0)]])
                    K.call('kfb_sma_fused_step_ex', self._local_struct(), pa, a.code, None, None, None, grad,
                           a.code, a.master.data_ptr(), a.mom.data_ptr(), a.count, 0.0, 1.0, C.byref(h), 2,
                           self.local_blocks, threads, K.stream_ptr(stream))
                    continue
                K.call('kfb_sma_fused_step_ex', cs, pa, a.code, sa,
                       a.params_buf.mc_ptr if use_mc else None, stage.mc_ptr if use_mc else None,
                       grad, a.code, a.master.data_ptr(), a.mom.data_ptr(), a.count,
                       float(self._kf_alpha), 1.0, C.byref(h), phase, blocks, threads, K.stream_ptr(stream))

    def _sma_blend_cross(self, a, avg):
        a.master.add_(avg - a.master, alpha=float(self._kf_alpha))

    def begin_step(self):
        """Start averaging the model on a side stream (fused optimizers on one
        NVSwitch domain). Called by zero_grad(); without it step() does the
        whole thing in one kernel after backward."""
        if not core.is_fused(self):
            return
        comm = self.comm
        if comm.np <= 1 or comm.cross or comm.emulated or self._kf_avg_pending:
            return
        if self._kf_side is None:
            self._kf_side = torch.cuda.Stream(device=comm.device, priority=-1)
        self._kf_side.wait_stream(torch.cuda.current_stream())
        self._sma_launch(1, stream=self._kf_side, channel=SMA_CHANNEL)
        self._kf_avg_pending = True

    def zero_grad(self, set_to_none=True):
        super(self.__class__, self).zero_grad(set_to_none=set_to_none)
        if getattr(self, 'overlap_average', True):
            self.begin_step()

    def _fused_step(self):
        if self._kf_avg_pending:
            torch.cuda.current_stream().wait_stream(self._kf_side)
            self._kf_avg_pending = False
            self._sma_launch(2)
        else:
            self._sma_launch(0)
        self._after_step()

    # -- any other optimizer -------------------------------------------------------------
    def _generic_arenas(self):
        from ...parallel import comm as pc
        from ...parallel.flat import build_arenas
        comm = pc.default_comm()
        if self._kf_garenas is None or self._kf_gcomm is not comm:
            named = [(n, p) for n, p in self._kf_named if p.is_cuda]
            self._kf_garenas = [(a, comm.alloc(a.nbytes, multicast=True))
                                for a in build_arenas(comm, named, need_grads=False)]
            self._kf_gcomm = comm
        return comm, self._kf_garenas

    def _generic_average(self):
        np_ = kfpy.current_cluster_size()
        if np_ <= 1:
            return
        alpha = float(self._kf_alpha)
        if any(p.is_cuda for _, p in self._kf_named):
            from ...ops import collective as coll
            comm, arenas = self._generic_arenas()
            for a, stage in arenas:
                if comm.cross:
                    stage.view(a.dtype)[:a.count].copy_(a.flat_params)
                    coll.hier_all_reduce(comm, stage, a.count, a.code, 'sum', scale=1.0 / comm.global_np)
                else:
                    coll.all_reduce(comm, a.params_buf, a.count, a.code, 'sum', scale=1.0 / comm.np, out=stage,
                                    algo=None if a.nbytes > coll.ONESHOT_MAX_BYTES else 0)
                a.flat_params.mul_(1 - alpha).add_(stage.view(a.dtype)[:a.count], alpha=alpha)
        for n, p in self._kf_named:
            if p.is_cuda:
                continue
            avg = tops.all_reduce_fn(p.data, 'sum', name='sma::' + n)
            p.data.mul_(1 - alpha).add_(avg, alpha=alpha / np_)

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
    gradients are applied (reference: optimizers/sma_sgd.py:50-74). With a fused optimizer the average runs on the
    NVSwitch under forward/backward (started by ``zero_grad()``), blend + update are one local kernel."""
    if named_parameters is None:
        named = [('p%d' % i, p) for i, p in enumerate(q for g in optimizer.param_groups for q in g['params'])]
    else:
        named = [(n, p) for n, p in named_parameters if p.requires_grad]
    return core.wrap(optimizer, _SynchronousAveraging, _kf_named=named, _kf_alpha=float(alpha))
