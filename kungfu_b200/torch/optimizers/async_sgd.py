"""Pair averaging (AD-PSGD): average with ONE randomly chosen peer per step.

Reference: /root/reference/srcs/python/kungfu/tensorflow/optimizers/async_sgd.py:13-142
 - target ~ U[0, np); if target == rank: (target + 1) % np        (:71-75)
 - first step: publish the model, barrier                           (:106-108)
 - v <- 0.5 (v + v_target); apply local gradients; publish          (:128-142)
With FusedSGD the pull is a one-sided load of the peer's published snapshot out
of its GPU memory (seqlock + double buffer), fused with the blend, the SGD
update and the publication of the new snapshot: one kernel, no barrier, no
NCCL (csrc/kernels/optim.cu:k_pair_avg). Other optimizers use the host p2p
store (request / save), like the reference.
"""
import ctypes as C
import random

import torch

from ... import python as kfpy
from ..ops import collective as tops
from . import core


class _PairAveraging(object):
    def _kf_setup(self):
        self._kf_snap = {}
        self._kf_started = False
        self._kf_rng = random.Random(self._kf_seed)
        self.torn_reads = None

    def _pick(self):
        np_, rank = kfpy.current_cluster_size(), kfpy.current_rank()
        if np_ <= 1:
            return -1
        t = self._kf_rng.randrange(np_)
        return (t + 1) % np_ if t == rank else t

    # -- fused path --------------------------------------------------------------------
    def _fused_init(self):
        from ...ops import _cuda as K
        comm = self.comm
        self.torn_reads = torch.zeros(1, dtype=torch.int32, device=comm.device)
        for g, arenas in self.arenas():
            for a in arenas:
                a.ensure_master()
                snap = comm.alloc(2 * a.nbytes)
                self._kf_snap[id(a)] = snap
                K.call('kfb_pair_publish', comm.cstruct, a.params_buf.ptrs[max(comm.rank, 0)],
                       snap.ptrs[max(comm.rank, 0)], a.nbytes, a.nbytes,
                       comm.default_blocks, comm.default_threads, K.stream_ptr())
        torch.cuda.current_stream().synchronize()
        kfpy.run_barrier()

    def _fused_step(self):
        from ...ops import _cuda as K
        comm = self.comm
        if len([a for _, ars in self.arenas() for a in ars]) > 1:
            raise RuntimeError('fused PairAveraging needs all parameters in one dtype (one snapshot sequence)')
        if not self._kf_started:
            self._fused_init()
            self._kf_started = True
        me = max(comm.rank, 0)
        device_select = getattr(self, 'device_select', False) and comm.np > 1
        target = -1 if device_select else self._pick()
        # one peer pull (NVLink) + a full-model local update (HBM): no CTA-indexed barrier in this kernel, so
        # it takes as many CTAs as the launch check allows instead of the NVLink-sized comm default
        blocks = self._blocks or max(comm.default_blocks, K.MAX_BLOCKS)
        threads = self._threads or comm.default_threads
        for g, arenas in self.arenas():
            h = self._hyper(g)
            for a in arenas:
                a.fuse_grads()
                snap = self._kf_snap[id(a)]
                if device_select:
                    # the kernel draws its own peer from (seed, step counter, rank): nothing in this launch
                    # depends on a host-side choice, so the step can be captured in a CUDA graph
                    if getattr(self, '_kf_step_t', None) is None:
                        self._kf_step_t = torch.zeros(1, dtype=torch.int64, device=comm.device)
                    sa, _keep = K.ptr_array(snap.ptrs)
                    K.call('kfb_pair_avg_step_dev', comm.cstruct, sa, a.nbytes, a.params_buf.ptrs[me], a.code,
                           a.grads_buf.ptrs[me], a.code, a.master.data_ptr(), a.mom.data_ptr(), a.count, 1.0,
                           C.byref(h), self.torn_reads.data_ptr(), self._kf_step_t.data_ptr(),
                           int(self._kf_seed) & 0xFFFFFFFFFFFFFFFF, blocks, threads, K.stream_ptr())
                    self._kf_step_t.add_(1)
                    continue
                K.call('kfb_pair_avg_step', comm.cstruct, target,
                       snap.ptrs[target] if target >= 0 else snap.ptrs[me], snap.ptrs[me], a.nbytes,
                       a.params_buf.ptrs[me], a.code, a.grads_buf.ptrs[me], a.code,
                       a.master.data_ptr(), a.mom.data_ptr(), a.count, 1.0, C.byref(h),
                       self.torn_reads.data_ptr(), blocks, threads, K.stream_ptr())
        self._steps += 1

    # -- generic path (host p2p store) -------------------------------------------------------
    def _fused_model(self):
        return torch.cat([p.data.detach().reshape(-1).float().cpu() for _, p in self._kf_named])

    def _generic_step(self):
        name = self._kf_model_name
        if not self._kf_started:
            tops.save_variable(self._fused_model(), name)
            kfpy.run_barrier()
            self._kf_started = True
        target = self._pick()
        if target >= 0:
            mine = self._fused_model()
            ok, other = tops.request_variable(target, name, mine)
            if ok:
                off = 0
                for _, p in self._kf_named:
                    n = p.numel()
                    o = other[off:off + n].view(p.shape).to(p.device, p.dtype)
                    p.data.add_(o).mul_(0.5)
                    off += n
        super(self.__class__, self).step()
        tops.save_variable(self._fused_model(), name)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if core.is_fused(self):
            self._fused_step()
        else:
            self._generic_step()
        return loss


def PairAveragingOptimizer(optimizer, named_parameters=None, fuse_requests=True,
                           fused_model_name=None, seed=None):
    """AD-PSGD: every step average the model with ONE randomly chosen peer (``v <- (v + v_peer) / 2``), then apply the
    local gradients; no barrier, no all-reduce. With FusedSGD the pull, blend, update and publication are one kernel
    over peer memory; set ``opt.device_select = True`` to let the kernel draw the peer (CUDA-graph capturable).
    ``seed`` seeds the peer choice (default: a function of the rank)."""
    if named_parameters is None:
        named = [('p%d' % i, p) for i, p in enumerate(q for g in optimizer.param_groups for q in g['params'])]
    else:
        named = [(n, p) for n, p in named_parameters if p.requires_grad]
    if seed is None:
        seed = 1000003 * (kfpy.current_rank() + 1)
    return core.wrap(optimizer, _PairAveraging, _kf_named=named,
                     _kf_model_name=fused_model_name or 'kungfu::fused-model', _kf_seed=seed)
