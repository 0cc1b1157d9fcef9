"""FusedSGD: torch.optim.SGD semantics on flat arenas with fp32 master weights.

One kernel per (param group, dtype) does the whole update; when driven by
``SynchronousSGDOptimizer`` the same kernel also performs the gradient
all-reduce over peer memory (reduce-scatter -> update of the owned shard ->
all-gather of the new parameters), see csrc/kernels/optim.cu:k_ssgd_fused.

Update rule = torch.optim.SGD (weight decay, momentum, dampening, nesterov),
which is what the reference's benchmarks wrap
(/root/reference/benchmarks/system/benchmark_kungfu_torch.py:76-83).
"""
import ctypes as C
import os

import torch

from ..ops import _cuda as K
from ..parallel import comm as _comm
from ..parallel.flat import build_arenas


class FusedSGD(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, momentum=0.0, dampening=0.0, weight_decay=0.0,
                 nesterov=False, comm=None, blocks=None, threads=None):
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError('Nesterov momentum requires a momentum and zero dampening')
        defaults = dict(lr=lr, momentum=momentum, dampening=dampening,
                        weight_decay=weight_decay, nesterov=nesterov)
        super().__init__(params, defaults)
        self._comm = comm
        self._arenas = None  # list of (group, [Arena])
        self._steps = 0
        self._blocks = blocks
        self.local_blocks = int(os.environ.get('KUNGFU_LOCAL_STEP_BLOCKS', '128'))
        self._threads = threads
        self._local = None
        self.sumsq_global = None  # ||averaged grad||^2 of the last sync step (device float)
        self.sumsq_local = None  # ||local grad||^2 of the last step (device float)
        self.monitor_norms = False
        self.use_nvls = os.environ.get('KUNGFU_FUSED_NVLS', '1') != '0'
        self.nvls_min_bytes = int(os.environ.get('KUNGFU_FUSED_NVLS_MIN_BYTES', str(256 << 10)))

    # -- arenas -------------------------------------------------------------------------
    @property
    def comm(self):
        if self._comm is None:
            self._comm = _comm.default_comm()
        return self._comm

    def arenas(self):
        if self._arenas is None:
            out = []
            for gi, g in enumerate(self.param_groups):
                named = [('g%d.p%d' % (gi, i), p) for i, p in enumerate(g['params']) if p.is_cuda]
                if len(named) != len(g['params']):
                    raise K.KungFuCudaError('FusedSGD only handles CUDA parameters')
                out.append((g, build_arenas(self.comm, named)))
            self._arenas = out
            dev = self.comm.device
            n = sum(len(a) for _, a in out)
            self._ss_local_parts = torch.zeros(max(n, 1), dtype=torch.float32, device=dev)
            self._ss_global_parts = torch.zeros(max(n, 1), dtype=torch.float32, device=dev)
        return self._arenas

    def rebuild(self, comm=None):
        """Re-create the arenas on a new communicator (after resize_cluster).
        Parameter values, master weights and momentum are carried over."""
        old = self._arenas
        self._comm = comm
        self._arenas = None
        self._local = None
        if hasattr(self, '_kf_stage'):
            self._kf_stage = {}  # SMA staging buffers belong to the old communicator
        # every member (joiners included) allocates its arenas here: symmetric
        # allocation is collective and must happen in the same order everywhere
        news = [a for _, ars in self.arenas() for a in ars]
        if old is None:
            return
        olds = [a for _, ars in old for a in ars]
        for o, n in zip(olds, news):
            if o.master is not None:
                n.ensure_master()
                n.master.copy_(o.master)
                n.mom.copy_(o.mom)

    def _hyper(self, g):
        h = K.SgdHyper()
        h.lr = float(g['lr'])
        h.momentum = float(g['momentum'])
        h.dampening = float(g['dampening'])
        h.weight_decay = float(g['weight_decay'])
        h.nesterov = 1 if g['nesterov'] else 0
        h.first_step = 1 if self._steps == 0 else 0
        return h

    def _local_struct(self):
        # np == 1 view of this rank for purely local steps
        if self._local is None:
            c = self.comm
            s = K.KfbComm()
            s.rank = 0
            s.np = 1
            s.pads[0] = c._struct.pads[max(c.rank, 0)]
            s.local_epoch = c._struct.local_epoch
            s.scratch = c._struct.scratch
            self._local = s
        return C.byref(self._local)

    # -- steps ------------------------------------------------------------------------------
    def _run(self, sync, grad_scale):
        comm = self.comm
        blocks = self._blocks or comm.default_blocks
        threads = self._threads or comm.default_threads
        k = 0
        for g, arenas in self.arenas():
            h = self._hyper(g)
            for a in arenas:
                a.ensure_master()
                ss_l = self._ss_local_parts[k:k + 1] if self.monitor_norms else None
                ss_g = self._ss_global_parts[k:k + 1] if self.monitor_norms else None
                a.fuse_grads(sumsq_out=ss_l)
                step_scale = grad_scale
                if sync and comm.cross:
                    # hosts are joined by the control plane: hierarchical all-reduce of
                    # the bucket first, then the purely local fused update
                    from ..ops import collective as coll
                    coll.hier_all_reduce(comm, a.grads_buf, a.count, a.code, 'sum', scale=grad_scale)
                    step_scale = 1.0
                use_nvls = (sync and comm.np > 1 and not comm.cross and a.grads_buf.mc_ptr and a.params_buf.mc_ptr
                            and self.use_nvls and a.nbytes >= self.nvls_min_bytes)
                if use_nvls:
                    # the switch reduces the gradient shard and multicasts the new parameters
                    K.call('kfb_ssgd_fused_step_nvls', comm.cstruct, a.grads_buf.mc_ptr, a.code,
                           a.params_buf.mc_ptr, a.code, a.master.data_ptr(), a.mom.data_ptr(), a.count,
                           float(step_scale), C.byref(h), ss_g.data_ptr() if ss_g is not None else None,
                           min(blocks, 32), threads, K.stream_ptr())
                    k += 1
                    continue
                step_blocks = blocks
                if sync and comm.np > 1 and not comm.cross:
                    ga, _k1 = a.grads_buf.ptr_array()
                    pa, _k2 = a.params_buf.ptr_array()
                    cs = comm.cstruct
                else:
                    me = max(comm.rank, 0)
                    ga, _k1 = K.ptr_array([a.grads_buf.ptrs[me]])
                    pa, _k2 = K.ptr_array([a.params_buf.ptrs[me]])
                    cs = self._local_struct()
                    # a purely local update is HBM-bound, not NVLink-bound: it needs every SM it can
                    # get (the comm default of 32 CTAs reaches ~1.7 TB/s; 128 is the flag-slot limit)
                    step_blocks = self.local_blocks
                K.call('kfb_ssgd_fused_step', cs, ga, a.code, pa, a.code,
                       a.master.data_ptr(), a.mom.data_ptr(), a.count, float(step_scale),
                       C.byref(h), ss_g.data_ptr() if ss_g is not None else None,
                       step_blocks, threads, K.stream_ptr())
                k += 1
        if self.monitor_norms:
            self.sumsq_local = self._ss_local_parts[:k].sum()
            self.sumsq_global = self._ss_global_parts[:k].sum()
        self._steps += 1

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        self._run(sync=False, grad_scale=1.0)
        return loss

    @torch.no_grad()
    def sync_step(self, average=True):
        """All-reduce gradients over peer memory and apply the update, fused."""
        n = self.comm.global_np if self.comm.cross else self.comm.np
        scale = 1.0 / n if average else 1.0
        self._run(sync=True, grad_scale=scale)

    # -- state ------------------------------------------------------------------------------
    def consolidate(self):
        """Make master weights / momentum complete on every rank (each rank only
        maintains the shard it owns during fused sync steps)."""
        comm = self.comm
        if comm.np <= 1 or self._arenas is None:
            return
        from ..ops import collective as coll
        for _, arenas in self._arenas:
            for a in arenas:
                if a.master is None:
                    continue
                npk = a.count // 8
                per = (npk + comm.np - 1) // comm.np * 8
                stage = comm.alloc(per * 4 * comm.np)
                sv = stage.view(torch.float32)
                for t in (a.master, a.mom):
                    lo = min(comm.rank * per, a.count)
                    hi = min(lo + per, a.count)
                    tmp = comm.alloc(per * 4)
                    tmp.view(torch.float32)[:hi - lo].copy_(t[lo:hi])
                    coll.all_gather(comm, tmp, stage, per * 4)
                    t.copy_(sv[:per * comm.np][:a.count])
        torch.cuda.synchronize()

    def broadcast_state(self):
        """Master weights, momentum and the step counter := rank 0's (after a
        resize: rank 0 consolidated before the membership changed, joiners
        start from its state)."""
        from ..torch.ops import collective as tops
        steps = torch.tensor([self._steps], dtype=torch.int64)
        self._steps = int(tops.all_reduce_fn(steps, 'max', name='kungfu::fused-sgd-steps')[0])
        if self.comm.np <= 1:
            return
        for gi, (_, arenas) in enumerate(self.arenas()):
            for ai, a in enumerate(arenas):
                a.ensure_master()
                tops.broadcast(a.master, name='kungfu::master.%d.%d' % (gi, ai))
                tops.broadcast(a.mom, name='kungfu::mom.%d.%d' % (gi, ai))
        torch.cuda.synchronize()
