"""Shared machinery of the distributed optimizer wrappers.

* ``wrap``      the reference's dynamic-subclass trick, so that
                ``isinstance(wrapped, type(original))`` keeps holding
                (/root/reference/srcs/python/kungfu/torch/optimizers/sync_sgd.py:29-32).
* ``GradSync``  bucketed gradient all-reduce over peer memory for ANY wrapped
                optimizer: gradients are gathered into flat symmetric buckets
                by one multi-tensor kernel (with ||g||^2 for free), reduced by
                the NVLink kernels with the 1/np scale folded in, and handed
                back as views (no scatter copy). With ``overlap=True`` buckets
                are launched from autograd hooks on a high-priority stream
                while backward is still running, in a fixed order on every
                rank (the reference needed a runtime order negotiation,
                /root/reference/srcs/cpp/src/nccl/scheduler.cpp:94-120).
"""
import torch

from ... import python as kfpy
from ..ops import collective as tops


def wrap(optimizer, mixin, **attrs):
    body = {k: v for k, v in mixin.__dict__.items() if k not in ('__dict__', '__weakref__')}
    clazz = type(optimizer.__class__.__name__, (optimizer.__class__, ), body)
    obj = clazz.__new__(clazz)
    obj.__dict__.update(optimizer.__dict__)
    for k, v in attrs.items():
        setattr(obj, k, v)
    obj._kf_setup()
    return obj


def is_fused(optimizer):
    from ...optim.fused import FusedOptimizer
    return isinstance(optimizer, FusedOptimizer)


class GradSync(object):
    BUCKET_BYTES = 32 << 20

    def __init__(self, named_parameters, overlap=False, bucket_bytes=None, average=True):
        self.named = [(n, p) for n, p in named_parameters if p.requires_grad]
        self.cuda_named = [(n, p) for n, p in self.named if p.is_cuda]
        self.cpu_named = [(n, p) for n, p in self.named if not p.is_cuda]
        self.overlap = overlap
        self.bucket_bytes = bucket_bytes or self.BUCKET_BYTES
        self._arenas = None
        self._buckets = None
        self._hooks = []
        self._comm_id = None
        self.sumsq_local = None
        self.sumsq_global = None
        self.monitor_norms = False
        # buckets are launched from autograd hooks, long before sync() is called:
        # the reduction's scale has to be known when the hooks fire
        self._average = bool(average)

    # -- lazily built per communicator (rebuilt after a resize) -----------------------
    def _ensure(self):
        from ...parallel import comm as pc
        from ...parallel.flat import Arena
        comm = pc.default_comm()
        if self._arenas is not None and self._comm_id == id(comm):
            return comm
        self._remove_hooks()
        groups = {}
        for n, p in self.cuda_named:
            groups.setdefault(p.dtype, []).append((n, p))
        self._arenas = [Arena(comm, groups[d], d, need_grads=True, repoint=False)
                        for d in sorted(groups, key=str)]
        self._comm_id = id(comm)
        # The bucket layout fixes the order in which gradients are reduced. The
        # reference negotiated that order at run time (rank 0's arrival order was
        # broadcast and adopted); here it is static, and every rank proves it holds
        # the same layout before the first collective.
        import hashlib
        layout = '|'.join('%s:%s:%s' % (n, tuple(p.shape), p.dtype) for n, p in self.cuda_named)
        digest = torch.tensor(list(hashlib.sha1(layout.encode()).digest()), dtype=torch.uint8)
        if kfpy.current_cluster_size() > 1 and not tops.consensus(digest, name='kungfu::grad-layout'):
            raise RuntimeError('workers registered different parameter lists (names / shapes / order differ)')
        dev = comm.device
        k = max(1, len(self._arenas))
        self._ss_local = torch.zeros(k, dtype=torch.float32, device=dev)
        self._ss_global = torch.zeros(k, dtype=torch.float32, device=dev)
        if self.overlap and comm.np > 1 and not comm.cross:
            # (several hosts: the cross-host leg runs on the host after backward, see _sync_cuda)
            self._build_buckets(comm)
        else:
            self._buckets = None
        return comm

    # -- plain (post-backward) path -------------------------------------------------------
    def sync(self, average=True):
        np_ = kfpy.current_cluster_size()
        handles = []
        for n, p in self.cpu_named:
            if p.grad is not None:
                handles.append(tops.inplace_all_reduce_async_op(p.grad, n, 'sum'))
        if self.cuda_named:
            comm = self._ensure()
            if self._buckets is not None:
                self._finish_overlap(average)
            else:
                self._sync_cuda(comm, average)
        tops.wait_all_handles(handles)
        if average and np_ > 1:
            for n, p in self.cpu_named:
                if p.grad is not None:
                    p.grad.div_(np_)

    def _sync_cuda(self, comm, average):
        from ...ops import collective as coll
        scale = 1.0 / comm.np if average else 1.0
        for k, a in enumerate(self._arenas):
            ss_l = self._ss_local[k:k + 1] if self.monitor_norms else None
            ss_g = self._ss_global[k:k + 1] if self.monitor_norms else None
            if a.fuse_grads(sumsq_out=ss_l) == 0:
                continue
            if comm.cross:
                gscale = 1.0 / comm.global_np if average else 1.0
                coll.hier_all_reduce(comm, a.grads_buf, a.count, a.code, 'sum', scale=gscale)
                if self.monitor_norms:
                    ss_g.copy_(a.flat_grads.float().pow(2).sum().reshape(1))
            elif comm.np > 1:
                coll.all_reduce(comm, a.grads_buf, a.count, a.code, 'sum', scale=scale, sumsq_out=ss_g)
            elif self.monitor_norms:
                ss_g.copy_(ss_l)
            a.point_grads_at_flat()
        if self.monitor_norms:
            self.sumsq_local = self._ss_local.sum()
            self.sumsq_global = self._ss_global.sum()

    # -- overlapped path -----------------------------------------------------------------------
    def _build_buckets(self, comm):
        self._stream = torch.cuda.Stream(device=comm.device, priority=-1)
        self._buckets = []  # (arena, [param idx], lo_elem, hi_elem)
        for a in self._arenas:
            esize = a.flat_grads.element_size()
            cur, lo, nbytes = [], None, 0
            # backward produces gradients roughly in reverse registration order
            for i in reversed(range(len(a.params))):
                cur.append(i)
                nbytes += a.params[i].numel() * esize
                if nbytes >= self.bucket_bytes or i == 0:
                    lo = a.offsets[min(cur)]
                    last = max(cur)
                    hi = a.offsets[last + 1] if last + 1 < len(a.params) else a.count
                    self._buckets.append({'arena': a, 'idx': list(cur), 'lo': lo, 'hi': hi,
                                          'pending': len(cur), 'launched': False, 'slot': len(self._buckets)})
                    cur, nbytes = [], 0
        self._next_bucket = 0
        self._where = {}
        for b_i, b in enumerate(self._buckets):
            for i in b['idx']:
                self._where[id(b['arena'].params[i])] = b_i
        for _, p in self.cuda_named:
            self._hooks.append(p.register_post_accumulate_grad_hook(self._on_grad))
        k = max(1, len(self._buckets))
        self._ss_local = torch.zeros(k, dtype=torch.float32, device=comm.device)
        self._ss_global = torch.zeros(k, dtype=torch.float32, device=comm.device)
        self._ss_out = torch.zeros(2, dtype=torch.float32, device=comm.device)
        self._fuse_scratch = torch.zeros(1024, dtype=torch.float32, device=comm.device)

    def _remove_hooks(self):
        for h in self._hooks:
            h.remove()
        self._hooks = []

    def _on_grad(self, p):
        b = self._buckets[self._where[id(p)]]
        b['pending'] -= 1
        self._launch_ready()

    def _launch_ready(self, force=False):
        from ...parallel import comm as pc
        comm = pc.default_comm()
        while self._next_bucket < len(self._buckets):
            b = self._buckets[self._next_bucket]
            if b['pending'] > 0 and not force:
                return
            self._launch_bucket(comm, b)
            self._next_bucket += 1

    def _launch_bucket(self, comm, b):
        from ...ops import collective as coll
        a = b['arena']
        self._stream.wait_stream(torch.cuda.current_stream())
        scale = 1.0 / comm.np if self._average else 1.0
        k = b['slot']
        ss_l = self._ss_local[k:k + 1] if self.monitor_norms else None
        ss_g = self._ss_global[k:k + 1] if self.monitor_norms else None
        a.fuse_grads(sumsq_out=ss_l, stream=self._stream, idx=b['idx'], scratch=self._fuse_scratch)
        esize = a.flat_grads.element_size()
        coll.all_reduce(comm, a.grads_buf, b['hi'] - b['lo'], a.code, 'sum', scale=scale,
                        byte_offset=b['lo'] * esize, stream=self._stream, sumsq_out=ss_g,
                        blocks=comm.channel_blocks(2, 16), channel=2)
        b['launched'] = True

    def _finish_overlap(self, average):
        if bool(average) != self._average:
            raise RuntimeError('GradSync(overlap=True) was built with average=%s; the buckets of this step '
                               'were already reduced with that scale' % self._average)
        self._launch_ready(force=True)
        torch.cuda.current_stream().wait_stream(self._stream)
        for a in self._arenas:
            a.point_grads_at_flat()
        for b in self._buckets:
            b['pending'] = len(b['idx'])
            b['launched'] = False
        self._next_bucket = 0
        if self.monitor_norms:
            from ...ops import _cuda as K
            n = len(self._buckets)
            K.call('kfb_sum_slots', self._ss_local.data_ptr(), n, self._ss_out.data_ptr(), K.stream_ptr())
            K.call('kfb_sum_slots', self._ss_global.data_ptr(), n, self._ss_out.data_ptr() + 4, K.stream_ptr())
            self.sumsq_local = self._ss_out[0]
            self.sumsq_global = self._ss_out[1]

    def arenas(self):
        self._ensure()
        return self._arenas
