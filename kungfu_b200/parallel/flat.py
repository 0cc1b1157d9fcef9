"""Flat parameter / gradient arenas in symmetric memory.

A model's parameters are re-pointed into one flat buffer per dtype ("arena")
that lives in peer-mapped memory, so that

* the S-SGD fused step can push updated parameters straight into every
  peer's model (reduce-scatter -> SGD -> all-gather in one kernel),
* SMA can average models by loading peers' parameters in place,
* PairAveraging can publish / pull whole-model snapshots with one pass,
* broadcast_parameters is one multicast store.

The reference fused tensors with tf.concat and sliced them back
(/root/reference/srcs/python/kungfu/tensorflow/ops/__init__.py:29-46); here the
tensors simply live inside the bucket, and gradients are gathered into their
bucket by one multi-tensor kernel that also yields ||g||^2.
"""
import torch

from ..ops import _cuda as K

ALIGN_ELEMS = 16


def _align(n, a=ALIGN_ELEMS):
    return (n + a - 1) // a * a


def _is_channels_last(t):
    return t.dim() == 4 and not t.is_contiguous() and t.is_contiguous(memory_format=torch.channels_last)


class Arena(object):
    """All parameters of one dtype, flattened.

    Attributes (R = n_local_ranks, 1 unless emulated):
      params_buf / grads_buf   SymmBuffers
      master, mom              fp32 local tensors [R * count]
      count                    padded element count (multiple of 16)
    """

    def __init__(self, comm, named_params, dtype, need_grads=True, multicast=True, repoint=True):
        self.comm = comm
        self.dtype = dtype
        self.code = K.dtype_code(dtype)
        self.names = [n for n, _ in named_params]
        self.params = [p for _, p in named_params]
        self.offsets = []
        off = 0
        for p in self.params:
            self.offsets.append(off)
            off += _align(p.numel())
        self.count = _align(off, ALIGN_ELEMS * 8)
        esize = torch.empty(0, dtype=dtype).element_size()
        self.nbytes = self.count * esize
        self._cl = [_is_channels_last(p.data) for p in self.params]
        self.params_buf = comm.alloc(self.nbytes, multicast=multicast) if repoint else None
        self.grads_buf = comm.alloc(self.nbytes, multicast=multicast) if need_grads else None
        self.flat_params = self.params_buf.view(dtype)[:self.count] if repoint else None
        self.flat_grads = self.grads_buf.view(dtype)[:self.count] if need_grads else None
        if repoint:
            self._repoint()
        self.master = None
        self.mom = None
        self.mom2 = None
        self._src_cache = None
        self.seg_ends = None
        # element ranges over which sharded optimizer state is split between the ranks
        # (rank r owns the r-th part of each range); one range unless the step is bucketed
        self.segments = [(0, self.count)]

    def _repoint(self):
        """Make every parameter a view into the flat buffer (values preserved)."""
        for p, off in zip(self.params, self.offsets):
            n = p.numel()
            flat = self.flat_params[off:off + n]
            if _is_channels_last(p.data):
                N, Cc, H, W = p.shape
                view = flat.view(N, H, W, Cc).permute(0, 3, 1, 2)
            else:
                view = flat.view(p.shape)
            view.copy_(p.data)
            p.data = view

    def is_cl(self, i):
        return self._cl[i]

    def grad_view(self, i):
        p, off = self.params[i], self.offsets[i]
        flat = self.flat_grads[off:off + p.numel()]
        if self._cl[i]:
            N, Cc, H, W = p.shape
            return flat.view(N, H, W, Cc).permute(0, 3, 1, 2)
        return flat.view(p.shape)

    def ensure_master(self, slots=1):
        """fp32 master copy of the parameters plus `slots` zeroed optimizer
        slots of the same layout (momentum; Adam: first and second moment)."""
        if self.master is not None and slots > 1 and self.mom2 is None:
            self.mom2 = torch.zeros_like(self.master)
        if self.master is None:
            R = self.comm.n_local_ranks
            if self.comm.emulated:
                m = torch.stack([self.params_buf.view(self.dtype, r)[:self.count].float() for r in range(R)]).reshape(-1)
            else:
                m = self.flat_params.float()
            self.master = m.contiguous()
            self.mom = torch.zeros_like(self.master)
            self.mom2 = torch.zeros_like(self.master) if slots > 1 else None
        return self.master, self.mom

    def segment_ends(self):
        if self.seg_ends is None:
            ends = [off + _align(p.numel()) for p, off in zip(self.params, self.offsets)]
            ends[-1] = self.count
            self.seg_ends = torch.tensor(ends, dtype=torch.int64, device=self.comm.device)
        return self.seg_ends

    def state_view(self, flat, i):
        """Parameter i's slice of any flat tensor laid out like this arena, in
        the parameter's logical shape (channels-last aware)."""
        p, off = self.params[i], self.offsets[i]
        v = flat[off:off + p.numel()]
        if self._cl[i]:
            N, Cc, H, W = p.shape
            return v.view(N, H, W, Cc).permute(0, 3, 1, 2)
        return v.view(p.shape)

    # ---- gradient gather (fuse) ------------------------------------------------------
    def fuse_grads(self, sumsq_out=None, stream=None, idx=None, scratch=None):
        """Copy every p.grad (or those of the parameters `idx`) into its slot of
        the flat gradient buffer (one or a few multi-tensor launches);
        optionally ||g||^2 -> sumsq_out[0]. A parameter without a gradient
        contributes zeros: its slot is cleared, so a stale gradient of an
        earlier step can never be reduced or applied again. Returns the number
        of gradients copied."""
        srcs, counts, offs, keep, temps = [], [], [], [], []
        which = range(len(self.params)) if idx is None else idx
        ctx = torch.cuda.stream(stream) if stream is not None else None
        for i in which:
            p, off = self.params[i], self.offsets[i]
            g = p.grad
            if g is None:
                if ctx is not None:
                    with ctx:
                        self.flat_grads[off:off + p.numel()].zero_()
                else:
                    self.flat_grads[off:off + p.numel()].zero_()
                continue
            if g.dtype != self.dtype:
                raise K.KungFuCudaError('gradient dtype %s != arena dtype %s' % (g.dtype, self.dtype))
            if self._cl[i]:
                if not g.is_contiguous(memory_format=torch.channels_last):
                    g = g.contiguous(memory_format=torch.channels_last)
            elif not g.is_contiguous():
                g = g.contiguous()
            # a contiguous temporary must outlive the launch below: freed earlier, the caching
            # allocator could hand its block to the next temporary and two sources would alias
            keep.append(g)
            if g is not p.grad:
                temps.append(g)
            srcs.append(g.data_ptr())
            counts.append(g.numel())
            offs.append(off)
        if not srcs:
            if sumsq_out is not None:
                if ctx is not None:
                    with ctx:
                        sumsq_out.zero_()
                else:
                    sumsq_out.zero_()
            return 0
        pa, _k1 = K.ptr_array(srcs)
        ca, _k2 = K.i64_array(counts)
        oa, _k3 = K.i64_array(offs)
        if scratch is None:
            scratch = self.comm._fuse_scratch
        K.call('kfb_fuse', pa, ca, oa, len(srcs), self.code, self.flat_grads.data_ptr(),
               self.code, sumsq_out.data_ptr() if sumsq_out is not None else None,
               scratch.data_ptr(), K.stream_ptr(stream))
        if stream is not None:
            for g in temps:
                g.record_stream(stream)
        del keep, temps
        return len(srcs)

    def point_grads_at_flat(self):
        """p.grad := view of the (reduced) flat gradient buffer, zero copy."""
        for i, p in enumerate(self.params):
            p.grad = self.grad_view(i)


def build_arenas(comm, named_parameters, need_grads=True, multicast=True):
    """Group trainable CUDA parameters by dtype, one Arena each (stable order)."""
    groups = {}
    for name, p in named_parameters:
        if not p.requires_grad:
            continue
        groups.setdefault(p.dtype, []).append((name, p))
    order = sorted(groups.keys(), key=lambda d: str(d))
    return [Arena(comm, groups[d], d, need_grads, multicast) for d in order]
