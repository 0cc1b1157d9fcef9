"""Peer-to-peer model ops over the host model store: selectors, fused-model
averaging, and the prefetching variants that keep the network off the step.

Reference: /root/reference/srcs/cpp/src/tensorflow/ops/cpu/peer_to_peer.cpp
(`ModelAveraging` :72-153, `AsyncModelAveraging` :156-258, `SaveModel` :260-326,
`RequestModel` :328-408, `AsyncRequestModel` :411-524, selectors :8-66) and
/root/reference/srcs/cpp/include/kungfu/tensorflow/model_buffer.hpp:13-53.

The store and the double buffer are native (csrc/ctrl/transport.cpp
`BlobStore`, csrc/capi/capi.cpp `kf_prefetch_*`): the request for the next step
runs on the runtime's worker threads while Python trains. These are host-path
ops (CPU tensors, or the fp32 host copy of a small model); models that live on
GPUs exchange snapshots through PairAveragingOptimizer's device path, which
never leaves NVLink.
"""
import random

import torch

from ... import python as kfpy
from . import local as _local

MODEL_NAME = 'kungfu::fused-model'


class PeerSelector(object):
    """`random` or `roundrobin` over a fixed list of ranks (peer_to_peer.cpp:8-66)."""

    def __init__(self, strategy, ranks, seed=None):
        self.ranks = [int(r) for r in ranks]
        if not self.ranks:
            raise ValueError('ranks must not be empty')
        if strategy not in ('random', 'roundrobin'):
            raise ValueError('unsupported selection strategy: %s' % strategy)
        self.strategy = strategy
        self._t = 0
        self._rng = random.Random(seed)

    def next(self):
        if self.strategy == 'random':
            return self.ranks[self._rng.randrange(len(self.ranks))]
        now = self._t
        self._t = (self._t + 1) % len(self.ranks)
        return self.ranks[now]

    __call__ = next


def _other_ranks():
    me = kfpy.current_rank()
    return [r for r in range(kfpy.current_cluster_size()) if r != me]


def _tensors(named_tensors):
    return [t for _, t in named_tensors] if named_tensors and isinstance(named_tensors[0], (tuple, list)) else list(named_tensors)


def _split(flat, tensors):
    out, off = [], 0
    for t in tensors:
        n = t.numel()
        out.append(flat[off:off + n].view(t.shape))
        off += n
    return out


class _ModelOp(object):
    def __init__(self, named_tensors, peer_selection_strategy='random', ranks=None, name=MODEL_NAME, seed=None):
        self.tensors = _tensors(named_tensors)
        self.name = name
        self.total = sum(t.numel() for t in self.tensors)
        self._ranks = ranks
        self._strategy = peer_selection_strategy
        self._seed = seed
        self._selector = None
        self._landing = None

    def _next_peer(self):
        # built lazily: the cluster may not be up when the op is constructed
        if self._selector is None:
            ranks = self._ranks if self._ranks is not None else _other_ranks()
            if not ranks:  # a cluster of one: its own snapshot
                ranks = [kfpy.current_rank()]
            self._selector = PeerSelector(self._strategy, ranks, self._seed)
        return self._selector.next()

    def _native(self):
        """fp32 CPU tensors stored densely: the runtime gathers, pulls and blends them in place (one pass each)."""
        return all(t.dtype == torch.float32 and not t.is_cuda and t.is_contiguous() for t in self.tensors)

    def _ptrs(self):
        return [t.data_ptr() for t in self.tensors], [t.numel() for t in self.tensors]

    def save(self):
        """Publish this worker's current model (SaveModel)."""
        if self._native():
            ptrs, counts = self._ptrs()
            kfpy.peer().save_parts(self.name, ptrs, [4 * c for c in counts])
            return self.total
        return _local.save_model([(None, t) for t in self.tensors], self.name)

    def _blend(self, flat):
        with torch.no_grad():
            if self._native():
                ptrs, counts = self._ptrs()
                kfpy.peer().blend_f32(ptrs, counts, flat.data_ptr())
                return
            for t, o in zip(self.tensors, _split(flat, self.tensors)):
                t.mul_(0.5).add_(o.to(device=t.device, dtype=t.dtype), alpha=0.5)


class ModelAveraging(_ModelOp):
    """Each call: pull the fused model of the next selected peer (blocking) and set
    every tensor to ``0.5 * (mine + theirs)`` in place. Returns the peer's rank, or
    -1 if it had nothing published yet (the tensors are then left alone)."""

    def __call__(self):
        target = self._next_peer()
        if self._landing is None:
            self._landing = torch.empty(self.total, dtype=torch.float32)  # kept: no 100 MB allocation per step
        flat = self._landing
        if not kfpy.peer().request(target, self.name, flat.data_ptr(), flat.numel() * 4, None):
            return -1
        self._blend(flat)
        return target


class RequestModel(_ModelOp):
    """Each call returns the selected peer's model as a list of tensors shaped like
    this worker's (``None`` if that peer has not published one)."""

    def __call__(self):
        ok, out = _local.request_model(self._next_peer(), [(None, t) for t in self.tensors], self.name)
        return out if ok else None


class _Prefetching(_ModelOp):
    def __init__(self, *args, **kw):
        super(_Prefetching, self).__init__(*args, **kw)
        self._pf = None
        self.fresh = 0  # completed background pulls so far

    def _pull(self):
        peer = kfpy.peer()
        if self._pf is None:
            self._pf = peer.prefetch_new(self.name, self.total * 4)
        flat = torch.empty(self.total, dtype=torch.float32)
        ok, fresh = peer.prefetch_next(self._pf, self._next_peer(), flat.data_ptr())
        if not ok:
            return None
        self.fresh = fresh
        return flat

    def close(self):
        if self._pf is not None:
            pf, self._pf = self._pf, None
            kfpy.peer().prefetch_free(pf)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class AsyncModelAveraging(_Prefetching):
    """ModelAveraging with a double buffer: only the first call waits for the network;
    every call starts the pull for the next one in the background and blends with
    the newest model that has fully arrived. Returns False while the first peer has
    nothing published."""

    def __call__(self):
        if self._native():
            peer = kfpy.peer()
            if self._pf is None:
                self._pf = peer.prefetch_new(self.name, self.total * 4)
            ptrs, counts = self._ptrs()
            with torch.no_grad():
                ok, fresh = peer.prefetch_blend_f32(self._pf, self._next_peer(), ptrs, counts)
            if ok:
                self.fresh = fresh
            return ok
        flat = self._pull()
        if flat is None:
            return False
        self._blend(flat)
        return True


class AsyncRequestModel(_Prefetching):
    """RequestModel with the same double buffer: returns the newest fully received
    peer model (list of tensors), ``None`` while the first peer has nothing published."""

    def __call__(self):
        flat = self._pull()
        if flat is None:
            return None
        return [o.to(device=t.device, dtype=t.dtype) for t, o in zip(self.tensors, _split(flat, self.tensors))]


def model_averaging(named_tensors, peer_selection_strategy='random', **kw):
    return ModelAveraging(named_tensors, peer_selection_strategy, **kw)


def async_model_averaging(named_tensors, peer_selection_strategy='random', **kw):
    return AsyncModelAveraging(named_tensors, peer_selection_strategy, **kw)


def async_request_model(named_tensors, peer_selection_strategy='random', **kw):
    return AsyncRequestModel(named_tensors, peer_selection_strategy, **kw)
