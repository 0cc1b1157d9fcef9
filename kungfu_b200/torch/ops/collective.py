"""Tensor collectives with the reference's torch API.

Reference: /root/reference/srcs/python/kungfu/torch/ops/collective.py:8-52 and
the pybind ops behind it (/root/reference/srcs/cpp/src/torch/ops/{cpu,cuda}/
collective.cpp). There a CUDA tensor was copied to the host, reduced over TCP
and copied back; here a CUDA tensor is reduced by one NVLink kernel out of a
symmetric staging buffer, and only CPU tensors use the host control plane.
All dtypes of the runtime are accepted (the reference's torch binding was
float32 only).
"""
import ctypes as C
import threading

import torch

from ... import _host as H
from ... import python as kfpy

_HOST_DTYPES = {
    torch.uint8: H.U8, torch.int8: H.I8, torch.int16: H.I16, torch.int32: H.I32,
    torch.int64: H.I64, torch.float16: H.F16, torch.float32: H.F32,
    torch.float64: H.F64, torch.bool: H.BOOL, torch.bfloat16: H.BF16,
}

_lock = threading.Lock()
_handles = {}
_next_handle = [1]
_anon = [0]
_staging = {}
_comm_stream = {}


def _host_dtype(t):
    try:
        return _HOST_DTYPES[t.dtype]
    except KeyError:
        raise TypeError('unsupported dtype %s' % t.dtype)


def _name(name, kind, concurrent=False):
    """Wire name of a collective. Messages of one name between two peers are
    FIFO, so unnamed synchronous ops can share a constant name (every rank must
    issue them in the same order anyway). Unnamed ASYNC ops may be in flight
    together and get a sequence number, which restarts whenever the membership
    changes so that joiners and survivors count from the same point."""
    if name is None:
        if concurrent:
            _anon[0] += 1
            name = 'kungfu::%s::%d' % (kind, _anon[0])
        else:
            name = 'kungfu::%s' % kind
    return name.encode()


def _reset_names():
    _anon[0] = 0


def _check(rc, what):
    if rc != 0:
        raise H.KungFuHostError('%s failed: %s' % (what, H.lib().kf_last_error().decode()))


# ---- CUDA path --------------------------------------------------------------------------
def _comm():
    from ...parallel import comm
    return comm.default_comm()


def _stage(comm, key, nbytes):
    """Symmetric staging buffer of at least nbytes (grown collectively: every
    rank sees the same sequence of sizes because collectives are called with
    matching tensors on every rank)."""
    k = (id(comm), key)
    buf = _staging.get(k)
    if buf is None or buf.nbytes < nbytes:
        cap = max(1 << 20, 1 << (int(nbytes) - 1).bit_length())
        buf = comm.alloc(cap, multicast=True)
        _staging[k] = buf
    return buf


def reset_staging():
    _staging.clear()
    _comm_stream.clear()


def drop_staging(comm):
    """Forget the staging buffers / side stream of one communicator (it is being closed)."""
    for k in [k for k in _staging if k[0] == id(comm)]:
        del _staging[k]
    _comm_stream.pop(id(comm), None)


ASYNC_CHANNEL = 1  # the asynchronous ops run on a side stream: their own CTA rows, epochs and staging buffers


def _cuda_all_reduce(x, y, op, channel=0):
    from ...ops import _cuda as K
    from ...ops import collective as coll
    comm = _comm()
    opc = K.OPS[op]
    tag = '' if channel == 0 else str(channel)
    if comm.cross:
        # job spans hosts: NVLink inside the host, control plane between host masters
        xc = x.contiguous()
        nbytes = xc.numel() * xc.element_size()
        src = _stage(comm, 'in', (nbytes + 15) // 16 * 16)
        src.local[:nbytes].copy_(xc.view(-1).view(torch.uint8))
        coll.hier_all_reduce(comm, src, xc.numel(), K.dtype_code(x.dtype), opc)
        y.copy_(src.local[:nbytes].view(x.dtype).view(xc.shape))
        return
    if comm.np == 1:
        if y.data_ptr() != x.data_ptr():
            y.copy_(x)
        return
    xc = x.contiguous()
    nbytes = xc.numel() * xc.element_size()
    pad = (nbytes + 15) // 16 * 16
    src = _stage(comm, 'in' + tag, pad)
    src.local[:nbytes].copy_(xc.view(-1).view(torch.uint8))
    code = K.dtype_code(x.dtype)
    algo = coll.choose_algo(comm, nbytes, src, opc, code)
    if algo == K.ALGO_ONESHOT:
        dst = _stage(comm, 'out' + tag, pad)
        coll.all_reduce(comm, src, xc.numel(), code, opc, algo=algo, out=dst, channel=channel)
    else:
        dst = src
        coll.all_reduce(comm, src, xc.numel(), code, opc, algo=algo, channel=channel)
    out = dst.local[:nbytes].view(x.dtype).view(xc.shape)
    y.copy_(out)


def _cuda_broadcast(x, channel=0):
    from ...ops import collective as coll
    comm = _comm()
    tag = '' if channel == 0 else str(channel)
    if comm.cross:
        # rare, latency-insensitive op: stage through the host control plane
        h = x.detach().contiguous().cpu().view(-1)
        _check(H.lib().kf_broadcast(kfpy.peer()._h, h.data_ptr(), h.data_ptr(), h.numel(), _host_dtype(h),
                                    b'kungfu::xhost-bcast'), 'broadcast')
        x.copy_(h.view(x.shape))
        return
    if comm.np == 1:
        return
    xc = x.contiguous()
    nbytes = xc.numel() * xc.element_size()
    buf = _stage(comm, 'in' + tag, (nbytes + 15) // 16 * 16)
    if comm.rank == 0:
        buf.local[:nbytes].copy_(xc.view(-1).view(torch.uint8))
    coll.broadcast(comm, buf, (nbytes + 15) // 16 * 16, root=0, channel=channel)
    x.copy_(buf.local[:nbytes].view(x.dtype).view(xc.shape))


def _cuda_all_gather(x, y):
    from ...ops import collective as coll
    comm = _comm()
    if comm.cross:
        h = x.detach().contiguous().cpu()
        out = torch.empty([kfpy.current_cluster_size()] + list(h.shape), dtype=h.dtype)
        _check(H.lib().kf_all_gather(kfpy.peer()._h, h.data_ptr(), out.data_ptr(), h.numel(), _host_dtype(h),
                                     b'kungfu::xhost-allgather'), 'all_gather')
        y.copy_(out)
        return
    if comm.np == 1:
        y.copy_(x.unsqueeze(0))
        return
    xc = x.contiguous()
    nbytes = xc.numel() * xc.element_size()
    slot = (nbytes + 15) // 16 * 16
    src = _stage(comm, 'in', slot)
    dst = _stage(comm, 'out', slot * comm.np)
    src.local[:nbytes].copy_(xc.view(-1).view(torch.uint8))
    coll.all_gather(comm, src, dst, slot)
    out = dst.local[:slot * comm.np].view(comm.np, slot)[:, :nbytes].contiguous()
    y.copy_(out.view(-1).view(x.dtype).view(y.shape))


# ---- public API ------------------------------------------------------------------------------
def inplace_all_reduce_op(x, op=None, name=None):
    """All-reduce ``x`` in place (CPU tensors: host runtime; CUDA tensors: peer-memory / NVLS kernels). Returns ``x``."""
    if op is None:
        op = 'sum'
    if x.is_cuda:
        _cuda_all_reduce(x, x, op)
        return
    if not x.is_contiguous():
        raise ValueError('in-place all-reduce needs a contiguous tensor')
    _check(H.lib().kf_all_reduce(kfpy.peer()._h, x.data_ptr(), x.data_ptr(), x.numel(),
                                 _host_dtype(x), H.OPS[op], _name(name, 'all_reduce')), 'all_reduce')


def all_reduce_fn(x, op=None, name=None):
    """Returns op-reduction of x over all workers (sum|min|max|prod)."""
    if op is None:
        op = 'sum'
    y = torch.empty_like(x, memory_format=torch.contiguous_format)
    if x.is_cuda:
        _cuda_all_reduce(x, y, op)
        return y
    xc = x.contiguous()
    _check(H.lib().kf_all_reduce(kfpy.peer()._h, xc.data_ptr(), y.data_ptr(), xc.numel(),
                                 _host_dtype(x), H.OPS[op], _name(name, 'all_reduce')), 'all_reduce')
    return y


def _new_handle(entry):
    with _lock:
        h = _next_handle[0]
        _next_handle[0] += 1
        _handles[h] = entry
    return h


def _stream_for(comm):
    s = _comm_stream.get(id(comm))
    if s is None:
        s = torch.cuda.Stream(device=comm.device, priority=-1)
        _comm_stream[id(comm)] = s
    return s


def inplace_all_reduce_async_op(x, name=None, op=None):
    """Starts an in-place all-reduce; returns a handle for wait_handle()."""
    if op is None:
        op = 'sum'
    if x.is_cuda:
        comm = _comm()
        s = _stream_for(comm)
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            _cuda_all_reduce(x, x, op, channel=ASYNC_CHANNEL)
            ev = torch.cuda.Event()
            ev.record(s)
        x.record_stream(s)
        return _new_handle(('cuda', ev))
    if not x.is_contiguous():
        raise ValueError('in-place all-reduce needs a contiguous tensor')
    h = H.lib().kf_all_reduce_async(kfpy.peer()._h, x.data_ptr(), x.data_ptr(), x.numel(),
                                    _host_dtype(x), H.OPS[op], _name(name, 'all_reduce', True))
    if h < 0:
        _check(-1, 'all_reduce_async')
    return _new_handle(('cpu', h, x))


def inplace_broadcast_async_op(x, name=None):
    """Starts an in-place broadcast from rank 0; returns a handle."""
    if x.is_cuda:
        comm = _comm()
        s = _stream_for(comm)
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            _cuda_broadcast(x, channel=ASYNC_CHANNEL)
            ev = torch.cuda.Event()
            ev.record(s)
        x.record_stream(s)
        return _new_handle(('cuda', ev))
    if not x.is_contiguous():
        raise ValueError('in-place broadcast needs a contiguous tensor')
    h = H.lib().kf_broadcast_async(kfpy.peer()._h, x.data_ptr(), x.data_ptr(), x.numel(),
                                   _host_dtype(x), _name(name, 'broadcast', True))
    if h < 0:
        _check(-1, 'broadcast_async')
    return _new_handle(('cpu', h, x))


def wait_handle(handle):
    """Block until the asynchronous op behind ``handle`` has finished."""
    with _lock:
        entry = _handles.pop(handle)
    if entry[0] == 'cuda':
        torch.cuda.current_stream().wait_event(entry[1])
    else:
        _check(H.lib().kf_wait_handle(kfpy.peer()._h, entry[1]), 'async collective')


def wait_all_handles(handles):
    """Block until every handle in ``handles`` has finished."""
    for h in handles:
        wait_handle(h)


def broadcast_parameters(state_dict):
    """Every tensor of state_dict := rank 0's (in place)."""
    handles = []
    for name, value in state_dict.items():
        if not torch.is_tensor(value):
            continue
        if not value.is_contiguous() or (value.dim() == 0 and not value.is_cuda):
            tmp = value.detach().clone().contiguous().view(-1)
            wait_handle(inplace_broadcast_async_op(tmp, name))
            value.copy_(tmp.view(value.shape))
            continue
        handles.append(inplace_broadcast_async_op(value.detach(), name))
    wait_all_handles(handles)


def all_gather(x, name=None):
    """Stack of every worker's x, shape [np] + x.shape."""
    n = kfpy.current_cluster_size()
    y = x.new_empty([n] + list(x.shape))
    if x.is_cuda:
        _cuda_all_gather(x, y)
        return y
    xc = x.contiguous()
    _check(H.lib().kf_all_gather(kfpy.peer()._h, xc.data_ptr(), y.data_ptr(), xc.numel(),
                                 _host_dtype(x), _name(name, 'all_gather')), 'all_gather')
    return y


def broadcast(x, name=None):
    """In-place broadcast from rank 0 (synchronous)."""
    wait_handle(inplace_broadcast_async_op(x, name))
    return x


def barrier():
    """Block until every worker has called it (CPU control plane)."""
    kfpy.run_barrier()


def _cuda_consensus(x):
    from ...ops import collective as coll
    comm = _comm()
    if comm.np == 1:
        return True
    xc = x.detach().contiguous()
    nbytes = xc.numel() * xc.element_size()
    pad = (nbytes + 15) // 16 * 16
    src = _stage(comm, 'in', pad)
    src.local[:pad].zero_()
    src.local[:nbytes].copy_(xc.view(-1).view(torch.uint8))
    agree = torch.zeros(1, dtype=torch.int32, device=x.device)
    coll.consensus(comm, src, pad, agree)
    return bool(int(agree))


def consensus(x, name=None):
    """True iff every worker holds identical bytes in x (CUDA tensors on one NVLink domain: one
    device kernel compares them in place, nothing is copied to the host)."""
    if x.is_cuda and not _comm().cross:
        # shapes first (control plane, a few bytes), then the bytes on the device
        meta = torch.tensor(list(x.shape) + [x.element_size(), x.dim()], dtype=torch.int64)
        if not consensus(meta, name=(name or 'kungfu::consensus') + '::shape'):
            return False
        return _cuda_consensus(x)
    xc = to_host(x.detach().contiguous())
    agree = C.c_int(0)
    _check(H.lib().kf_consensus(kfpy.peer()._h, xc.data_ptr(), xc.numel() * xc.element_size(),
                                _name(name, 'consensus'), C.byref(agree)), 'consensus')
    return bool(agree.value)


def _cuda_monitored_all_reduce(x, tree, op):
    """The strategy walk on the device: the graphs active in the runtime (set by -strategy, set_tree or
    the `tree` argument) are walked by ONE kernel over peer memory, timed with CUDA events, and the
    bytes / device time land in the same statistics calc_stats() and check_interference() read."""
    from ...parallel.planner import planner_for
    comm = _comm()
    if tree is not None and len(tree) > 0:
        kfpy.set_tree([int(t) for t in tree])
    if comm.np == 1:
        return x.clone()
    from ...ops import _cuda as K
    xc = x.contiguous()
    nbytes = xc.numel() * xc.element_size()
    pad = (nbytes + 15) // 16 * 16
    buf = _stage(comm, 'in', pad)
    buf.local[:nbytes].copy_(xc.view(-1).view(torch.uint8))
    planner_for(comm).monitored_all_reduce(buf, xc.numel(), K.dtype_code(x.dtype), op)
    return buf.local[:nbytes].view(x.dtype).view(xc.shape).clone()


def monitored_all_reduce(x, tree=None, op=None, name=None):
    """All-reduce that feeds the strategy throughput stats (CPU tensors: host runtime; CUDA tensors:
    device graph executor);
    `tree` (father array) installs a new tree first
    (/root/reference/srcs/python/kungfu/tensorflow/ops/collective.py:40-60)."""
    if op is None:
        op = 'sum'
    if x.is_cuda and not _comm().cross:
        return _cuda_monitored_all_reduce(x, tree, op)
    xc = to_host(x.contiguous())
    y = torch.empty_like(xc)
    arr, n = None, 0
    if tree is not None and len(tree) > 0:
        n = len(tree)
        arr = (C.c_int32 * n)(*[int(t) for t in tree])
    _check(H.lib().kf_monitored_all_reduce(kfpy.peer()._h, xc.data_ptr(), y.data_ptr(), xc.numel(),
                                           _host_dtype(xc), H.OPS[op], arr, n,
                                           _name(name, 'monitored_all_reduce')), 'monitored_all_reduce')
    return y.to(x.device)


def to_host(x):
    """Host copy of a tensor (the one place where the host paths stage device data)."""
    return x if not x.is_cuda else x.to('cpu')


def save_variable(x, name, version=None):
    """Publish a tensor in this worker's p2p store (host path)."""
    xc = to_host(x.detach().contiguous())
    kfpy.peer().save(name, xc.data_ptr(), xc.numel() * xc.element_size(),
                     None if version is None else str(version))


def request_variable(target, name, like, version=None):
    """Fetch `name` from worker `target`'s store; returns (ok, tensor)."""
    # (the store is a host object, like the reference's; models that live on GPUs exchange snapshots
    # through PairAveragingOptimizer's device path instead)
    out = torch.empty(like.shape, dtype=like.dtype)
    ok = kfpy.peer().request(int(target), name, out.data_ptr(), out.numel() * out.element_size(),
                             None if version is None else str(version))
    return ok, out.to(like.device)
