"""ctypes binding of libkungfu_host.so (C++ control plane / planner / elastic).

See include/kungfu_host.h for the C ABI. The reference bound its Go runtime
the same way (/root/reference/srcs/python/kungfu/loader.py:11-14,
/root/reference/srcs/cpp/src/python/init.cpp:10-41).
"""
import ctypes as C
import os
import socket
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'lib', 'libkungfu_host.so')

_lib = None
_lock = threading.Lock()

# KungFu_Datatype-compatible codes (/root/reference/srcs/cpp/include/kungfu/dtype.h:21-39)
def _tc(c, b):
    return (c << 16) | (b << 8) | 8


U8, U16, U32, U64 = _tc(0, 1), _tc(0, 2), _tc(0, 4), _tc(0, 8)
I8, I16, I32, I64 = _tc(1, 1), _tc(1, 2), _tc(1, 4), _tc(1, 8)
F16, F32, F64 = _tc(2, 2), _tc(2, 4), _tc(2, 8)
BOOL = _tc(3, 1)
BF16 = _tc(4, 2)  # extension: not in the reference
OPS = {None: 0, 'sum': 0, 'min': 1, 'max': 2, 'prod': 3}


class KungFuHostError(RuntimeError):
    pass


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise KungFuHostError('%s is missing: run `python -m kungfu_b200._build`' % LIB_PATH)
                l = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
                _declare(l)
                _lib = l
    return _lib


_VP = C.c_void_p


def _declare(l):
    def f(name, res, *args):
        fn = getattr(l, name)
        fn.restype = res
        fn.argtypes = list(args)

    f('kf_last_error', C.c_char_p)
    f('kf_peer_create_from_env', _VP)
    f('kf_peer_start', C.c_int, _VP)
    f('kf_peer_close', None, _VP)
    for n in ('kf_rank', 'kf_size', 'kf_local_rank', 'kf_local_size', 'kf_host_count', 'kf_detached',
              'kf_barrier', 'kf_calc_stats', 'kf_log_stats', 'kf_print_strategy_stats',
              'kf_check_interference', 'kf_cluster_version'):
        f(n, C.c_int, _VP)
    f('kf_uid', C.c_uint64, _VP)
    f('kf_all_reduce', C.c_int, _VP, _VP, _VP, C.c_int64, C.c_int, C.c_int, C.c_char_p)
    f('kf_monitored_all_reduce', C.c_int, _VP, _VP, _VP, C.c_int64, C.c_int, C.c_int, C.POINTER(C.c_int32), C.c_int, C.c_char_p)
    f('kf_reduce', C.c_int, _VP, _VP, _VP, C.c_int64, C.c_int, C.c_int, C.c_char_p)
    f('kf_broadcast', C.c_int, _VP, _VP, _VP, C.c_int64, C.c_int, C.c_char_p)
    f('kf_all_gather', C.c_int, _VP, _VP, _VP, C.c_int64, C.c_int, C.c_char_p)
    f('kf_gather', C.c_int, _VP, _VP, _VP, C.c_int64, C.c_int, C.c_char_p)
    f('kf_local_reduce', C.c_int, _VP, _VP, _VP, C.c_int64, C.c_int, C.c_int, C.c_char_p)
    f('kf_local_broadcast', C.c_int, _VP, _VP, _VP, C.c_int64, C.c_int, C.c_char_p)
    f('kf_cross_all_reduce', C.c_int, _VP, _VP, _VP, C.c_int64, C.c_int, C.c_int, C.c_char_p)
    f('kf_consensus', C.c_int, _VP, _VP, C.c_int64, C.c_char_p, C.POINTER(C.c_int))
    f('kf_all_reduce_async', C.c_int64, _VP, _VP, _VP, C.c_int64, C.c_int, C.c_int, C.c_char_p)
    f('kf_broadcast_async', C.c_int64, _VP, _VP, _VP, C.c_int64, C.c_int, C.c_char_p)
    f('kf_wait_handle', C.c_int, _VP, C.c_int64)
    f('kf_save', C.c_int, _VP, C.c_char_p, C.c_char_p, _VP, C.c_int64)
    f('kf_request', C.c_int, _VP, C.c_int, C.c_char_p, C.c_char_p, _VP, C.c_int64)
    f('kf_request_async', C.c_int64, _VP, C.c_int, C.c_char_p, C.c_char_p, _VP, C.c_int64)
    f('kf_prefetch_new', _VP, _VP, C.c_char_p, C.c_int64)
    f('kf_prefetch_next', C.c_int, _VP, C.c_int, _VP, C.POINTER(C.c_int64))
    f('kf_prefetch_free', None, _VP)
    f('kf_save_parts', C.c_int, _VP, C.c_char_p, C.c_char_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int64), C.c_int)
    f('kf_blend_f32', None, C.POINTER(C.c_void_p), C.POINTER(C.c_int64), C.c_int, _VP)
    f('kf_prefetch_blend_f32', C.c_int, _VP, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int64), C.c_int, C.POINTER(C.c_int64))
    f('kf_resize_cluster', C.c_int, _VP, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int))
    f('kf_resize_cluster_from_url', C.c_int, _VP, C.POINTER(C.c_int), C.POINTER(C.c_int))
    f('kf_peek_cluster_change', C.c_int, _VP, C.POINTER(C.c_int))
    f('kf_propose_new_size', C.c_int, _VP, C.c_int)
    f('kf_set_tree', C.c_int, _VP, C.POINTER(C.c_int32), C.c_int)
    f('kf_set_strategy', C.c_int, _VP, C.c_char_p)
    f('kf_current_strategies', C.c_int, _VP, C.c_char_p, C.c_int)
    f('kf_device_plan', C.c_int, _VP, C.POINTER(C.c_int32), C.c_int)
    f('kf_device_plan_for', C.c_int, C.c_char_p, C.c_char_p, C.c_int, C.POINTER(C.c_int32), C.c_int)
    f('kf_current_strategy_name', C.c_int, _VP, C.c_char_p, C.c_int)
    f('kf_get_peer_latencies', C.c_int, _VP, C.POINTER(C.c_float), C.c_int)
    f('kf_egress_rates', C.c_int, _VP, C.POINTER(C.c_float), C.c_int)
    f('kf_send_fd', C.c_int, _VP, C.c_int, C.c_char_p, C.c_int)
    f('kf_recv_fd', C.c_int, _VP, C.c_int, C.c_char_p, C.POINTER(C.c_int))
    f('kf_peer_addr', C.c_int, _VP, C.c_int, C.c_char_p, C.c_int)


def _check(rc, what):
    if rc != 0:
        raise KungFuHostError('%s failed: %s' % (what, lib().kf_last_error().decode()))


def _find_free_port_block(n):
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _adopt_torchrun_env():
    """torchrun (bench.py's launcher) sets RANK/WORLD_SIZE/MASTER_PORT rather
    than the KUNGFU_* contract: derive an equivalent single-host peer list."""
    if 'KUNGFU_SELF_SPEC' in os.environ:
        return
    ws = int(os.environ.get('WORLD_SIZE', '1'))
    if ws <= 1 or 'RANK' not in os.environ:
        return
    rank = int(os.environ['RANK'])
    base = int(os.environ.get('KUNGFU_PORT_BASE', str(20000 + (int(os.environ.get('MASTER_PORT', '29500')) * 7) % 12000)))  # below the ephemeral range
    peers = ','.join('127.0.0.1:%d' % (base + r) for r in range(ws))
    os.environ['KUNGFU_SELF_SPEC'] = '127.0.0.1:%d' % (base + rank)
    os.environ['KUNGFU_INIT_PEERS'] = peers
    os.environ.setdefault('KUNGFU_INIT_CLUSTER_VERSION', '0')
    os.environ.setdefault('KUNGFU_ALLREDUCE_STRATEGY', 'AUTO')


def device_plan_table(v):
    """Unpack {n_plans, n_rounds, n_reduce_rounds, n_edges, round_begin[], edge_plan[], edge_src[]}
    (kf_device_plan, include/kungfu_host.h) into the table the device executor walks."""
    n_plans, n_rounds, n_reduce, n_edges = v[:4]
    begin = v[4:4 + n_rounds + 1]
    plan = v[4 + n_rounds + 1:4 + n_rounds + 1 + n_edges]
    src = v[4 + n_rounds + 1 + n_edges:4 + n_rounds + 1 + 2 * n_edges]
    rounds = [[(plan[e], src[e]) for e in range(begin[r], begin[r + 1])] for r in range(n_rounds)]
    return {'n_plans': n_plans, 'rounds': rounds, 'n_reduce': n_reduce, 'begin': list(begin), 'plan': list(plan),
            'src': list(src)}


def device_plan_for(peers_csv, strategy, rank):
    """Compile a named strategy over an explicit peer list for `rank` (tests, tools)."""
    buf = (C.c_int32 * 8192)()
    n = lib().kf_device_plan_for(peers_csv.encode(), strategy.encode(), int(rank), buf, len(buf))
    if n < 0:
        raise KungFuHostError('device_plan_for: %s' % lib().kf_last_error().decode())
    return device_plan_table(list(buf[:n]))


class Exchange(object):
    """What the device layer needs from the control plane."""

    def __init__(self, peer):
        self._peer = peer
        self.rank = peer.rank()
        self.size = peer.size()
        self._n = 0

    def barrier(self):
        self._peer.barrier()

    def all_gather_bytes(self, b):
        n = len(b)
        self._n += 1
        send = C.create_string_buffer(bytes(b), n)
        recv = C.create_string_buffer(n * self.size)
        _check(lib().kf_all_gather(self._peer._h, send, recv, n, U8, ('xchg::ag%d' % self._n).encode()), 'all_gather')
        raw = recv.raw
        return [raw[i * n:(i + 1) * n] for i in range(self.size)]

    def exchange_fds(self, fd, tag):
        """Everyone sends its fd to everyone (unix sockets, SCM_RIGHTS)."""
        out = [None] * self.size
        for k in range(1, self.size):
            dst = (self.rank + k) % self.size
            _check(lib().kf_send_fd(self._peer._h, dst, tag.encode(), fd), 'send_fd')
        for k in range(1, self.size):
            src = (self.rank - k) % self.size
            got = C.c_int(-1)
            _check(lib().kf_recv_fd(self._peer._h, src, tag.encode(), C.byref(got)), 'recv_fd')
            out[src] = got.value
        return out

    def broadcast_fd(self, fd, tag, root=0):
        if self.rank == root:
            for r in range(self.size):
                if r != root:
                    _check(lib().kf_send_fd(self._peer._h, r, tag.encode(), fd), 'send_fd')
            return fd
        got = C.c_int(-1)
        _check(lib().kf_recv_fd(self._peer._h, root, tag.encode(), C.byref(got)), 'recv_fd')
        return got.value


class Peer(object):
    """One per process. Thin handle over the C++ kungfu::Peer equivalent."""

    def __init__(self, handle):
        self._h = handle
        self._xchg = None
        self._xchg_version = None

    @classmethod
    def from_env(cls):
        _adopt_torchrun_env()
        l = lib()
        h = l.kf_peer_create_from_env()
        if not h:
            raise KungFuHostError('cannot create peer: %s' % l.kf_last_error().decode())
        _check(l.kf_peer_start(h), 'peer start')
        return cls(h)

    def close(self):
        if self._h:
            lib().kf_peer_close(self._h)
            self._h = None

    def rank(self):
        return lib().kf_rank(self._h)

    def size(self):
        return lib().kf_size(self._h)

    def local_rank(self):
        return lib().kf_local_rank(self._h)

    def local_size(self):
        return lib().kf_local_size(self._h)

    def host_count(self):
        return lib().kf_host_count(self._h)

    def uid(self):
        return lib().kf_uid(self._h)

    def detached(self):
        return bool(lib().kf_detached(self._h))

    def cluster_version(self):
        return lib().kf_cluster_version(self._h)

    def barrier(self):
        _check(lib().kf_barrier(self._h), 'barrier')

    def exchange(self):
        # One Exchange per MEMBERSHIP: its message names are numbered, and a joiner starts counting
        # at 1, so a survivor must start over too. Size and rank alone do not identify a membership
        # (1 -> 2 -> 1 -> 2 comes back to the same size and rank with a new peer): the cluster
        # version does.
        v = self.cluster_version()
        if (self._xchg is None or self._xchg_version != v or self._xchg.size != self.size()
                or self._xchg.rank != self.rank()):
            self._xchg = Exchange(self)
            self._xchg_version = v
        return self._xchg

    # -- elastic ----------------------------------------------------------------------
    def resize_cluster(self, n):
        ch, det = C.c_int(0), C.c_int(0)
        _check(lib().kf_resize_cluster(self._h, n, C.byref(ch), C.byref(det)), 'resize_cluster')
        return bool(ch.value), bool(det.value)

    def peek_cluster_change(self):
        d = C.c_int(0)
        _check(lib().kf_peek_cluster_change(self._h, C.byref(d)), 'peek_cluster_change')
        return bool(d.value)

    def resize_cluster_from_url(self):
        ch, det = C.c_int(0), C.c_int(0)
        _check(lib().kf_resize_cluster_from_url(self._h, C.byref(ch), C.byref(det)), 'resize_cluster_from_url')
        return bool(ch.value), bool(det.value)

    def propose_new_size(self, n):
        _check(lib().kf_propose_new_size(self._h, n), 'propose_new_size')

    # -- adaptation / monitoring ---------------------------------------------------
    def check_interference(self):
        return bool(lib().kf_check_interference(self._h))

    def calc_stats(self):
        lib().kf_calc_stats(self._h)

    def log_stats(self):
        lib().kf_log_stats(self._h)

    def print_strategy_stats(self):
        lib().kf_print_strategy_stats(self._h)

    def set_tree(self, forest):
        arr = (C.c_int32 * len(forest))(*forest)
        _check(lib().kf_set_tree(self._h, arr, len(forest)), 'set_tree')

    def set_strategy(self, name):
        _check(lib().kf_set_strategy(self._h, name.encode()), 'set_strategy')

    def strategies(self):
        """Active (reduce, bcast) graph pairs: list of dicts with 'loops' (set of
        ranks that reduce), 'reduce' and 'bcast' edge lists."""
        import re
        buf = C.create_string_buffer(1 << 20)
        if lib().kf_current_strategies(self._h, buf, len(buf)) < 0:
            raise KungFuHostError(lib().kf_last_error().decode())
        out = []
        for line in buf.value.decode().strip().split('\n'):
            red, bc = line.split(' bcast=')
            edges = lambda t: [(int(a), int(b)) for a, b in re.findall(r'\((\d+)->(\d+)\)', t)]
            out.append({'loops': set(int(a) for a in re.findall(r'\((\d+)\)', red)),
                        'reduce': edges(red), 'bcast': edges(bc)})
        return out

    def device_plan(self):
        """This rank's compiled edge table of the active strategies (see device_plan_table)."""
        buf = (C.c_int32 * 8192)()
        n = lib().kf_device_plan(self._h, buf, len(buf))
        if n < 0:
            raise KungFuHostError('device_plan: %s' % lib().kf_last_error().decode())
        return device_plan_table(list(buf[:n]))

    def strategy_name(self):
        buf = C.create_string_buffer(64)
        lib().kf_current_strategy_name(self._h, buf, 64)
        return buf.value.decode()

    def peer_latencies(self):
        n = self.size()
        arr = (C.c_float * n)()
        _check(lib().kf_get_peer_latencies(self._h, arr, n), 'get_peer_latencies')
        return list(arr)

    def egress_rates(self):
        n = self.size()
        arr = (C.c_float * n)()
        _check(lib().kf_egress_rates(self._h, arr, n), 'egress_rates')
        return list(arr)

    # -- p2p store -------------------------------------------------------------------
    def save(self, name, buf_ptr, nbytes, version=None):
        v = version.encode() if version is not None else None
        _check(lib().kf_save(self._h, v, name.encode(), buf_ptr, nbytes), 'save')

    def request(self, target, name, buf_ptr, nbytes, version=None):
        v = version.encode() if version is not None else None
        rc = lib().kf_request(self._h, target, v, name.encode(), buf_ptr, nbytes)
        return rc == 0

    def request_async(self, target, name, buf_ptr, nbytes, version=None):
        """Start a request on the runtime's workers; `wait_handle(h)` is True iff the blob arrived.
        The caller keeps the buffer alive until then."""
        v = version.encode() if version is not None else None
        h = lib().kf_request_async(self._h, target, v, name.encode(), buf_ptr, nbytes)
        if h < 0:
            raise KungFuHostError('request_async failed: %s' % lib().kf_last_error().decode())
        return h

    def wait(self, handle):
        """Block until an async op finishes; True iff it succeeded."""
        return lib().kf_wait_handle(self._h, handle) == 0

    def prefetch_new(self, name, nbytes):
        pf = lib().kf_prefetch_new(self._h, name.encode(), nbytes)
        if not pf:
            raise KungFuHostError('prefetch_new failed: %s' % lib().kf_last_error().decode())
        return pf

    @staticmethod
    def prefetch_next(pf, target, buf_ptr):
        """(ok, completed background pulls so far); see kf_prefetch_next in include/kungfu_host.h."""
        fresh = C.c_int64(0)
        rc = lib().kf_prefetch_next(pf, target, buf_ptr, C.byref(fresh))
        return rc == 0, int(fresh.value)

    @staticmethod
    def prefetch_free(pf):
        lib().kf_prefetch_free(pf)

    @staticmethod
    def _parts(ptrs, counts):
        n = len(ptrs)
        return (C.c_void_p * n)(*ptrs), (C.c_int64 * n)(*counts), n

    def save_parts(self, name, ptrs, nbytes, version=None):
        """Publish the concatenation of several byte ranges as one blob (gathered natively, one copy)."""
        pa, ca, n = self._parts(ptrs, nbytes)
        v = version.encode() if version is not None else None
        _check(lib().kf_save_parts(self._h, v, name.encode(), pa, ca, n), 'save_parts')

    @classmethod
    def blend_f32(cls, ptrs, counts, other_ptr):
        """t_i <- 0.5 * (t_i + other[...]) for fp32 tensors given by pointer / element count."""
        pa, ca, n = cls._parts(ptrs, counts)
        lib().kf_blend_f32(pa, ca, n, other_ptr)

    @classmethod
    def prefetch_blend_f32(cls, pf, target, ptrs, counts):
        """prefetch_next fused with the blend: (ok, completed background pulls so far)."""
        pa, ca, n = cls._parts(ptrs, counts)
        fresh = C.c_int64(0)
        rc = lib().kf_prefetch_blend_f32(pf, target, pa, ca, n, C.byref(fresh))
        return rc == 0, int(fresh.value)
