"""Device communicators over symmetric (peer-mapped) memory.

Three flavours share one interface:

* ``LocalComm``     np == 1, plain device memory (kernels skip the barriers).
* ``EmulatedComm``  np "ranks" on ONE GPU: every collective is a single
                    cooperative launch whose blockIdx.y plays the rank. Used
                    by GPU tests when fewer GPUs than ranks are available
                    (rank-waiting kernels must never be run as separate
                    launches on one GPU).
* ``PeerComm``      one process per GPU; buffers are CUDA VMM allocations whose
                    POSIX fds are passed over the control plane's unix sockets
                    and mapped by every peer (CUDA-IPC handles as fallback),
                    plus an optional NVLS multicast alias.

This replaces the reference's NCCLController/NCCLHelper/CudaStream stack
(/root/reference/srcs/cpp/src/nccl/controller.cpp:8-53,
/root/reference/srcs/cpp/src/nccl/gpu_collective.cpp:59-152): there is no
communicator library underneath, only mapped pointers and flags.
"""
import ctypes as C
import os

import torch

from ..ops import _cuda as K


def _align(n, a):
    return (n + a - 1) // a * a


class SymmBuffer(object):
    """A buffer that exists at the same size on every rank.

    ``local``  uint8 tensor aliasing this rank's copy (emulation: rank 0's)
    ``ptrs``   device addresses of every rank's copy as mapped HERE
    ``mc_ptr`` multicast alias (NVLS) or None
    """

    def __init__(self, comm, nbytes, locals_, ptrs, mc_ptr=None, handles=None):
        self.comm = comm
        self.nbytes = nbytes
        self.locals = locals_  # list of uint8 tensors (len np for emulation, else 1)
        self.local = locals_[0]
        self.ptrs = ptrs
        self.mc_ptr = mc_ptr
        self._handles = handles or []
        self._mc = None  # multicast object handle (PeerComm._bind_multicast)

    def view(self, dtype, rank=None):
        """Typed tensor over this rank's copy (or emulated rank `rank`)."""
        t = self.local if rank is None else self.locals[rank]
        return t.view(dtype)

    def ptr_array(self, byte_offset=0):
        return K.ptr_array([p + byte_offset for p in self.ptrs])

    def release(self):
        """Unmap and release this buffer: the multicast alias, every peer's
        import and this rank's own physical block. The caller guarantees that no
        kernel still touches it (PeerComm.close synchronises first)."""
        lib = K.lib()
        if self._mc is not None:
            lib.kfb_mc_free(self._mc)
            self._mc = None
            self.mc_ptr = None
        for h in self._handles:
            if isinstance(h, tuple):
                kind, ptr = h
                if kind == 'ipc':
                    lib.kfb_ipc_free(C.c_void_p(ptr))
                elif kind == 'ipc-peer':
                    lib.kfb_ipc_close(C.c_void_p(ptr))
            elif h is not None:
                lib.kfb_symm_free(h)
        self._handles = []
        self.ptrs = []
        self.locals = []
        self.local = None


class _CommBase(object):
    rank = 0
    np = 1
    emulated = False
    cross = False        # True: more hosts exist; collectives need a cross-host leg
    global_np = None     # job-wide worker count when cross
    is_local_root = True

    def __init__(self, device):
        self.device = torch.device(device)
        self._struct = K.KfbComm()
        self._keep = []
        self._channels = {}
        self.closed = False
        from .devstats import DeviceStats
        self.stats = DeviceStats(self)  # bytes / device time per kernel variant (calc_stats, egress_rates, ...)
        self.planner = None             # parallel.planner.DevicePlanner, attached on first use
        self.default_blocks = min(64, int(os.environ.get('KUNGFU_COMM_BLOCKS', '32')))
        self.default_threads = int(os.environ.get('KUNGFU_COMM_THREADS', '512'))

    def _init_struct(self, rank, np, pad_ptrs, n_state):
        s = self._struct
        s.rank = rank
        s.np = np
        for r in range(K.MAX_RANKS):
            s.pads[r] = pad_ptrs[r] if r < len(pad_ptrs) else None
        self._epoch = torch.zeros(K.STATE_EPOCH_WORDS * n_state, dtype=torch.int32, device=self.device)
        self._scratch = torch.zeros(K.SCRATCH_FLOATS * n_state, dtype=torch.float32, device=self.device)
        self._fuse_scratch = torch.zeros(K.FUSE_SCRATCH_FLOATS, dtype=torch.float32, device=self.device)
        s.local_epoch = self._epoch.data_ptr()
        s.scratch = self._scratch.data_ptr()
        s.mc_pad = None

    @property
    def cstruct(self):
        return C.byref(self._struct)

    # Collectives keep their barrier epochs, flags and reduction scratch per
    # CTA index, so two kernels of one communicator may only be in flight at the
    # same time if they use disjoint CTA rows. A channel is such a slice: the
    # caller's stream uses channel 0, the asynchronous torch ops channel 1, the
    # gradient buckets that overlap backward channel 2. Kernels of ONE channel
    # must be stream-ordered.
    CHANNELS = {0: (0, 64), 1: (64, 32), 2: (96, 32)}  # channel -> (first CTA row, rows)

    def channel(self, ch):
        """ctypes reference to the communicator as seen by channel `ch`."""
        if ch == 0:
            return self.cstruct
        cs = self._channels.get(ch)
        if cs is None:
            base, _ = self.CHANNELS[ch]
            s = K.KfbComm()
            s.rank, s.np = self._struct.rank, self._struct.np
            for r in range(K.MAX_RANKS):
                p = self._struct.pads[r]
                s.pads[r] = (p + base * K.MAX_RANKS * 4) if p else None
            s.local_epoch = self._struct.local_epoch + base * 4
            scratch = torch.zeros(K.SCRATCH_FLOATS * max(1, self.n_local_ranks), dtype=torch.float32,
                                  device=self.device)
            s.scratch = scratch.data_ptr()
            s.mc_pad = None
            cs = (s, scratch)
            self._channels[ch] = cs
        return C.byref(cs[0])

    def channel_blocks(self, ch, want=None):
        return max(1, min(want or self.default_blocks, self.CHANNELS[ch][1]))

    @property
    def n_local_ranks(self):
        """How many ranks' worth of state live in this process (np when emulated)."""
        return self.np if self.emulated else 1

    def alloc(self, nbytes, multicast=False):
        raise NotImplementedError

    def barrier(self, stream=None):
        if self.np > 1:
            K.call('kfb_barrier', self.cstruct, K.stream_ptr(stream))

    def close(self, collective=True):
        self.closed = True
        self._keep = []
        self._channels = {}


class LocalComm(_CommBase):
    def __init__(self, device='cuda'):
        super().__init__(device)
        self._pad = torch.zeros(K.PAD_BYTES, dtype=torch.uint8, device=self.device)
        self._init_struct(0, 1, [self._pad.data_ptr()], 1)

    def alloc(self, nbytes, multicast=False):
        nbytes = _align(nbytes, 256)
        t = torch.zeros(nbytes, dtype=torch.uint8, device=self.device)
        return SymmBuffer(self, nbytes, [t], [t.data_ptr()])


class EmulatedComm(_CommBase):
    emulated = True

    def __init__(self, np, device='cuda'):
        super().__init__(device)
        assert 1 <= np <= K.MAX_RANKS
        self.np = np
        self.rank = -1
        self._pads = [torch.zeros(K.PAD_BYTES, dtype=torch.uint8, device=self.device) for _ in range(np)]
        self._init_struct(-1, np, [p.data_ptr() for p in self._pads], np)
        # every (block, rank) CTA must be co-resident
        sms = torch.cuda.get_device_properties(self.device).multi_processor_count
        self.default_blocks = max(1, min(self.default_blocks, sms // np))

    def alloc(self, nbytes, multicast=False):
        nbytes = _align(nbytes, 256)
        ts = [torch.zeros(nbytes, dtype=torch.uint8, device=self.device) for _ in range(self.np)]
        return SymmBuffer(self, nbytes, ts, [t.data_ptr() for t in ts])


class PeerComm(_CommBase):
    """Real multi-GPU communicator. `xchg` is the control-plane exchange:

        xchg.rank, xchg.size
        xchg.all_gather_bytes(b: bytes) -> [bytes] * size
        xchg.exchange_fds(fd: int, tag: str) -> [fd or None] * size   (own slot None)
        xchg.barrier()
    """

    def __init__(self, xchg, device, mode=None):
        super().__init__(device)
        self.xchg = xchg
        self.rank = xchg.rank
        self.np = xchg.size
        assert self.np <= K.MAX_RANKS, 'one NVSwitch domain holds at most %d ranks' % K.MAX_RANKS
        self.mode = mode or os.environ.get('KUNGFU_SYMM_MODE', 'vmm')
        self._seq = 0
        self.multicast_ok = False
        if os.environ.get('KUNGFU_DISABLE_NVLS', '0') != '1' and self.mode == 'vmm':
            # (always exchanged, even by a lone worker of a host: the exchange is job-collective)
            ok = K.lib().kfb_multicast_supported(self.device.index or 0)
            oks = xchg.all_gather_bytes(bytes([1 if ok else 0]))
            self.multicast_ok = self.np > 1 and all(b == b'\x01' for b in oks)
        self._pad_buf = self.alloc(K.PAD_BYTES)
        self._init_struct(self.rank, self.np, self._pad_buf.ptrs, 1)

    def _alloc_vmm(self, nbytes):
        lib = K.lib()
        dev = self.device.index or 0
        h = C.c_void_p()
        K.check(lib.kfb_symm_alloc(dev, nbytes, C.byref(h)))
        size = lib.kfb_symm_size(h)
        fd = C.c_int(-1)
        K.check(lib.kfb_symm_export_fd(h, C.byref(fd)))
        self._seq += 1
        fds = self.xchg.exchange_fds(fd.value, 'symm%d' % self._seq)
        handles, ptrs = [], []
        for r in range(self.np):
            if r == self.rank:
                handles.append(h)
                ptrs.append(lib.kfb_symm_ptr(h))
            else:
                ph = C.c_void_p()
                K.check(lib.kfb_symm_import_fd(dev, fds[r], size, C.byref(ph)))
                lib.kfb_close_fd(fds[r])
                handles.append(ph)
                ptrs.append(lib.kfb_symm_ptr(ph))
        lib.kfb_close_fd(fd.value)
        return size, handles, ptrs

    def _alloc_ipc(self, nbytes):
        lib = K.lib()
        dev = self.device.index or 0
        p = C.c_void_p()
        hb = C.create_string_buffer(64)
        K.check(lib.kfb_ipc_alloc(dev, nbytes, C.byref(p), hb))
        all_h = self.xchg.all_gather_bytes(hb.raw)
        ptrs = []
        for r in range(self.np):
            if r == self.rank:
                ptrs.append(p.value)
            else:
                q = C.c_void_p()
                K.check(lib.kfb_ipc_open(dev, all_h[r], C.byref(q)))
                ptrs.append(q.value)
        return nbytes, [('ipc', p.value)] + [('ipc-peer', q) for r, q in enumerate(ptrs) if r != self.rank], ptrs

    def alloc(self, nbytes, multicast=False):
        """Collective: every rank must call with the same size, in the same order."""
        gran = 2 << 20
        nbytes = _align(max(nbytes, 1), gran)
        # collective-call hygiene: a mismatch here would otherwise surface as an
        # obscure driver error when a peer maps a block of another size
        want = self.xchg.all_gather_bytes(int(nbytes).to_bytes(8, 'little'))
        if any(w != want[0] for w in want):
            raise K.KungFuCudaError('symmetric alloc mismatch: ranks asked for %s bytes; allocations are collective '
                                    'and must be issued in the same order on every rank' %
                                    [int.from_bytes(w, 'little') for w in want])
        if self.mode == 'vmm':
            size, handles, ptrs = self._alloc_vmm(nbytes)
        else:
            size, handles, ptrs = self._alloc_ipc(nbytes)
        mc_ptr, mc_obj = None, None
        if multicast and self.multicast_ok:
            mc_ptr, mc_obj = self._bind_multicast(size, handles[self.rank])
        elif multicast and self.np > 1 and require_nvls():
            raise K.KungFuCudaError('KUNGFU_REQUIRE_NVLS=1 but no multicast (NVLS) alias can be bound on this '
                                    'machine / in this mode (%s)' % self.mode)
        local = K.tensor_from_ptr(ptrs[self.rank], size, self.device, owner=None)
        self.xchg.barrier()
        buf = SymmBuffer(self, size, [local], ptrs, mc_ptr, handles)
        buf._mc = mc_obj
        self._keep.append(buf)
        return buf

    def free(self, buf):
        """Collective: release one buffer on every rank (nobody may still use it)."""
        torch.cuda.synchronize(self.device)
        self.xchg.barrier()
        if buf in self._keep:
            self._keep.remove(buf)
        buf.release()

    def _bind_multicast(self, size, my_handle):
        lib = K.lib()
        dev = self.device.index or 0
        mc = C.c_void_p()
        self._seq += 1
        tag = 'mc%d' % self._seq
        try:
            if self.rank == 0:
                fd = C.c_int(-1)
                K.check(lib.kfb_mc_create(dev, self.np, size, C.byref(mc), C.byref(fd)))
                self.xchg.broadcast_fd(fd.value, tag, root=0)
                lib.kfb_close_fd(fd.value)
            else:
                rfd = self.xchg.broadcast_fd(None, tag, root=0)
                K.check(lib.kfb_mc_import(dev, self.np, rfd, size, C.byref(mc)))
                lib.kfb_close_fd(rfd)
            K.check(lib.kfb_mc_add_device(mc))
            self.xchg.barrier()
            p = C.c_void_p()
            K.check(lib.kfb_mc_bind(mc, my_handle, C.byref(p)))
            self.xchg.barrier()
            return p.value, mc
        except K.KungFuCudaError:
            # a failed bind must fail on every rank or on none; ranks agree below
            self.multicast_ok = False
            raise

    def close(self, collective=True):
        """Give everything back: unmap the peers' memory (their GPUs can then
        free it), release this rank's blocks and multicast objects. The
        reference dropped the connections to removed peers and destroyed its
        NCCL communicators at this point
        (/root/reference/srcs/go/rchannel/client/connection_pool.go:29-51,
        /root/reference/srcs/cpp/src/nccl/gpu_collective.cpp:84-94). With
        collective=False no peer is consulted (a detached worker on its way
        out, or a membership that no longer exists)."""
        if self.closed:
            return
        self.closed = True
        torch.cuda.synchronize(self.device)
        if collective:
            try:
                self.xchg.barrier()
            except Exception:
                pass
        for buf in self._keep:
            buf.release()
        self._keep = []
        self._channels = {}


def require_nvls():
    """KUNGFU_REQUIRE_NVLS=1: a missing multicast alias is an error instead of a
    silent fall-back to the peer-to-peer kernels (multi-GPU tests set it, so a
    box whose multicast setup fails cannot pass as if NVLS had been tested)."""
    return os.environ.get('KUNGFU_REQUIRE_NVLS', '0') == '1'


_default = None
_retired = []


def default_comm():
    """The process-wide communicator, created on first use from the runtime's
    view of the cluster (np == 1 -> LocalComm)."""
    global _default
    if _default is None:
        from .. import python as kfpy
        dev = torch.device('cuda', torch.cuda.current_device())
        if kfpy.current_cluster_size() <= 1:
            _default = LocalComm(dev)
        elif kfpy.host_count() > 1:
            # several hosts: peer memory only spans this host's GPUs; the
            # cross-host leg goes through the control plane (hierarchical
            # all-reduce, /root/reference/srcs/cpp/src/nccl/controller.cpp:8-40)
            lx = LocalExchange(kfpy.peer())
            _default = PeerComm(lx, dev)  # also for a lone worker: allocation stays job-collective
            _default.cross = True
            _default.global_np = kfpy.current_cluster_size()
            _default.is_local_root = lx.rank == 0
        else:
            _default = PeerComm(kfpy.control_exchange(), dev)
        if isinstance(_default, PeerComm) and not _default.cross:
            # measured bus bandwidth decides which all-reduce variant serves which size; the probe is
            # collective, and creating the communicator is a point every rank passes
            from .planner import planner_for
            planner_for(_default).probe()
    return _default


class LocalExchange(object):
    """The control-plane exchange restricted to the workers of this host.
    Calls are still issued by every worker of the job in lockstep (symmetric
    allocation is collective), so job-wide primitives can be reused and sliced."""

    def __init__(self, peer):
        from .. import _host as H
        self._peer = peer
        self._g = peer.exchange()
        n = peer.size()
        buf = C.create_string_buffer(64)
        hosts = []
        for r in range(n):
            H.lib().kf_peer_addr(peer._h, r, buf, 64)
            hosts.append(buf.value.decode().split(':')[0])
        mine = hosts[peer.rank()]
        self.members = [r for r in range(n) if hosts[r] == mine]  # global ranks on this host
        self.rank = self.members.index(peer.rank())
        self.size = len(self.members)

    def barrier(self):
        # host-local: reduce to the host master, then broadcast back (hosts may
        # take different code paths, e.g. only multi-GPU hosts bind multicast)
        from .. import _host as H
        a, b = C.c_uint8(1), C.c_uint8(0)
        H._check(H.lib().kf_local_reduce(self._peer._h, C.byref(a), C.byref(b), 1, H.U8, 0, b'kungfu::lx-barrier'), 'local barrier')
        H._check(H.lib().kf_local_broadcast(self._peer._h, C.byref(b), C.byref(b), 1, H.U8, b'kungfu::lx-barrier'), 'local barrier')

    def all_gather_bytes(self, b):
        # job-wide gather, sliced: only used at points every worker reaches
        everyone = self._g.all_gather_bytes(b)
        return [everyone[r] for r in self.members]

    def exchange_fds(self, fd, tag):
        from .. import _host as H
        out = [None] * self.size
        for k in range(1, self.size):
            dst = self.members[(self.rank + k) % self.size]
            H._check(H.lib().kf_send_fd(self._peer._h, dst, tag.encode(), fd), 'send_fd')
        for k in range(1, self.size):
            i = (self.rank - k) % self.size
            got = C.c_int(-1)
            H._check(H.lib().kf_recv_fd(self._peer._h, self.members[i], tag.encode(), C.byref(got)), 'recv_fd')
            out[i] = got.value
        return out

    def broadcast_fd(self, fd, tag, root=0):
        from .. import _host as H
        if self.rank == root:
            for i in range(self.size):
                if i != root:
                    H._check(H.lib().kf_send_fd(self._peer._h, self.members[i], tag.encode(), fd), 'send_fd')
            return fd
        got = C.c_int(-1)
        H._check(H.lib().kf_recv_fd(self._peer._h, self.members[root], tag.encode(), C.byref(got)), 'recv_fd')
        return got.value


def set_default_comm(comm):
    global _default
    _default = comm


def reset_default_comm():
    """Retire the communicator (after a resize the next use maps peer memory for
    the new membership; surviving processes are not restarted). Its buffers
    stay mapped until release_retired(): optimizers copy their state out of the
    old arenas into the new ones first."""
    global _default
    if _default is not None:
        _retired.append(_default)
    _default = None


def release_retired():
    """Unmap and free every buffer of communicators retired by a membership
    change. Call once nothing refers to their memory any more (the fused
    optimizers do after rebuild(); ElasticState after the state hand-over)."""
    from ..torch.ops import collective as tops
    while _retired:
        c = _retired.pop()
        tops.drop_staging(c)
        c.close(collective=False)


def shutdown():
    """Release all device communication resources of this process (a detached
    worker calls this before it exits so that the survivors' GPUs get the
    peer mappings of its memory back)."""
    global _default
    release_retired()
    if _default is not None:
        from ..torch.ops import collective as tops
        tops.drop_staging(_default)
        _default.close(collective=False)
        _default = None
