"""Fused optimizers on flat arenas with fp32 master weights.

``FusedSGD`` (torch.optim.SGD semantics), ``FusedAdamW`` and ``FusedAdam``
(torch.optim.AdamW / Adam semantics) share one engine:

* parameters and gradients live in flat symmetric arenas (one per dtype);
* one kernel per (param group, arena[, bucket]) does the whole update; driven
  by ``SynchronousSGDOptimizer`` the same kernel also performs the gradient
  all-reduce over peer memory (reduce-scatter -> update of the owned shard ->
  all-gather of the new parameters; through the NVSwitch with
  multimem.ld_reduce / multimem.st when a multicast alias exists), see
  csrc/kernels/optim.cu:k_ssgd_fused(_nvls);
* with ``overlap=True`` the arena is cut into buckets in reverse registration
  order and every bucket's gather + fused step is launched from the autograd
  hooks of its parameters on a high-priority stream, so the exchange and the
  update of the late layers run under the backward pass of the early ones.
  The reference needed a dedicated NCCL thread and a run-time order
  negotiation for this (/root/reference/srcs/cpp/src/nccl/scheduler.cpp:38-120);
  here the order is static and identical on every rank.

The reference wraps framework optimizers
(/root/reference/srcs/python/kungfu/tensorflow/optimizers/core.py:6-72,
/root/reference/benchmarks/system/benchmark_kungfu_torch.py:76-83 uses SGD;
BERT-style workloads use Adam).
"""
import ctypes as C
import os

import torch

from ..ops import _cuda as K
from ..parallel import comm as _comm
from ..parallel.flat import build_arenas

OVERLAP_CHANNEL = 2


def _record_step(comm, algo, nbytes, stream, ev0=None):
    """Feed the device traffic statistics (parallel/devstats.py): a fused step moves the same bytes
    over NVLink as the all-reduce variant it is built on."""
    st = getattr(comm, 'stats', None)
    if st is not None:
        st.end(ev0, algo, nbytes, stream)


def _begin_step(comm, stream):
    st = getattr(comm, 'stats', None)
    return st.begin(stream) if st is not None else None


class _Bucket(object):
    __slots__ = ('group', 'arena', 'idx', 'lo', 'hi', 'slot', 'pending')

    def __init__(self, group, arena, idx, lo, hi, slot):
        self.group, self.arena, self.idx, self.lo, self.hi, self.slot = group, arena, idx, lo, hi, slot
        self.pending = len(idx)


class FusedOptimizer(torch.optim.Optimizer):
    """Engine shared by FusedSGD / FusedAdamW / FusedAdam (see module docstring)."""
    KIND = 0  # kfb_sgd_hyper.kind
    BUCKET_BYTES = 12 << 20

    def __init__(self, params, defaults, comm=None, blocks=None, threads=None, overlap=False, bucket_bytes=None):
        super().__init__(params, defaults)
        self._comm = comm
        self._arenas = None  # list of (group, [Arena])
        self._steps = 0
        self._step_t = None  # device copy of _steps (Adam's bias corrections read it inside the kernels)
        self._blocks = blocks
        self.local_blocks = int(os.environ.get('KUNGFU_LOCAL_STEP_BLOCKS', '128'))
        self._threads = threads
        self._local = None
        self._stage_buf = None
        self.sumsq_global = None  # ||averaged grad||^2 of the last sync step (device float)
        self.sumsq_local = None  # ||local grad||^2 of the last step (device float)
        self.monitor_norms = False
        self.use_nvls = os.environ.get('KUNGFU_FUSED_NVLS', '1') != '0'
        self.nvls_min_bytes = int(os.environ.get('KUNGFU_FUSED_NVLS_MIN_BYTES', str(256 << 10)))
        self.last_algo = None  # 'nvls' | 'p2p' | 'local': what the last sync step ran
        self.step_algos = []
        self._algos = set()
        # overlapped (bucketed) synchronous step
        self.overlap = bool(overlap)
        self.bucket_bytes = int(bucket_bytes or os.environ.get('KUNGFU_BUCKET_BYTES', self.BUCKET_BYTES))
        self._ov_average = True
        self._ov_sync = False
        self._ov = None
        # CTAs of a bucket kernel that runs UNDER backward. They may sit at the entry barrier for as
        # long as the ranks are skewed, and every resident CTA takes a slot away from the one-wave
        # grids of the model kernels, so hidden buckets get few CTAs (they only have to finish
        # before backward does); the last bucket, whose time is exposed, gets the full set.
        self.hidden_blocks = int(os.environ.get('KUNGFU_OVERLAP_BLOCKS', '8'))
        # how many autograd hooks after a bucket's arrival signal its exchange kernel is launched
        self.defer_hooks = int(os.environ.get('KUNGFU_OVERLAP_DEFER_HOOKS', '6'))

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
                    raise K.KungFuCudaError('%s only handles CUDA parameters' % type(self).__name__)
                out.append((g, build_arenas(self.comm, named)))
            self._arenas = out
            dev = self.comm.device
            self._step_t = torch.full((1, ), self._steps, dtype=torch.int64, device=dev)
            self._size_norm_slots(sum(len(a) for _, a in out))
            if self.overlap:
                self._setup_overlap()
        return self._arenas

    def _size_norm_slots(self, n):
        dev = self.comm.device
        n = max(n, 1)
        self._ss_local_parts = torch.zeros(n, dtype=torch.float32, device=dev)
        self._ss_global_parts = torch.zeros(n, dtype=torch.float32, device=dev)
        self._ss_out = torch.zeros(2, dtype=torch.float32, device=dev)

    def _all_arenas(self):
        return [a for _, ars in self.arenas() for a in ars]

    def rebuild(self, comm=None):
        """Re-create the arenas on a new communicator (after resize_cluster).
        Parameter values, master weights and optimizer slots are carried over;
        the old communicator's buffers are released afterwards. Collective:
        every member of the new cluster calls it (joiners included)."""
        old = self._arenas
        self._teardown_overlap(consolidate=False)  # the caller consolidated before the membership changed
        self._comm = comm
        self._arenas = None
        self._local = None
        self._stage_buf = None
        if hasattr(self, '_kf_stage'):
            self._kf_stage = {}  # SMA staging buffers belong to the old communicator
        # every member (joiners included) allocates its arenas here: symmetric
        # allocation is collective and must happen in the same order everywhere
        news = [a for _, ars in self.arenas() for a in ars]
        if old is not None:
            olds = [a for _, ars in old for a in ars]
            for o, n in zip(olds, news):
                if o.master is not None:
                    n.ensure_master(slots=self._n_slots())
                    n.master.copy_(o.master)
                    n.mom.copy_(o.mom)
                    if o.mom2 is not None:
                        n.mom2.copy_(o.mom2)
        del old
        torch.cuda.synchronize()
        from .. import python as kfpy
        kfpy.run_barrier()
        _comm.release_retired()

    # -- hyper-parameters ---------------------------------------------------------------
    def _n_slots(self):
        return 1

    def _fill_hyper(self, h, g):
        raise NotImplementedError

    def _hyper(self, g, arena=None, lo=0):
        """kfb_sgd_hyper for one launch over arena[lo:...] of param group g."""
        h = K.SgdHyper()
        h.kind = self.KIND
        h.first_step = 1 if self._steps == 0 else 0
        self._fill_hyper(h, g)
        if self.KIND != 0:
            h.step = self._step_t.data_ptr()
            if arena is not None and arena.mom2 is not None:
                h.exp_avg_sq = arena.mom2.data_ptr() + lo * 4
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

    # -- one fused launch over [lo, hi) of an arena ----------------------------------------
    def _launch_range(self, g, a, lo, hi, sync, grad_scale, ss_g, cstruct, blocks, threads, stream,
                      local_blocks=None):
        comm = self.comm
        a.ensure_master(slots=self._n_slots())
        h = self._hyper(g, a, lo)
        count = hi - lo
        esz = a.flat_grads.element_size()
        go, mo = lo * esz, lo * 4
        ssp = ss_g.data_ptr() if ss_g is not None else None
        multi = sync and comm.np > 1 and not comm.cross
        from ..ops import collective as coll
        use_nvls = (multi and a.grads_buf.mc_ptr and a.params_buf.mc_ptr and self.use_nvls
                    and count * esz >= self.nvls_min_bytes
                    and coll.PREFERRED_LARGE in (None, K.ALGO_NVLS))  # the planner may have voted NVLS out
        if (multi and _comm.require_nvls() and count * esz >= self.nvls_min_bytes and self.use_nvls
                and not (a.grads_buf.mc_ptr and a.params_buf.mc_ptr)):
            raise K.KungFuCudaError('KUNGFU_REQUIRE_NVLS=1 but the fused step has no multicast alias to run on')
        if use_nvls:
            # the switch reduces the gradient shard and multicasts the new parameters
            self._algos.add('nvls')
            ev0 = _begin_step(comm, stream)
            K.call('kfb_ssgd_fused_step_nvls', cstruct, a.grads_buf.mc_ptr + go, a.code,
                   a.params_buf.mc_ptr + go, a.code, a.master.data_ptr() + mo, a.mom.data_ptr() + mo, count,
                   float(grad_scale), C.byref(h), ssp, 1 if coll.SLOW_VARIANT == K.ALGO_NVLS else min(blocks, 32),
                   threads, K.stream_ptr(stream))
            _record_step(comm, K.ALGO_NVLS, count * esz, stream, ev0)
            return
        if multi:
            self._algos.add('p2p')
            ga, _k1 = a.grads_buf.ptr_array(go)
            pa, _k2 = a.params_buf.ptr_array(go)
            cs = cstruct
            step_blocks = 1 if coll.SLOW_VARIANT == K.ALGO_TWOSHOT else blocks
        else:
            self._algos.add('local')
            me = max(comm.rank, 0)
            ga, _k1 = K.ptr_array([a.grads_buf.ptrs[me] + go])
            pa, _k2 = K.ptr_array([a.params_buf.ptrs[me] + go])
            cs = self._local_struct()
            # a purely local update is HBM-bound, not NVLink-bound: it needs every SM it can
            # get (the comm default of 32 CTAs reaches ~1.7 TB/s; 128 is the flag-slot limit)
            step_blocks = local_blocks or self.local_blocks
        if comm.emulated and (lo != 0 or hi != a.count):
            raise K.KungFuCudaError('bucketed steps are not available under single-GPU rank emulation')
        ev0 = _begin_step(comm, stream) if multi else None
        K.call('kfb_ssgd_fused_step', cs, ga, a.code, pa, a.code, a.master.data_ptr() + mo,
               a.mom.data_ptr() + mo, count, float(grad_scale), C.byref(h), ssp, step_blocks, threads,
               K.stream_ptr(stream))
        if multi:
            _record_step(comm, K.ALGO_TWOSHOT, count * esz, stream, ev0)

    # -- monolithic step ------------------------------------------------------------------
    def _run(self, sync, grad_scale):
        comm = self.comm
        blocks = self._blocks or comm.default_blocks
        threads = self._threads or comm.default_threads
        k = 0
        for g, arenas in self.arenas():
            for a in arenas:
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
                self._launch_range(g, a, 0, a.count, sync, step_scale, ss_g, comm.cstruct, blocks, threads, None)
                k += 1
        if self.monitor_norms:
            self._sum_norm_slots(k)
        self._after_step()

    def _sum_norm_slots(self, k):
        # fixed slots summed in index order by one thread block: reproducible
        K.call('kfb_sum_slots', self._ss_local_parts.data_ptr(), k, self._ss_out.data_ptr(), K.stream_ptr())
        K.call('kfb_sum_slots', self._ss_global_parts.data_ptr(), k, self._ss_out.data_ptr() + 4, K.stream_ptr())
        self.sumsq_local = self._ss_out[0]
        self.sumsq_global = self._ss_out[1]

    def _after_step(self):
        # the most capable path any launch of this step took
        for name in ('nvls', 'p2p', 'local'):
            if name in self._algos:
                self.last_algo = name
                break
        self.step_algos = sorted(self._algos)
        self._algos = set()
        self._steps += 1
        self._step_t.add_(1)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if self._ov is not None and self._ov['armed']:
            if self._ov_sync:
                raise RuntimeError('the overlapped step was armed for synchronous steps; use sync_step() '
                                   '(SynchronousSGDOptimizer does) or enable_overlap(sync=False)')
            # hooks are installed: buckets were (or are now) launched without any exchange
            self._finish_overlap()
        else:
            self._run(sync=False, grad_scale=1.0)
        return loss

    @torch.no_grad()
    def sync_step(self, average=True):
        """All-reduce gradients over peer memory and apply the update, fused."""
        self.arenas()
        if self._ov is not None and self._ov['armed']:
            if not self._ov_sync or bool(average) != bool(self._ov_average):
                raise RuntimeError('the overlapped step was armed with sync=%s, average=%s: call '
                                   'enable_overlap(sync=True, average=%s) first' %
                                   (self._ov_sync, self._ov_average, bool(average)))
            self._finish_overlap()
            return
        n = self.comm.global_np if self.comm.cross else self.comm.np
        scale = 1.0 / n if average else 1.0
        self._run(sync=True, grad_scale=scale)

    # -- overlapped, bucketed step ----------------------------------------------------------
    def enable_overlap(self, sync=True, average=True, bucket_bytes=None):
        """Launch every bucket's gather + fused step from autograd hooks while
        backward is still running. ``sync``: the step is a synchronous S-SGD
        step (gradients all-reduced) rather than a purely local update.
        One backward pass per step() is assumed (no gradient accumulation);
        call disable_overlap() around anything else."""
        if bucket_bytes:
            self.bucket_bytes = int(bucket_bytes)
        self.overlap = True
        self._ov_sync = bool(sync)
        self._ov_average = bool(average)
        if self._arenas is not None:
            self._setup_overlap()

    def rebind_overlap(self):
        """Re-register the autograd hooks on this object (an optimizer wrapper
        re-creates the object around the same state)."""
        if self._ov is not None:
            self._teardown_overlap()
            self._setup_overlap()

    def disable_overlap(self):
        self.overlap = False
        self._teardown_overlap()

    def _teardown_overlap(self, consolidate=True):
        if self._ov is None:
            return
        if consolidate and self._arenas is not None and self.comm.np > 1:
            self.consolidate()  # ownership of the optimizer state is about to change shape
        for h in self._ov['hooks']:
            h.remove()
        for a in self._all_arenas() if self._arenas is not None else []:
            a.segments = [(0, a.count)]
        if self._arenas is not None:
            self._size_norm_slots(sum(len(a) for _, a in self._arenas))
        self._ov = None

    def _setup_overlap(self):
        comm = self.comm
        if self._ov is not None:
            self._teardown_overlap()
        if comm.cross or comm.emulated:
            # several hosts: the cross-host leg runs on the host; rank emulation: one cooperative
            # launch per collective. Both keep the single post-backward step.
            return
        if any(a.master is not None for a in self._all_arenas()) and comm.np > 1:
            self.consolidate()
        buckets, where = [], {}
        for g, arenas in self._arenas:
            for a in arenas:
                esize = a.flat_grads.element_size()
                cur, nbytes, segs = [], 0, []
                # backward produces gradients roughly in reverse registration order
                for i in reversed(range(len(a.params))):
                    cur.append(i)
                    nbytes += a.params[i].numel() * esize
                    if nbytes >= self.bucket_bytes or i == 0:
                        lo = a.offsets[min(cur)]
                        last = max(cur)
                        hi = a.offsets[last + 1] if last + 1 < len(a.params) else a.count
                        b = _Bucket(g, a, list(cur), lo, hi, len(buckets))
                        for j in cur:
                            where[id(a.params[j])] = b
                        buckets.append(b)
                        segs.append((lo, hi))
                        cur, nbytes = [], 0
                a.segments = sorted(segs)
        hooks = []
        for b in buckets:
            for j in b.idx:
                hooks.append(b.arena.params[j].register_post_accumulate_grad_hook(self._on_grad))
        stream = torch.cuda.Stream(device=comm.device, priority=-1)
        self._ov = {'buckets': buckets, 'where': where, 'hooks': hooks, 'stream': stream, 'fused': 0, 'stepped': 0,
                    'since': 0, 'armed': True,
                    'fuse_scratch': torch.zeros(K.FUSE_SCRATCH_FLOATS, dtype=torch.float32, device=comm.device)}
        self._size_norm_slots(len(buckets))
        if comm.np > 1:
            # per-bucket norm partials of this rank's shards, summed over the ranks once per step
            nb4 = (len(buckets) + 3) // 4 * 4
            self._ov['norm_in'] = comm.alloc(nb4 * 4)
            self._ov['norm_out'] = comm.alloc(nb4 * 4)
            self._ss_partial = self._ov['norm_in'].view(torch.float32)[:nb4]
            self._ss_global_parts = self._ov['norm_out'].view(torch.float32)[:nb4]
        else:
            self._ss_partial = self._ss_global_parts

    def _on_grad(self, p):
        ov = self._ov
        if ov is None or not ov['armed']:
            return
        ov['where'][id(p)].pending -= 1
        ov['since'] += 1
        self._progress()

    def _progress(self, force=False):
        """Buckets go through two launches, in bucket order, on the side stream:
        (1) gather + arrival signal as soon as all gradients of the bucket exist;
        (2) the fused exchange + update a few autograd hooks later (or when the
        next bucket is complete), ordered behind the main stream at that point.
        By then every peer's arrival is normally in, so the kernel does not sit
        on SMs waiting for a slower rank while backward needs them."""
        ov = self._ov
        bks = ov['buckets']
        nb = len(bks)
        while True:
            if ov['stepped'] < ov['fused']:
                nxt_ready = ov['fused'] < nb and bks[ov['fused']].pending <= 0
                if force or nxt_ready or ov['since'] >= self.defer_hooks:
                    self._step_bucket(bks[ov['stepped']])
                    ov['stepped'] += 1
                    continue
                return
            if ov['fused'] < nb and (bks[ov['fused']].pending <= 0 or force):
                self._fuse_bucket(bks[ov['fused']])
                ov['fused'] += 1
                ov['since'] = 0
                continue
            return

    def _bucket_blocks(self, b):
        comm, ov = self.comm, self._ov
        blocks = min(comm.channel_blocks(OVERLAP_CHANNEL, self._blocks), 32)
        if b.slot != len(ov['buckets']) - 1:
            blocks = min(blocks, self.hidden_blocks)
        return blocks

    def _multi(self):
        comm = self.comm
        return self._ov_sync and comm.np > 1 and not comm.cross

    def _fuse_bucket(self, b):
        comm, ov = self.comm, self._ov
        side = ov['stream']
        # everything this bucket reads was produced on the stream that is current now
        side.wait_stream(torch.cuda.current_stream())
        ss_l = self._ss_local_parts[b.slot:b.slot + 1] if self.monitor_norms else None
        b.arena.fuse_grads(sumsq_out=ss_l, stream=side, idx=b.idx, scratch=ov['fuse_scratch'])
        if self._multi():
            K.call('kfb_peer_signal', comm.channel(OVERLAP_CHANNEL), self._bucket_blocks(b), K.stream_ptr(side))

    def _step_bucket(self, b):
        comm, ov = self.comm, self._ov
        side = ov['stream']
        # ordered behind the main stream as of NOW: this is what keeps the kernel from starting
        # (and spinning) right behind its own signal
        side.wait_stream(torch.cuda.current_stream())
        a = b.arena
        ss_g = self._ss_partial[b.slot:b.slot + 1] if self.monitor_norms else None
        sync = self._ov_sync
        scale = (1.0 / comm.np) if (sync and self._ov_average) else 1.0
        last = b.slot == len(ov['buckets']) - 1
        blocks = self._bucket_blocks(b)
        threads = self._threads or comm.default_threads
        if self._multi():
            flags = K.STEP_PRESIGNALED | K.STEP_PARTIAL_NORM
            if not last:
                flags |= K.STEP_NO_EXIT_BARRIER
            K.lib().kfb_set_step_flags(flags)
        self._launch_range(b.group, a, b.lo, b.hi, sync, scale, ss_g, comm.channel(OVERLAP_CHANNEL), blocks,
                           threads, side, local_blocks=None if last else self.hidden_blocks * 4)

    def _finish_overlap(self):
        ov = self._ov
        self._progress(force=True)
        torch.cuda.current_stream().wait_stream(ov['stream'])
        for b in ov['buckets']:
            b.pending = len(b.idx)
        ov['fused'] = ov['stepped'] = ov['since'] = 0
        if self.monitor_norms:
            n = len(ov['buckets'])
            if self._multi():
                # every slot holds this rank's shard partial: add the ranks up (rank order: identical everywhere)
                from ..ops import collective as coll
                coll.all_reduce(self.comm, ov['norm_in'], (n + 3) // 4 * 4, K.F32, 'sum', algo=K.ALGO_ONESHOT,
                                out=ov['norm_out'])
            elif self._ss_partial is not self._ss_global_parts:
                self._ss_global_parts.copy_(self._ss_partial)
            self._sum_norm_slots(n)
        self._after_step()

    # -- state ------------------------------------------------------------------------------
    def _stage(self, nbytes):
        """One symmetric staging buffer per optimizer, grown on demand (collective)."""
        if self._stage_buf is None or self._stage_buf.nbytes < nbytes:
            if self._stage_buf is not None:
                self.comm.free(self._stage_buf)
            self._stage_buf = self.comm.alloc(nbytes, multicast=True)
        return self._stage_buf

    def consolidate(self):
        """Make master weights and optimizer slots complete on every rank (during
        synchronous fused steps a rank only maintains the shards it owns). One
        all-reduce per tensor over a reused staging buffer: every element is
        live on exactly one rank, the others contribute zeros."""
        comm = self.comm
        if comm.np <= 1 or self._arenas is None or comm.cross:
            return
        from ..ops import collective as coll
        need = max((a.count * 4 for a in self._all_arenas() if a.master is not None), default=0)
        if need == 0:
            return
        stage = self._stage(need)
        sv = stage.view(torch.float32)
        R = comm.n_local_ranks
        for a in self._all_arenas():
            if a.master is None:
                continue
            for t in [a.master, a.mom] + ([a.mom2] if a.mom2 is not None else []):
                for r in range(R):
                    rank = r if comm.emulated else comm.rank
                    src = t[r * a.count:(r + 1) * a.count]
                    dst = stage.view(torch.float32, r if comm.emulated else None)[:a.count]
                    dst.zero_()
                    for lo, hi in a.segments:
                        npk = (hi - lo) // 8
                        per = (npk + comm.np - 1) // comm.np * 8
                        s0 = min(lo + rank * per, hi)
                        s1 = min(s0 + per, hi)
                        dst[s0:s1].copy_(src[s0:s1])
                coll.all_reduce(comm, stage, a.count, K.F32, 'sum', algo=K.ALGO_TWOSHOT)
                for r in range(R):
                    t[r * a.count:(r + 1) * a.count].copy_(
                        stage.view(torch.float32, r if comm.emulated else None)[:a.count])
        del sv
        torch.cuda.synchronize()

    def broadcast_state(self):
        """Master weights, optimizer slots and the step counter := rank 0's
        (after a resize: rank 0 consolidated before the membership changed,
        joiners start from its state)."""
        from ..torch.ops import collective as tops
        steps = torch.tensor([self._steps], dtype=torch.int64)
        self._steps = int(tops.all_reduce_fn(steps, 'max', name='kungfu::fused-steps')[0])
        self.arenas()
        self._step_t.fill_(self._steps)
        if self.comm.np <= 1 and not self.comm.cross:
            return
        for gi, (_, arenas) in enumerate(self.arenas()):
            for ai, a in enumerate(arenas):
                a.ensure_master(slots=self._n_slots())
                tops.broadcast(a.master, name='kungfu::master.%d.%d' % (gi, ai))
                tops.broadcast(a.mom, name='kungfu::mom.%d.%d' % (gi, ai))
                if a.mom2 is not None:
                    tops.broadcast(a.mom2, name='kungfu::mom2.%d.%d' % (gi, ai))
        torch.cuda.synchronize()

    SLOT_NAMES = ('momentum_buffer', )

    def state_dict(self):
        """Checkpoint with the fp32 master weights and the optimizer slots (the
        reference checkpoints optimizer slots with the model,
        /root/reference/srcs/python/kungfu/tensorflow/hooks/elastic.py:80-87).
        Collective when np > 1 (the sharded state is consolidated first).
        Layout: torch's {'state': {index: {...}}, 'param_groups': [...]} with
        'master' (fp32 weights) next to the usual slot names, all in the
        parameter's logical shape."""
        self.arenas()
        self.consolidate()
        state, groups, base = {}, [], 0
        for g, arenas in self._arenas:
            idx_of = {id(p): base + i for i, p in enumerate(g['params'])}
            for a in arenas:
                if a.master is None:
                    continue
                slots = [a.mom] + ([a.mom2] if a.mom2 is not None else [])
                for i, p in enumerate(a.params):
                    e = {'master': a.state_view(a.master[:a.count], i).clone().contiguous()}
                    for name, t in zip(self.SLOT_NAMES, slots):
                        e[name] = a.state_view(t[:a.count], i).clone().contiguous()
                    state[idx_of[id(p)]] = e
            pg = {k: v for k, v in g.items() if k != 'params'}
            pg['params'] = list(range(base, base + len(g['params'])))
            groups.append(pg)
            base += len(g['params'])
        return {'state': state, 'param_groups': groups, 'kungfu': {'steps': self._steps, 'kind': self.KIND}}

    def load_state_dict(self, sd):
        """Inverse of state_dict(): hyper-parameters, step counter, fp32 masters
        and slots; the low-precision parameters are re-derived from the masters."""
        if len(sd['param_groups']) != len(self.param_groups):
            raise ValueError('loaded state dict has a different number of parameter groups')
        for g, sg in zip(self.param_groups, sd['param_groups']):
            if len(g['params']) != len(sg['params']):
                raise ValueError("loaded state dict contains a parameter group that doesn't match the size of "
                                 "optimizer's group")
            g.update({k: v for k, v in sg.items() if k != 'params'})
        self._steps = int(sd.get('kungfu', {}).get('steps', 0))
        self.arenas()
        self._step_t.fill_(self._steps)
        base = 0
        st = sd['state']
        for g, arenas in self._arenas:
            idx_of = {id(p): base + i for i, p in enumerate(g['params'])}
            for a in arenas:
                slots = None
                for i, p in enumerate(a.params):
                    e = st.get(idx_of[id(p)], st.get(str(idx_of[id(p)])))
                    if e is None:
                        continue
                    a.ensure_master(slots=self._n_slots())
                    if slots is None:
                        slots = [a.mom] + ([a.mom2] if a.mom2 is not None else [])
                    R = a.comm.n_local_ranks
                    for r in range(R):
                        a.state_view(a.master[r * a.count:(r + 1) * a.count], i).copy_(e['master'])
                        for name, t in zip(self.SLOT_NAMES, slots):
                            if name in e:
                                a.state_view(t[r * a.count:(r + 1) * a.count], i).copy_(e[name])
                    p.data.copy_(e['master'])
            base += len(g['params'])
        torch.cuda.synchronize()


class FusedSGD(FusedOptimizer):
    """torch.optim.SGD semantics (weight decay, momentum, dampening, nesterov)."""
    KIND = 0

    def __init__(self, params, lr=1e-3, momentum=0.0, dampening=0.0, weight_decay=0.0,
                 nesterov=False, comm=None, blocks=None, threads=None, overlap=False, bucket_bytes=None):
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError('Nesterov momentum requires a momentum and zero dampening')
        defaults = dict(lr=lr, momentum=momentum, dampening=dampening,
                        weight_decay=weight_decay, nesterov=nesterov)
        super().__init__(params, defaults, comm, blocks, threads, overlap, bucket_bytes)

    def _fill_hyper(self, h, g):
        h.lr = float(g['lr'])
        h.momentum = float(g['momentum'])
        h.dampening = float(g['dampening'])
        h.weight_decay = float(g['weight_decay'])
        h.nesterov = 1 if g['nesterov'] else 0


class FusedAdamW(FusedOptimizer):
    """torch.optim.AdamW semantics (decoupled weight decay; no amsgrad) with
    fp32 master weights and moments. The bias corrections are computed inside
    the kernels from a device-side step counter, so a captured CUDA graph of
    the training step stays valid from one replay to the next."""
    KIND = 1
    SLOT_NAMES = ('exp_avg', 'exp_avg_sq')

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, comm=None,
                 blocks=None, threads=None, overlap=False, bucket_bytes=None):
        if not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0:
            raise ValueError('invalid betas %s' % (betas, ))
        defaults = dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults, comm, blocks, threads, overlap, bucket_bytes)

    def _n_slots(self):
        return 2

    def _fill_hyper(self, h, g):
        h.lr = float(g['lr'])
        h.beta1, h.beta2 = float(g['betas'][0]), float(g['betas'][1])
        h.eps = float(g['eps'])
        h.weight_decay = float(g['weight_decay'])
        t = self._steps + 1
        h.bias_corr1 = 1.0 - h.beta1 ** t
        h.bias_corr2 = 1.0 - h.beta2 ** t


class FusedAdam(FusedAdamW):
    """torch.optim.Adam semantics (weight decay as an L2 term of the gradient)."""
    KIND = 2

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, **kw):
        super().__init__(params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, **kw)
