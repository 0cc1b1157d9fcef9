"""Strategy planner on the device: graph executor + online variant chooser.

The reference walks a (reduce graph, broadcast graph) pair on the host for
every 1 MiB chunk (/root/reference/srcs/go/kungfu/session/session.go:218-313)
and switches strategy when measured throughput drops
(/root/reference/srcs/go/kungfu/session/adaptiveStrategies.go:61-121,
/root/reference/srcs/go/kungfu/session/adaptation.go:8-28).

On one NVSwitch domain topology shaping buys no bandwidth, so the planner has
two jobs here:

* ``GraphExecutor``: API parity for -strategy / set_tree / monitored
  all-reduce on CUDA tensors. The runtime compiles a strategy's graphs into
  rounds (nodes of equal depth) of pull edges for this rank; one kernel walks
  all rounds with peer barriers in between (csrc/kernels/collectives.cu:
  k_graph_allreduce). The buffer is cut into one chunk per active graph pair
  (RING/CLIQUE/MULTI_* have several), like the reference's chunk hashing.
* ``VariantChooser``: the B200 meaning of "adaptive strategy selection" -- it
  times the fused variants {one-shot, two-shot, NVLS} with CUDA events at a few
  sizes, agrees on the result across ranks (max over ranks, control plane) and
  installs the size thresholds the collectives use. ``check_interference``
  re-measures the active variant against its reference window (< 80 % => vote,
  majority => switch to the next-best variant).
"""
import torch

from .. import python as kfpy
from ..ops import _cuda as K
from ..ops import collective as coll


class GraphExecutor(object):
    """Runs the active (reduce, broadcast) graph pairs on the device.

    The graphs are compiled by the runtime (csrc/core/graph.cpp:compile_device_plan, `kf_device_plan`) into
    this rank's edge table: for every round, which chunk to pull from which rank (see kfb_graph_allreduce
    in include/kungfu_b200.h). On a real communicator the whole walk is ONE kernel launch; the same table
    can be walked from Python with one kernel per edge (`launcher='python'`), which is what the host
    simulation in tests/test_planner_cpu.py drives with threads and numpy."""

    def __init__(self, comm, table=None, launcher=None):
        self.comm = comm
        self.table = table if table is not None else kfpy.peer().device_plan()
        self.stats = []  # (seconds, bytes) of monitored runs
        self.launcher = launcher or ('device' if hasattr(comm, 'cstruct') and not getattr(comm, 'emulated', False)
                                     else 'python')

    def all_reduce(self, buf, count, dtype, op='sum', byte_offset=0, monitored=False, stream=None):
        """In-place all-reduce of buf[byte_offset : +count] following the active graphs."""
        comm = self.comm
        if comm.np == 1:
            return
        code = dtype if isinstance(dtype, int) else K.dtype_code(dtype)
        es = coll._ESIZE[code]
        k = self.table['n_plans']
        # chunk i -> strategy i (aligned to 16 bytes)
        per = ((count + k - 1) // k * es + 15) // 16 * 16 // es
        chunks = []
        for i in range(k):
            lo = i * per
            chunks.append((byte_offset + lo * es, max(0, min(per, count - lo))))
        ev0 = ev1 = None
        if monitored:
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record(stream)
        if self.launcher == 'device':
            self._walk_device(buf, chunks, code, es, op, stream)
        else:
            self._walk_python(buf, chunks, code, es, op, stream)
        if monitored:
            ev1.record(stream)
        st = getattr(comm, 'stats', None)
        if st is not None:
            from .devstats import ALGO_GRAPH
            st.record(ALGO_GRAPH, count * es, stream, events=(ev0, ev1) if monitored else None)
        if monitored:
            ev1.synchronize()
            self.stats.append((ev0.elapsed_time(ev1) / 1e3, count * es))

    def _walk_device(self, buf, chunks, code, es, op, stream):
        comm, t = self.comm, self.table
        bufs, _k0 = buf.ptr_array()
        lo, _k1 = K.i64_array([c[0] for c in chunks])
        nb, _k2 = K.i64_array([c[1] * es for c in chunks])
        begin, _k3 = K.i32_array(t['begin'])
        plan, _k4 = K.i32_array(t['plan'])
        src, _k5 = K.i32_array(t['src'])
        K.call('kfb_graph_allreduce', comm.cstruct, bufs, code, op if isinstance(op, int) else K.OPS[op],
               len(chunks), lo, nb, len(t['rounds']), t['n_reduce'], begin, plan, src,
               comm.default_blocks, comm.default_threads, K.stream_ptr(stream))

    def _walk_python(self, buf, chunks, code, es, op, stream):
        comm, t = self.comm, self.table
        me = comm.rank
        for r, edges in enumerate(t['rounds']):
            comm.barrier(stream)  # inputs of this round are final on every rank
            for i, src in edges:
                off, n = chunks[i]
                if n <= 0:
                    continue
                if r < t['n_reduce']:
                    coll.edge_reduce(buf.ptrs[me] + off, buf.ptrs[src] + off, n, code, op, stream=stream)
                else:
                    coll.edge_copy(buf.ptrs[me] + off, buf.ptrs[src] + off, n * es, stream=stream)
        comm.barrier(stream)  # nobody overwrites its buffer while a child still pulls

    def throughput(self):
        """bytes/s over the monitored runs since the last call (the reference's StrategyStat window)."""
        if not self.stats:
            return 0.0
        t = sum(s for s, _ in self.stats)
        b = sum(n for _, n in self.stats)
        self.stats = []
        return b / t if t > 0 else 0.0


class VariantChooser(object):
    """Pick {one-shot, two-shot, NVLS} per message size from measured bus bandwidth."""
    PROBE_SIZES = (16 << 10, 256 << 10, 2 << 20, 32 << 20)
    THRESHOLD = 0.8

    def __init__(self, comm, dtype=torch.bfloat16):
        self.comm = comm
        self.dtype = dtype
        self.table = {}      # size -> {algo: ms}
        self.best = {}       # size -> algo
        self.reference = {}  # algo -> busbw at the largest probe size
        self.disabled = set()
        self._bufs = None    # probe buffers (symmetric allocations are collective: made once)
        self.switches = 0

    def _time(self, fn, iters=8):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        e1.synchronize()
        ms = e0.elapsed_time(e1) / iters
        from ..torch.ops import collective as tops
        return float(tops.all_reduce_fn(torch.tensor([ms], dtype=torch.float64), 'max')[0])

    def probe(self, sizes=None):
        comm = self.comm
        if comm.np == 1:
            return
        sizes = sizes or self.PROBE_SIZES
        es = torch.empty(0, dtype=self.dtype).element_size()
        cap = max(max(sizes), max(self.PROBE_SIZES))
        if self._bufs is None or self._bufs[0].nbytes < cap:
            self._bufs = (comm.alloc(cap, multicast=True), comm.alloc(cap, multicast=True))
        buf, out = self._bufs
        code = K.dtype_code(self.dtype)
        algos = [K.ALGO_ONESHOT, K.ALGO_TWOSHOT] + ([K.ALGO_NVLS] if buf.mc_ptr else [])
        for nbytes in sizes:
            self.table[nbytes] = {}
            for algo in algos:
                if algo in self.disabled or (algo == K.ALGO_ONESHOT and nbytes > (8 << 20)):
                    continue
                self.table[nbytes][algo] = self._time(
                    lambda: coll.all_reduce(comm, buf, nbytes // es, code, 'sum', algo=algo, out=out))
            if not self.table[nbytes]:
                del self.table[nbytes]  # nothing left that can serve this size
                continue
            self.best[nbytes] = min(self.table[nbytes], key=self.table[nbytes].get)
        if not self.table:
            return
        big = max(self.table)
        for algo, ms in self.table[big].items():
            self.reference[algo] = big / ms / 1e6 * 2 * (comm.np - 1) / comm.np
        self._install()

    def _install(self):
        """Turn the probe table into the threshold used by ops.collective.choose_algo."""
        sizes = sorted(self.best)
        one_max = 0
        for s in sizes:
            if self.best[s] == K.ALGO_ONESHOT:
                one_max = s
            else:
                break
        coll.ONESHOT_MAX_BYTES = max(one_max, 4 << 10)
        bigs = [self.best[s] for s in sizes if s > coll.ONESHOT_MAX_BYTES]
        coll.PREFERRED_LARGE = max(set(bigs), key=bigs.count) if bigs else None

    def choose(self, nbytes):
        if not self.best:
            return None
        s = min(self.best, key=lambda k: abs(k - nbytes))
        return self.best[s]

    def busbw(self, algo, nbytes=None):
        nbytes = nbytes or max(self.table)
        ms = self.table[nbytes][algo]
        return nbytes / ms / 1e6 * 2 * (self.comm.np - 1) / self.comm.np

    def check_interference(self):
        """Re-measure the active large-message variant; if it fell below 80 %
        of its reference on a majority of ranks, switch to the next best.
        Returns True when a switch happened (same contract as the reference's
        check_interference + set_tree pair)."""
        comm = self.comm
        if comm.np == 1 or not self.reference:
            return False
        big = max(self.table)
        active = self.best[big]
        old_ref = self.reference[active]
        self.probe(sizes=(big, ))
        now = self.busbw(active, big)
        from ..torch.ops import collective as tops
        vote = 1 if now < self.THRESHOLD * old_ref else 0
        total = int(tops.all_reduce_fn(torch.tensor([vote], dtype=torch.int32), 'sum')[0])
        self.reference[active] = old_ref
        if total > comm.np // 2:
            self.disabled.add(active)
            self.probe(sizes=(big, ))
            return True
        return False


class DevicePlanner(object):
    """What `-strategy`, set_tree, monitored_all_reduce, calc_stats and
    check_interference mean for CUDA tensors. One per communicator
    (``planner_for(comm)``); created together with the communicator so that
    the variant probe, which is collective, runs at a point every rank reaches.

    Two adaptive mechanisms, both with the reference's rule (a window under 80 %
    of the best window seen is a vote; a majority of votes switches,
    /root/reference/srcs/go/kungfu/session/adaptiveStrategies.go:61-121):

    * kernel variants {one-shot, two-shot, NVLS}: chosen per size from the probe,
      re-voted from the device-timed windows of the TRAINING collectives
      (DeviceStats), switched by disabling the slow variant and re-probing;
    * strategy graphs: the executor follows whatever set_tree / set_strategy
      installed in the runtime (digest-checked there), so the reference's
      "switch to a star rooted at rank 1" demo works unchanged on GPU tensors.
    """

    def __init__(self, comm):
        self.comm = comm
        self.chooser = VariantChooser(comm)
        self._executor = None
        self._digest = None

    def probe(self):
        import os
        if self.comm.np > 1 and not self.comm.emulated and os.environ.get('KUNGFU_PROBE_VARIANTS', '1') != '0':
            self.chooser.probe(sizes=(64 << 10, 1 << 20, 16 << 20))

    def executor(self):
        """Graph executor for the strategies currently active in the runtime."""
        table = kfpy.peer().device_plan()
        digest = repr(table)
        if self._executor is None or digest != self._digest:
            self._executor = GraphExecutor(self.comm, table)
            self._digest = digest
        return self._executor

    def monitored_all_reduce(self, buf, count, dtype, op='sum', stream=None):
        self.comm.stats.timing = True
        self.executor().all_reduce(buf, count, dtype, op, monitored=True, stream=stream)

    def calc_stats(self):
        self.comm.stats.calc()

    def check_interference(self):
        """Vote on the active kernel variant from the windows calc_stats() closed;
        a majority switches to the next-best variant. Returns True on a switch."""
        comm = self.comm
        if comm.np <= 1:
            return False
        from ..torch.ops import collective as tops
        vote = 1 if comm.stats.below_reference(VariantChooser.THRESHOLD) else 0
        total = int(tops.all_reduce_fn(torch.tensor([vote], dtype=torch.int32), 'sum')[0])
        if total * 2 <= comm.np:
            return False
        active = comm.stats.active
        large = [K.ALGO_TWOSHOT] + ([K.ALGO_NVLS] if getattr(comm, 'multicast_ok', False) else [])
        alternatives = [a for a in large if a != active and a not in self.chooser.disabled]
        if active in large and not alternatives:
            # the vote passed but there is nothing to switch to (no multicast: two-shot is the only
            # large-message variant); the caller still learns that the workers see interference
            return True
        if active in (K.ALGO_ONESHOT, K.ALGO_TWOSHOT, K.ALGO_NVLS):
            self.chooser.disabled.add(active)
            self.chooser.probe(sizes=(1 << 20, 16 << 20))
            self.chooser.switches += 1
            # the new variant starts a fresh reference window
            for st in comm.stats.by_variant.values():
                st.reference = 0.0
            return True
        return False


def planner_for(comm):
    if comm.planner is None:
        comm.planner = DevicePlanner(comm)
    return comm.planner
