"""Strategy planner on the device: graph executor + online variant chooser.

The reference walks a (reduce graph, broadcast graph) pair on the host for
every 1 MiB chunk (/root/reference/srcs/go/kungfu/session/session.go:218-313)
and switches strategy when measured throughput drops
(/root/reference/srcs/go/kungfu/session/adaptiveStrategies.go:61-121,
/root/reference/srcs/go/kungfu/session/adaptation.go:8-28).

On one NVSwitch domain topology shaping buys no bandwidth, so the planner has
two jobs here:

* ``GraphExecutor``: API parity for -strategy / set_tree / monitored
  all-reduce on CUDA tensors. A strategy's graphs are compiled into rounds
  (nodes of equal depth); a round is a device barrier plus single-edge
  pull kernels (csrc/kernels/collectives.cu:k_edge_reduce / k_edge_copy) on the
  caller's stream. The buffer is cut into one chunk per active graph pair
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


def _depths(n, edges):
    """depth[v] = longest path from any source to v along `edges`."""
    prevs = {v: [] for v in range(n)}
    for a, b in edges:
        prevs[b].append(a)
    depth = {}

    def d(v, seen=()):
        if v in depth:
            return depth[v]
        if v in seen:
            raise ValueError('cycle in strategy graph')
        depth[v] = 0 if not prevs[v] else 1 + max(d(u, seen + (v, )) for u in prevs[v])
        return depth[v]

    for v in range(n):
        d(v)
    return depth, prevs


class GraphExecutor(object):
    def __init__(self, comm, strategies=None):
        self.comm = comm
        self.strategies = strategies if strategies is not None else kfpy.peer().strategies()
        self.stats = []  # (seconds, bytes) of monitored runs
        self._plans = [self._compile(s) for s in self.strategies]

    def _compile(self, s):
        n = self.comm.np
        rank = self.comm.rank
        rdepth, rprev = _depths(n, s['reduce'])
        bdepth, bprev = _depths(n, s['bcast'])
        rounds_r = max(rdepth.values()) if rdepth else 0
        rounds_b = max(bdepth.values()) if bdepth else 0
        return {
            'reduce_rounds': rounds_r, 'bcast_rounds': rounds_b,
            'my_rdepth': rdepth[rank], 'my_rprev': rprev[rank] if rank in s['loops'] else [],
            'my_bdepth': bdepth[rank], 'my_bprev': bprev[rank],
        }

    def all_reduce(self, buf, count, dtype, op='sum', byte_offset=0, monitored=False, stream=None):
        """In-place all-reduce of buf[byte_offset : +count] following the active graphs."""
        comm = self.comm
        if comm.np == 1:
            return
        code = dtype if isinstance(dtype, int) else K.dtype_code(dtype)
        es = coll._ESIZE[code]
        k = len(self._plans)
        # chunk i -> strategy i (aligned to 16 bytes)
        per = ((count + k - 1) // k * es + 15) // 16 * 16 // es
        ev0 = ev1 = None
        if monitored:
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record(stream)
        max_r = max(p['reduce_rounds'] for p in self._plans)
        max_b = max(p['bcast_rounds'] for p in self._plans)
        me = comm.rank
        for rnd in range(1, max_r + 1):
            comm.barrier(stream)  # inputs of this round are final on every rank
            for i, p in enumerate(self._plans):
                lo = i * per
                n = min(per, count - lo)
                if n <= 0 or p['my_rdepth'] != rnd:
                    continue
                off = byte_offset + lo * es
                for src in p['my_rprev']:
                    coll.edge_reduce(buf.ptrs[me] + off, buf.ptrs[src] + off, n, code, op, stream=stream)
        for rnd in range(1, max_b + 1):
            comm.barrier(stream)
            for i, p in enumerate(self._plans):
                lo = i * per
                n = min(per, count - lo)
                if n <= 0 or p['my_bdepth'] != rnd:
                    continue
                off = byte_offset + lo * es
                for src in p['my_bprev']:
                    coll.edge_copy(buf.ptrs[me] + off, buf.ptrs[src] + off, n * es, stream=stream)
        comm.barrier(stream)  # nobody overwrites its buffer while a child still pulls
        if monitored:
            ev1.record(stream)
            ev1.synchronize()
            self.stats.append((ev0.elapsed_time(ev1) / 1e3, count * es))

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
        cap = max(sizes)
        buf = comm.alloc(cap, multicast=True)
        out = comm.alloc(cap, multicast=True)
        code = K.dtype_code(self.dtype)
        algos = [K.ALGO_ONESHOT, K.ALGO_TWOSHOT] + ([K.ALGO_NVLS] if buf.mc_ptr else [])
        for nbytes in sizes:
            self.table[nbytes] = {}
            for algo in algos:
                if algo in self.disabled or (algo == K.ALGO_ONESHOT and nbytes > (8 << 20)):
                    continue
                self.table[nbytes][algo] = self._time(
                    lambda: coll.all_reduce(comm, buf, nbytes // es, code, 'sum', algo=algo, out=out))
            self.best[nbytes] = min(self.table[nbytes], key=self.table[nbytes].get)
        big = max(sizes)
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
