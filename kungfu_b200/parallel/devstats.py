"""Traffic statistics of the device data plane.

In the reference the monitored all-reduce is the training all-reduce: the
session accumulates bytes and wall time per strategy
(/root/reference/srcs/go/kungfu/session/monitoring.go:15-31,
/root/reference/srcs/go/kungfu/session/strategy.go:15-56), `calc_stats` folds
them into a throughput window, `check_interference` votes on it, and
`egress_rates` reports the bytes each peer received per second
(/root/reference/srcs/go/monitor/counters.go:13-164).

Here gradients never touch the host, so the same API is fed from the kernels'
launch sites: every device collective records (variant, bytes) and, while
timing is on, a pair of CUDA events around the launch. Nothing is
synchronised on the hot path: events are read when statistics are asked for.

Per-GPU bytes per direction for S bytes reduced:
one-shot (n-1) S ingress, two-shot 2 (n-1)/n S, NVLS S (1 + 1/n).
"""
import time

import torch

from ..ops import _cuda as K

VARIANTS = {K.ALGO_ONESHOT: 'oneshot', K.ALGO_TWOSHOT: 'twoshot', K.ALGO_NVLS: 'nvls', 3: 'graph'}
ALGO_GRAPH = 3


def wire_bytes(algo, nbytes, np_):
    """Bytes one GPU sends (= receives) over NVLink for an all-reduce of nbytes."""
    if np_ <= 1:
        return 0.0
    if algo == K.ALGO_ONESHOT:
        return float(np_ - 1) * nbytes
    if algo == K.ALGO_NVLS:
        return nbytes * (1.0 + 1.0 / np_)
    return 2.0 * (np_ - 1) / np_ * nbytes


class VariantStat(object):
    """The reference's StrategyStat for one kernel variant."""

    def __init__(self, name):
        self.name = name
        self.acc_bytes = 0      # since the last calc()
        self.acc_seconds = 0.0  # device time since the last calc()
        self.calls = 0
        self.total_bytes = 0
        self.throughput = 0.0   # bytes / s of the last window
        self.reference = 0.0    # best window seen (the reference's `reff`)
        self.history = []

    def fold(self):
        if self.acc_seconds > 0:
            self.throughput = self.acc_bytes / self.acc_seconds
            self.reference = max(self.reference, self.throughput)
            self.history.append(self.throughput)
        self.acc_bytes, self.acc_seconds = 0, 0.0


class DeviceStats(object):
    def __init__(self, comm):
        self.comm = comm
        self.timing = False       # CUDA events around every collective (monitor=True, calc_stats, egress_rates)
        self.by_variant = {}
        self._pending = []        # (variant, bytes, ev0, ev1)
        self._egress_bytes = 0.0  # wire bytes since the last egress sample
        self._egress_t0 = time.time()
        self._last_rates = None
        self.active = None        # variant of the most recent large collective

    def _stat(self, algo):
        st = self.by_variant.get(algo)
        if st is None:
            st = VariantStat(VARIANTS.get(algo, str(algo)))
            self.by_variant[algo] = st
        return st

    def record(self, algo, nbytes, stream=None, events=None):
        st = self._stat(algo)
        st.calls += 1
        st.total_bytes += nbytes
        self._egress_bytes += wire_bytes(algo, nbytes, self.comm.np)
        self.active = algo
        if events is not None:
            self._pending.append((algo, nbytes, events[0], events[1]))

    def begin(self, stream=None):
        """Start event of a timed collective, or None when timing is off or a CUDA graph is being captured."""
        if not self.timing or torch.cuda.is_current_stream_capturing():
            return None
        ev0 = torch.cuda.Event(enable_timing=True)
        ev0.record(stream)
        return ev0

    def end(self, ev0, algo, nbytes, stream=None):
        if ev0 is None:
            self.record(algo, nbytes, stream)
            return
        ev1 = torch.cuda.Event(enable_timing=True)
        ev1.record(stream)
        self.record(algo, nbytes, stream, events=(ev0, ev1))

    def _drain(self, wait=False):
        keep = []
        for algo, nbytes, ev0, ev1 in self._pending:
            if wait:
                ev1.synchronize()
            if ev1.query():
                st = self._stat(algo)
                st.acc_bytes += nbytes
                st.acc_seconds += ev0.elapsed_time(ev1) / 1e3
            else:
                keep.append((algo, nbytes, ev0, ev1))
        self._pending = keep

    # -- the reference's API surface ---------------------------------------------------
    def calc(self):
        """calc_stats: close the current window of every variant."""
        self.timing = True
        self._drain(wait=True)
        for st in self.by_variant.values():
            st.fold()

    def lines(self):
        out = []
        for algo in sorted(self.by_variant):
            st = self.by_variant[algo]
            out.append('device variant %-8s calls %6d  total %10.1f MiB  window %8.2f GB/s  best %8.2f GB/s' %
                       (st.name, st.calls, st.total_bytes / 2**20, st.throughput / 1e9, st.reference / 1e9))
        return out

    def below_reference(self, threshold=0.8):
        """True if the active variant's last window fell under threshold x its best window."""
        st = self.by_variant.get(self.active)
        if st is None or st.reference <= 0 or st.throughput <= 0:
            return False
        return st.throughput < threshold * st.reference

    def egress_rates(self):
        """float[np]: NVLink bytes per second this GPU sent to every peer since the
        previous call (the switch spreads a variant's traffic evenly over the peers)."""
        self.timing = True
        now = time.time()
        dt = max(now - self._egress_t0, 1e-6)
        n = self.comm.np
        per_peer = self._egress_bytes / dt / max(n - 1, 1)
        self._egress_bytes, self._egress_t0 = 0.0, now
        me = max(self.comm.rank, 0)
        self._last_rates = [0.0 if r == me else per_peer for r in range(n)]
        return self._last_rates
