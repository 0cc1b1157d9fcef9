"""Device traffic statistics (kungfu_b200.parallel.devstats) without a GPU: the byte model of the variants, the
window bookkeeping of calc_stats and the 80 % rule of the interference vote
(/root/reference/srcs/go/kungfu/session/adaptiveStrategies.go:61-121)."""
import pytest

from kungfu_b200.ops import _cuda as K
from kungfu_b200.parallel import devstats as ds


class _Comm(object):
    def __init__(self, np_, rank=0):
        self.np, self.rank = np_, rank


def test_wire_bytes_follow_the_survey_byte_model():
    S = 1 << 20
    assert ds.wire_bytes(K.ALGO_ONESHOT, S, 8) == 7 * S          # every rank pulls everybody's buffer
    assert ds.wire_bytes(K.ALGO_TWOSHOT, S, 8) == 2 * 7 / 8 * S  # reduce-scatter + all-gather
    assert ds.wire_bytes(K.ALGO_NVLS, S, 8) == S * (1 + 1 / 8)   # the switch reduces and multicasts
    assert ds.wire_bytes(K.ALGO_NVLS, S, 1) == 0.0
    # NVLS moves less than two-shot from 4 ranks on, more at 2: what the probe finds on real hardware
    assert ds.wire_bytes(K.ALGO_NVLS, S, 2) > ds.wire_bytes(K.ALGO_TWOSHOT, S, 2)
    assert ds.wire_bytes(K.ALGO_NVLS, S, 4) < ds.wire_bytes(K.ALGO_TWOSHOT, S, 4)


def test_windows_reference_and_vote():
    st = ds.DeviceStats(_Comm(8, rank=3))
    v = st._stat(K.ALGO_NVLS)
    # window 1: 10 x 50 MB in 1 ms
    v.acc_bytes, v.acc_seconds = 10 * 50e6, 1e-3
    st.active = K.ALGO_NVLS
    for s in st.by_variant.values():
        s.fold()
    assert v.throughput == pytest.approx(5e11) and v.reference == pytest.approx(5e11)
    assert not st.below_reference(0.8)
    # window 2: same bytes, 1.2 ms (83 % of the best window): still fine
    v.acc_bytes, v.acc_seconds = 10 * 50e6, 1.2e-3
    v.fold()
    assert not st.below_reference(0.8) and v.reference == pytest.approx(5e11)
    # window 3: 2 ms (50 %): a vote
    v.acc_bytes, v.acc_seconds = 10 * 50e6, 2e-3
    v.fold()
    assert st.below_reference(0.8)
    # an empty window does not change anything
    v.fold()
    assert v.throughput == pytest.approx(2.5e11) and len(v.history) == 3
    assert any('nvls' in line for line in st.lines())


def test_egress_rates_spread_the_wire_bytes_over_the_peers():
    st = ds.DeviceStats(_Comm(4, rank=1))
    st.record(K.ALGO_TWOSHOT, 8 << 20)
    st.record(K.ALGO_NVLS, 8 << 20)
    st._egress_t0 -= 1.0  # pretend one second has passed
    rates = st.egress_rates()
    assert len(rates) == 4 and rates[1] == 0.0
    total = ds.wire_bytes(K.ALGO_TWOSHOT, 8 << 20, 4) + ds.wire_bytes(K.ALGO_NVLS, 8 << 20, 4)
    assert sum(rates) == pytest.approx(total, rel=0.05)
    assert st.egress_rates()[0] == 0.0  # counters restart with every sample
