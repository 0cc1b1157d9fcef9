"""The device strategy executor (kungfu_b200.parallel.planner.GraphExecutor) simulated on the host: every rank is a
thread, the device barrier is a threading.Barrier, the two edge kernels act on numpy arrays. Checks, for every named
strategy and np = 2..8 (one and two hosts), that the compiled rounds leave the full sum on every rank - i.e. that the
round / depth compilation of the reference's (reduce, broadcast) graph pairs is right - without a GPU."""
import ctypes as C
import re
import threading

import numpy as np
import pytest

from kungfu_b200 import _host as H

STRATEGIES = ['STAR', 'RING', 'CLIQUE', 'TREE', 'BINARY_TREE', 'BINARY_TREE_STAR', 'MULTI_BINARY_TREE_STAR', 'AUTO']
ES = 4  # float32


def _graphs(peers, strategy):
    l = H.lib()
    l.kf_gen_strategy_graphs.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int]
    buf = C.create_string_buffer(1 << 18)
    assert l.kf_gen_strategy_graphs(peers.encode(), strategy.encode(), buf, len(buf)) >= 0, l.kf_last_error()
    out = []
    for line in buf.value.decode().strip().split('\n'):
        red, bc = line.split(' bcast=')
        edges = lambda t: [(int(a), int(b)) for a, b in re.findall(r'\((\d+)->(\d+)\)', t)]  # noqa: E731
        out.append({'loops': set(int(a) for a in re.findall(r'\((\d+)\)', red)), 'reduce': edges(red), 'bcast': edges(bc)})
    return out


class _FakeComm(object):
    def __init__(self, rank, np_, barrier):
        self.rank, self.np, self._b = rank, np_, barrier

    def barrier(self, stream=None):
        self._b.wait(timeout=30)


class _FakeBuf(object):
    """ptrs[r] = byte address of rank r's buffer inside one shared arena"""

    def __init__(self, np_, stride):
        self.ptrs = [r * stride for r in range(np_)]


def _simulate(peers, strategy, count=1000):
    from kungfu_b200.ops import collective as coll
    from kungfu_b200.ops import _cuda as K
    from kungfu_b200.parallel import planner
    np_ = len(peers.split(','))
    strategies = _graphs(peers, strategy)
    stride = (count * ES + 255) // 256 * 256
    arena = np.zeros(np_ * stride // ES, dtype=np.float32)
    for r in range(np_):
        arena[r * stride // ES:r * stride // ES + count] = np.arange(count, dtype=np.float32) * 0.001 + (r + 1)
    want = sum(np.arange(count, dtype=np.float32) * 0.001 + (r + 1) for r in range(np_))
    view = lambda ptr, n: arena[ptr // ES:ptr // ES + n]  # noqa: E731
    lock = threading.Lock()

    def edge_reduce(dst, src, n, dtype, op='sum', **_):
        with lock:
            view(dst, n)[:] += view(src, n)

    def edge_copy(dst, src, nbytes, **_):
        with lock:
            view(dst, nbytes // ES)[:] = view(src, nbytes // ES)

    barrier = threading.Barrier(np_)
    buf = _FakeBuf(np_, stride)
    errors = []

    def run(rank):
        try:
            # the edge table comes from the runtime's compiler, exactly what the device kernel is given
            ex = planner.GraphExecutor(_FakeComm(rank, np_, barrier), H.device_plan_for(peers, strategy, rank))
            ex.all_reduce(buf, count, K.F32)
        except Exception as e:  # noqa: BLE001
            errors.append((rank, repr(e)))
            barrier.abort()

    old = coll.edge_reduce, coll.edge_copy
    coll.edge_reduce, coll.edge_copy = edge_reduce, edge_copy
    try:
        ths = [threading.Thread(target=run, args=(r, )) for r in range(np_)]
        [t.start() for t in ths]
        [t.join(60) for t in ths]
    finally:
        coll.edge_reduce, coll.edge_copy = old
    assert not errors, errors
    for r in range(np_):
        got = arena[r * stride // ES:r * stride // ES + count]
        assert np.allclose(got, want, rtol=1e-6), (strategy, np_, r, got[:3], want[:3])
    return len(strategies)


@pytest.mark.parametrize('strategy', STRATEGIES)
@pytest.mark.parametrize('np_', [2, 3, 4, 5, 8])
def test_device_graph_executor_on_one_host(strategy, np_):
    peers = ','.join('127.0.0.1:%d' % (10000 + i) for i in range(np_))
    k = _simulate(peers, strategy)
    if strategy in ('RING', 'CLIQUE'):
        assert k == np_  # one rotation / root per rank: the buffer is cut into np chunks


@pytest.mark.parametrize('strategy', ['BINARY_TREE_STAR', 'MULTI_BINARY_TREE_STAR', 'STAR', 'RING', 'AUTO'])
def test_device_graph_executor_on_two_and_three_hosts(strategy):
    _simulate('10.0.0.1:10000,10.0.0.1:10001,10.0.0.2:10000,10.0.0.2:10001,10.0.0.3:10000', strategy, count=777)
    _simulate('10.0.0.1:10000,10.0.0.1:10001,10.0.0.1:10002,10.0.0.2:10000,10.0.0.2:10001,10.0.0.2:10002', strategy, count=64)


def test_variant_chooser_turns_a_probe_table_into_thresholds():
    """VariantChooser bookkeeping with a hand-written probe table (no GPU): the one-shot ceiling is the largest probe
    size up to which one-shot kept winning, the large-message preference is the most frequent winner above it."""
    from kungfu_b200.ops import _cuda as K
    from kungfu_b200.ops import collective as coll
    from kungfu_b200.parallel.planner import VariantChooser
    saved = coll.ONESHOT_MAX_BYTES, coll.PREFERRED_LARGE
    try:
        vc = VariantChooser(_FakeComm(0, 8, None))
        vc.table = {16 << 10: {K.ALGO_ONESHOT: 0.010, K.ALGO_TWOSHOT: 0.014},
                    256 << 10: {K.ALGO_ONESHOT: 0.012, K.ALGO_TWOSHOT: 0.015, K.ALGO_NVLS: 0.013},
                    2 << 20: {K.ALGO_ONESHOT: 0.060, K.ALGO_TWOSHOT: 0.030, K.ALGO_NVLS: 0.025},
                    32 << 20: {K.ALGO_TWOSHOT: 0.110, K.ALGO_NVLS: 0.075}}
        vc.best = {s: min(t, key=t.get) for s, t in vc.table.items()}
        vc._install()
        assert coll.ONESHOT_MAX_BYTES == 256 << 10
        assert coll.PREFERRED_LARGE == K.ALGO_NVLS
        assert vc.choose(20 << 10) == K.ALGO_ONESHOT and vc.choose(100 << 20) == K.ALGO_NVLS
        # busBW = 2 (n-1)/n * bytes / time: 32 MiB in 75 us on 8 ranks
        assert abs(vc.busbw(K.ALGO_NVLS) - (32 << 20) / 0.075 / 1e6 * 2 * 7 / 8) < 1e-6
        # one-shot never wins: the ceiling falls back to its 4 KiB floor
        vc.best = {s: K.ALGO_TWOSHOT for s in vc.table}
        vc._install()
        assert coll.ONESHOT_MAX_BYTES == 4 << 10 and coll.PREFERRED_LARGE == K.ALGO_TWOSHOT
    finally:
        coll.ONESHOT_MAX_BYTES, coll.PREFERRED_LARGE = saved


def test_choose_algo_decision_table():
    from kungfu_b200.ops import _cuda as K
    from kungfu_b200.ops import collective as coll

    class Buf(object):
        def __init__(self, mc):
            self.mc_ptr = mc
    saved = coll.ONESHOT_MAX_BYTES, coll.PREFERRED_LARGE
    try:
        coll.ONESHOT_MAX_BYTES, coll.PREFERRED_LARGE = 256 << 10, None
        c1, c8 = _FakeComm(0, 1, None), _FakeComm(0, 8, None)
        assert coll.choose_algo(c1, 1 << 30) == K.ALGO_ONESHOT                       # nothing to exchange
        assert coll.choose_algo(c8, 4 << 10, Buf(1)) == K.ALGO_ONESHOT               # latency-bound sizes
        assert coll.choose_algo(c8, 1 << 20, Buf(0)) == K.ALGO_TWOSHOT               # no multicast alias
        assert coll.choose_algo(c8, 1 << 20, Buf(1), K.OP_SUM, K.BF16) == K.ALGO_NVLS
        assert coll.choose_algo(c8, 1 << 20, Buf(1), K.OP_MAX, K.BF16) == K.ALGO_TWOSHOT   # the switch only adds
        assert coll.choose_algo(c8, 1 << 20, Buf(1), K.OP_SUM, K.I32) == K.ALGO_TWOSHOT    # ... floating point
        coll.PREFERRED_LARGE = K.ALGO_TWOSHOT                                         # what a probe may install
        assert coll.choose_algo(c8, 1 << 20, Buf(1), K.OP_SUM, K.BF16) == K.ALGO_TWOSHOT
        coll.PREFERRED_LARGE = K.ALGO_NVLS
        assert coll.choose_algo(c8, 1 << 20, Buf(0)) == K.ALGO_TWOSHOT               # preferred but impossible here
    finally:
        coll.ONESHOT_MAX_BYTES, coll.PREFERRED_LARGE = saved
