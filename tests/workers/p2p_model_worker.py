"""Fused-model p2p ops on every worker: selectors, ModelAveraging, RequestModel and the
prefetching (double-buffered) variants, checked against values computed by hand.
Reference semantics: /root/reference/srcs/cpp/src/tensorflow/ops/cpu/peer_to_peer.cpp:72-524."""
import time

import torch

import kungfu_b200.python as kfpy
import kungfu_b200.torch as kf
from kungfu_b200.torch.ops import p2p

rank, np_ = kf.current_rank(), kf.current_cluster_size()


def model(value):
    return [('w', torch.full((3, 4), float(value))), ('b', torch.full((5, ), float(value) + 0.5))]


# ---- selectors ----
rr = p2p.PeerSelector('roundrobin', [3, 5, 7])
assert [rr.next() for _ in range(7)] == [3, 5, 7, 3, 5, 7, 3]
rnd = p2p.PeerSelector('random', [1, 2, 4], seed=0)
draws = [rnd() for _ in range(200)]
assert set(draws) == {1, 2, 4}
for bad in (lambda: p2p.PeerSelector('nearest', [1]), lambda: p2p.PeerSelector('random', [])):
    try:
        bad()
        raise SystemExit('selector accepted a bad argument')
    except ValueError:
        pass

# ---- nothing published yet: a request fails cleanly, the model is untouched ----
m = model(rank)
avg = p2p.ModelAveraging(m, 'roundrobin')
if np_ > 1:
    assert avg() == -1 and float(m[0][1][0, 0]) == rank
    assert p2p.RequestModel(m, 'roundrobin')() is None
    assert p2p.AsyncModelAveraging(m, 'roundrobin')() is False
kf.run_barrier()

# ---- ModelAveraging: round robin over the other ranks, 0.5 * (mine + theirs) ----
avg.save()  # publishes `rank` / `rank + 0.5`
kf.run_barrier()
others = [r for r in range(np_) if r != rank] or [rank]
mine = float(rank)
avg = p2p.ModelAveraging(m, 'roundrobin')
for i in range(2 * len(others)):
    t = avg()
    assert t == others[i % len(others)], (t, others)
    mine = 0.5 * (mine + t)
    assert torch.allclose(m[0][1], torch.full((3, 4), mine)) and torch.allclose(m[1][1], torch.full((5, ), 0.5 * 0 + mine + 0.5)), (m, mine)
kf.run_barrier()

# ---- RequestModel returns tensors shaped like mine, values of the selected peer ----
req = p2p.RequestModel(model(0), 'roundrobin')
got = req()
assert got[0].shape == (3, 4) and got[1].shape == (5, )
assert float(got[0][0, 0]) == others[0] and float(got[1][0]) == others[0] + 0.5

# ---- prefetching variants: the first call blocks, later calls hand out the newest complete model ----
kf.run_barrier()
areq = p2p.AsyncRequestModel(model(0), 'roundrobin', ranks=[others[0]])
got = areq()
assert float(got[0][0, 0]) == others[0]
kf.run_barrier()
# every worker publishes a new model; the background pull picks it up within a few calls
p2p.ModelAveraging(model(100 + rank)).save()
kf.run_barrier()
deadline = time.time() + 20
seen_new = False
while time.time() < deadline and not seen_new:
    got = areq()
    v = float(got[0][0, 0])
    assert v in (others[0], 100 + others[0]), v  # never a torn mixture
    assert float(got[1][0]) == v + 0.5            # both tensors come from ONE snapshot
    seen_new = v == 100 + others[0]
    time.sleep(0.01)
assert seen_new and areq.fresh >= 1, areq.fresh
areq.close()

am = model(0)
aavg = p2p.AsyncModelAveraging(am, 'roundrobin', ranks=[others[0]])
assert aavg() is True
assert float(am[0][1][0, 0]) == 0.5 * (100 + others[0])
for _ in range(50):  # many calls in flight back to back: the double buffer never hands out a partial model
    aavg()
    assert abs(float(am[1][1][0]) - float(am[0][1][0, 0]) - 0.5 * 0) <= 0.5 + 1e-6
aavg.close()
aavg.close()  # idempotent

# ---- models that are not dense fp32 take the generic path (fp32 on the wire, cast back on arrival) ----
kf.run_barrier()
mixed = [('w', torch.full((4, 6), float(rank), dtype=torch.float64).t()), ('b', torch.full((3, ), float(rank) + 0.5, dtype=torch.float16))]
assert not mixed[0][1].is_contiguous()
gen = p2p.ModelAveraging(mixed, 'roundrobin', name='mixed-model')
assert not gen._native() and avg._native()
gen.save()
kf.run_barrier()
t = gen()
assert t == others[0] and mixed[0][1].dtype == torch.float64 and mixed[1][1].dtype == torch.float16
assert torch.allclose(mixed[0][1], torch.full((6, 4), 0.5 * (rank + t), dtype=torch.float64))
assert torch.allclose(mixed[1][1].float(), torch.full((3, ), 0.5 * (rank + t) + 0.5))
agen = p2p.AsyncModelAveraging(mixed, 'roundrobin', name='mixed-model')
assert agen() is True
agen.close()
kf.run_barrier()

# ---- the raw async request handle ----
buf = torch.empty(17, dtype=torch.float32)
h = kfpy.peer().request_async(others[0], p2p.MODEL_NAME, buf.data_ptr(), 17 * 4)
assert kfpy.peer().wait(h) is True and float(buf[0]) == 100 + others[0]
h = kfpy.peer().request_async(others[0], 'no-such-blob', buf.data_ptr(), 17 * 4)
assert kfpy.peer().wait(h) is False
kf.run_barrier()
print('OK p2p-model %d/%d' % (rank, np_), flush=True)
