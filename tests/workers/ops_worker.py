"""Collective / p2p / adaptation checks run on every worker (CPU tensors).
Mirrors /root/reference/tests/python/integration/test_operators.py,
/root/reference/tests/cpp/integration/fake_agent.cpp:15-205 and
/root/reference/tests/go/cmd/kungfu-test-public-apis."""
import random
import sys

import torch

import kungfu_b200.python as kfpy
import kungfu_b200.torch as kf

rank, np_ = kf.current_rank(), kf.current_cluster_size()


def check_all_reduce():
    for dt in (torch.float32, torch.float64, torch.int32, torch.int64, torch.float16, torch.bfloat16, torch.uint8):
        x = (torch.arange(20) % 3 + 1).to(dt) if dt == torch.uint8 else (torch.arange(20) + rank).to(dt)
        y = kf.ops.all_reduce_fn(x, 'sum')
        exp = (torch.arange(20) % 3 + 1).double() * np_ if dt == torch.uint8 else sum((torch.arange(20) + r).double() for r in range(np_))
        # 16-bit sums are rounded at every hop of the graph walk: the bound grows with the length of the chain
        assert torch.allclose(y.double(), exp, rtol=1e-2 * max(1, np_ / 4)), (dt, y, exp)
    x = torch.tensor([float(rank + 1)] * 5)
    assert kf.ops.all_reduce_fn(x, 'max')[0] == np_
    assert kf.ops.all_reduce_fn(x, 'min')[0] == 1
    prod = 1.0
    for r in range(np_):
        prod *= r + 1
    assert abs(float(kf.ops.all_reduce_fn(x, 'prod')[0]) - prod) <= 1e-6 * prod  # np! leaves fp32's exact range at np = 14
    # larger than the 1 MiB chunk: exercises chunking + strategy hashing
    big = torch.full((700000, ), float(rank + 1))
    out = kf.ops.all_reduce_fn(big)
    assert torch.all(out == np_ * (np_ + 1) / 2)
    # async, many in flight, in place
    ts = [torch.full((1000 + i, ), float(rank)) for i in range(20)]
    hs = [kf.ops.inplace_all_reduce_async_op(t, 'async%d' % i) for i, t in enumerate(ts)]
    kf.ops.wait_all_handles(hs)
    for t in ts:
        assert torch.all(t == sum(range(np_)))


def check_gather_broadcast_consensus():
    x = torch.arange(6, dtype=torch.float32).view(2, 3) + 10 * rank
    y = kf.ops.all_gather(x)
    assert y.shape == (np_, 2, 3)
    for r in range(np_):
        assert torch.equal(y[r], torch.arange(6, dtype=torch.float32).view(2, 3) + 10 * r)
    b = torch.full((17, ), float(rank))
    kf.ops.broadcast(b)
    assert torch.all(b == 0)
    sd = {'w': torch.full((3, 3), float(rank)), 'n': torch.tensor(rank)}
    kf.broadcast_parameters(sd)
    assert torch.all(sd['w'] == 0) and int(sd['n']) == 0
    assert kf.ops.consensus(torch.tensor([1, 2, 3]))
    assert kf.ops.consensus(torch.tensor([rank])) == (np_ == 1)
    kf.run_barrier()


def check_p2p():
    # versioned save + request, and an expected failure for an unknown name
    x = torch.full((100, ), float(rank))
    kf.ops.save_variable(x, 'model')
    kf.ops.save_variable(x + 100, 'model', version=7)
    kf.run_barrier()
    target = (rank + 1) % np_
    ok, got = kf.ops.request_variable(target, 'model', x)
    assert ok and torch.all(got == target)
    ok, got = kf.ops.request_variable(target, 'model', x, version=7)
    assert ok and torch.all(got == target + 100)
    ok, _ = kf.ops.request_variable(target, 'model!', x)
    assert not ok
    ok, _ = kf.ops.request_variable(target, 'model', x, version=8)
    assert not ok
    kf.run_barrier()


def check_set_tree_and_monitoring():
    rng = random.Random(42)  # same trees on every rank
    for _ in range(6):
        # random tree as a father array: father[i] < i after a random relabelling
        perm = list(range(np_))
        rng.shuffle(perm)
        father = [0] * np_
        father[perm[0]] = perm[0]
        for i in range(1, np_):
            father[perm[i]] = perm[rng.randrange(i)]
        kfpy.set_tree(father)
        x = torch.ones(300000)
        y = kf.ops.monitored_all_reduce(x)
        assert torch.all(y == np_)
    y = kf.ops.monitored_all_reduce(torch.ones(10), tree=[0] * np_)
    assert torch.all(y == np_)
    kfpy.calc_stats()
    kfpy.log_stats()
    assert kfpy.check_interference() in (False, True)
    for s in ('RING', 'BINARY_TREE', 'STAR'):
        kfpy.set_strategy(s)
        assert torch.all(kf.ops.all_reduce_fn(torch.ones(5000)) == np_)
    lat = kfpy.get_peer_latencies()
    assert len(lat) == np_ and all(l >= 0 for l in lat)
    assert len(kfpy.egress_rates()) == np_


check_all_reduce()
check_gather_broadcast_consensus()
check_p2p()
check_set_tree_and_monitoring()
kf.run_barrier()
print('OK rank %d/%d' % (rank, np_))
