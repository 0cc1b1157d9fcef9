"""Launch order of the overlapped, bucketed fused step (kungfu_b200.optim.fused.FusedOptimizer._progress), driven
on the host with recording stand-ins for the two launches of a bucket. The invariants the device protocol relies on:

* a bucket is gathered + signalled only when all its gradients exist, buckets in bucket order;
* the exchange kernel of bucket b is launched after its own signal and BEFORE bucket b+1 is signalled: the arrival
  flags are epoch counters per CTA row, a wait-only kernel waits for "the current epoch", so a later signal in
  between would make it wait for the wrong one;
* the exchange is deferred by a few hooks (or until the next bucket is complete), never dropped: step() forces
  whatever is left, in order.
"""
import random

import pytest

from kungfu_b200.optim.fused import FusedSGD, _Bucket


class _P(object):
    pass


def _make(n_buckets, per_bucket, defer):
    opt = object.__new__(FusedSGD)
    params, buckets, where = [], [], {}
    for b in range(n_buckets):
        idx = list(range(b * per_bucket, (b + 1) * per_bucket))
        bk = _Bucket(None, None, idx, 0, 0, b)
        buckets.append(bk)
        for _ in idx:
            p = _P()
            params.append(p)
            where[id(p)] = bk
    opt._ov = {'buckets': buckets, 'where': where, 'fused': 0, 'stepped': 0, 'since': 0, 'armed': True}
    opt.defer_hooks = defer
    events = []
    opt._fuse_bucket = lambda b: events.append(('fuse', b.slot))
    opt._step_bucket = lambda b: events.append(('step', b.slot))
    return opt, params, events


def _check(events, n_buckets):
    assert [e for e in events if e[0] == 'fuse'] == [('fuse', b) for b in range(n_buckets)]
    assert [e for e in events if e[0] == 'step'] == [('step', b) for b in range(n_buckets)]
    for b in range(n_buckets):
        assert events.index(('fuse', b)) < events.index(('step', b))
        if b + 1 < n_buckets:
            assert events.index(('step', b)) < events.index(('fuse', b + 1)), 'bucket %d stepped after the next signal' % b


@pytest.mark.parametrize('defer', [0, 3, 1000])
def test_in_order_gradients(defer):
    opt, params, events = _make(4, 5, defer)
    for p in params:  # autograd delivers the buckets' gradients in bucket order
        opt._on_grad(p)
    # nothing may be gathered before its bucket is complete
    assert ('fuse', 3) in events
    opt._progress(force=True)
    _check(events, 4)


def test_exchange_is_deferred_by_hooks_not_launched_right_behind_its_signal():
    opt, params, events = _make(2, 8, 3)
    for p in params[:8]:
        opt._on_grad(p)
    assert events == [('fuse', 0)]          # signalled, exchange not yet launched
    opt._on_grad(params[8])
    opt._on_grad(params[9])
    assert events == [('fuse', 0)]
    opt._on_grad(params[10])                # third hook after the signal
    assert events == [('fuse', 0), ('step', 0)]
    for p in params[11:]:
        opt._on_grad(p)
    assert events == [('fuse', 0), ('step', 0), ('fuse', 1)]
    opt._progress(force=True)               # step(): the last bucket's exchange is exposed
    _check(events, 2)


@pytest.mark.parametrize('seed', range(20))
def test_out_of_order_gradients_and_unused_parameters(seed):
    rng = random.Random(seed)
    nb, per = rng.randint(1, 6), rng.randint(1, 7)
    opt, params, events = _make(nb, per, rng.choice([0, 1, 2, 6]))
    order = list(params)
    rng.shuffle(order)
    unused = set(id(p) for p in rng.sample(order, k=rng.randint(0, len(order) // 3)))
    for p in order:
        if id(p) not in unused:           # parameters without a gradient never fire their hook
            opt._on_grad(p)
    opt._progress(force=True)
    _check(events, nb)
