#!/usr/bin/env python
"""Elastic MNIST written as a plain eager training loop: the loop itself calls
resize(new_size) at chosen steps and, when the membership changed, re-syncs the
step counter and the model before it goes on.

    kungfu-run -w -np 1 -H 127.0.0.1:4 python examples/mnist_elastic_eager.py

Reference: /root/reference/examples/mnist_elastic_eager/main.py:84-116 (TF2
eager loop; `resize`, then all-reduce(max) of the step and a broadcast of the
variables).
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))  # run from a source checkout

import torch
import torch.nn.functional as F

import kungfu_b200.torch as kf
from kungfu_b200.datasets import ElasticShardedDataset, load_mnist
from kungfu_b200.models import SLP

GLOBAL_BATCH = 48
MAX_STEP = 40
SCHEDULE = {8: 2, 16: 4, 24: 3, 32: 1}  # step -> new cluster size

x, y, xt, yt = load_mnist(n_train=4800, n_test=500)
model = SLP(zero_init=False)
opt = kf.optimizers.SynchronousSGDOptimizer(torch.optim.SGD(model.parameters(), lr=0.1), model.named_parameters())


def sync(step):
    """What a worker does after every membership change (and when it joins)."""
    step = int(kf.ops.all_reduce_fn(torch.tensor([step], dtype=torch.int64), 'max', name='sync-step')[0])
    kf.broadcast_parameters(model.state_dict())
    return step


step = sync(0)
data = ElasticShardedDataset(x, y, GLOBAL_BATCH, kf.current_rank(), kf.current_cluster_size(), offset=step * GLOBAL_BATCH)
while step < MAX_STEP:
    np_ = kf.current_cluster_size()
    xb, yb = data.next_batch()
    opt.zero_grad()
    # each worker holds 1/np of the global batch and the optimizer averages gradients over workers
    loss = F.cross_entropy(model(xb), yb, reduction='sum') / GLOBAL_BATCH * np_
    loss.backward()
    opt.step()
    step += 1
    if step in SCHEDULE and SCHEDULE[step] != np_:
        changed, detached = kf.ops.resize(SCHEDULE[step])
        if detached:
            print('worker %x detached at step %d' % (kf.uid(), step), flush=True)
            sys.exit(0)
        if changed:
            step = sync(step)
            data.reshard(kf.current_rank(), kf.current_cluster_size(), offset=step * GLOBAL_BATCH)
            if kf.current_rank() == 0:
                print('step %d: cluster is now %d workers' % (step, kf.current_cluster_size()), flush=True)

with torch.no_grad():
    acc = float((model(xt).argmax(1) == yt).float().mean())
print('worker %x finished at step %d with np=%d, test accuracy %.3f' % (kf.uid(), step, kf.current_cluster_size(), acc),
      flush=True)
