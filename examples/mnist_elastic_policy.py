"""Elastic MNIST driven by a policy (resize at fixed steps).

    kungfu-run -w -builtin-config-port 9100 -np 1 -H 127.0.0.1:4 python examples/mnist_elastic_policy.py

Reference: /root/reference/examples/mnist_elastic_policy/{main.py,elastic_policy.py,run.sh}
(resizes at steps {10:2, 20:3, 30:4, 40:1})."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))  # run from a source checkout

import torch
import torch.nn.functional as F

import kungfu_b200.torch as kf
from kungfu_b200.datasets import ElasticShardedDataset, load_mnist
from kungfu_b200.models import SLP
from kungfu_b200.torch.policy import PolicyRunner, ScheduledElasticPolicy

BATCH = 48  # global batch, divisible by 1, 2, 3, 4
x, y, _, _ = load_mnist(n_train=4800, n_test=10)
model = SLP(zero_init=False)
opt = kf.optimizers.SynchronousSGDOptimizer(torch.optim.SGD(model.parameters(), lr=0.1), model.named_parameters())
runner = PolicyRunner([ScheduledElasticPolicy({10: 2, 20: 3, 30: 4, 40: 1})], epoch_size=2400, epoch_num=1,
                      model=model, batch_size=BATCH // kf.current_cluster_size())
data = ElasticShardedDataset(x, y, BATCH, kf.current_rank(), kf.current_cluster_size(),
                             offset=runner.state.trained_samples)
while True:
    np_ = kf.current_cluster_size()
    data.reshard(kf.current_rank(), np_, offset=runner.state.step * BATCH)
    xb, yb = data.next_batch()
    runner.before_step()
    opt.zero_grad()
    loss = F.cross_entropy(model(xb), yb, reduction='sum') / BATCH * np_
    loss.backward()
    opt.step()
    from kungfu_b200 import variables as V
    V.set_value(V.GraphKeys.BATCH_SIZE, BATCH // np_)
    if kf.current_rank() == 0 and runner.state.step % 5 == 0:
        print('step %d np %d loss %.4f' % (runner.state.step, np_, float(loss) / np_))
    if not runner.after_step():
        break
print('worker %x done at step %d, resizes: %s' % (kf.uid(), runner.state.step, runner.state.profiler.report()))
