"""Elastic training loop: follows a size schedule with resize_cluster; joiners
pick up the step counter and the model from the survivors (live state
transfer, no restart of surviving workers). Mirrors
/root/reference/benchmarks/adaptation/adaptive_trainer.py and
/root/reference/srcs/python/kungfu/tensorflow/hooks/elastic.py:40-70."""
import os
import sys

import torch

import kungfu_b200.torch as kf
from kungfu_b200.utils import step_based_schedule

schedule = sys.argv[1]           # e.g. "1:3,2:3,4:3,2:3,1:3"
max_step = int(sys.argv[2])
log_path = sys.argv[3]

model = torch.nn.Linear(8, 4)
torch.manual_seed(kf.uid() % 1000)  # joiners start with DIFFERENT weights on purpose
torch.nn.init.normal_(model.weight)
opt = kf.optimizers.SynchronousSGDOptimizer(torch.optim.SGD(model.parameters(), lr=0.01), model.named_parameters())
pid = os.getpid()


def sync_state(step):
    step = int(kf.ops.all_reduce_fn(torch.tensor([step], dtype=torch.int64), 'max')[0])
    kf.broadcast_parameters(model.state_dict())
    return step


step = sync_state(0)
while step < max_step:
    np_ = kf.current_cluster_size()
    x = torch.ones(2, 8)
    opt.zero_grad()
    model(x).sum().backward()
    opt.step()
    # replicas must be identical at every step, whatever the membership history
    ws = kf.ops.all_gather(model.weight.data)
    assert torch.allclose(ws[0], ws[-1]), 'replicas diverged at step %d' % step
    with open(log_path, 'a') as f:
        f.write('%d %d %d %d\n' % (step, kf.current_rank(), np_, pid))
    step += 1
    want = step_based_schedule(schedule, step, default=np_, strict=False)
    if want != np_:
        changed, detached = kf.resize_cluster(want)
        if detached:
            with open(log_path, 'a') as f:
                f.write('detached %d %d\n' % (step, pid))
            sys.exit(0)
        if changed:
            step = sync_state(step)
print('OK elastic worker rank %d done at step %d' % (kf.current_rank(), step))
