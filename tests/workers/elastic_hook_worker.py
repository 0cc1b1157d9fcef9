"""Training loop driven by ElasticHook: after every step the hook polls the config server, follows membership
changes (progress counters by all-reduce(max), model by broadcast) and stops on detach or after the sample budget."""
import sys
import time

import torch

import kungfu_b200.torch as kf
from kungfu_b200.torch.elastic import ElasticHook

torch.manual_seed(0)
model = torch.nn.Linear(4, 2)
opt = kf.optimizers.SynchronousSGDOptimizer(torch.optim.SGD(model.parameters(), lr=0.01), model.named_parameters())
hook = ElasticHook(local_batch_size=4, epochs=1, epoch_size=int(sys.argv[1]), model=model)
sizes = set()
while True:
    sizes.add(kf.current_cluster_size())
    opt.zero_grad()
    model(torch.ones(4, 4)).sum().backward()
    opt.step()
    time.sleep(0.05)
    if not hook.after_step():
        break
if hook.exit_reason == 'finished':
    digest = kf.ops.all_reduce_fn(model.weight.detach().clone(), 'max', name='digest')
    assert torch.equal(digest, model.weight.detach()), 'replicas diverged'
print('OK elastic hook: %s after %d samples in %d steps, sizes seen %s, resizes %s' %
      (hook.exit_reason, hook.state.trained_samples, hook.state.step, sorted(sizes), hook.state.profiler.report()), flush=True)
