"""Elastic training on GPUs: FusedSGD S-SGD over peer memory while the cluster
follows a size schedule. After every step all replicas must hold identical
parameters; surviving processes are never restarted (their pids persist)."""
import os
import sys

import torch
import torch.nn.functional as F

import kungfu_b200.torch as kf
from kungfu_b200 import models
from kungfu_b200.optim import FusedSGD
from kungfu_b200.torch.elastic import ElasticTrainHook

schedule, max_step, log_path = sys.argv[1], int(sys.argv[2]), sys.argv[3]
torch.cuda.set_device(kf.get_cuda_index())
dev = torch.device('cuda', torch.cuda.current_device())
torch.manual_seed(kf.uid() % 1000)  # joiners start from different weights on purpose
model = models.to_b200_layout(models.resnet18(num_classes=10, fused_bn=True).to(dev))
opt = kf.optimizers.SynchronousSGDOptimizer(FusedSGD(model.parameters(), lr=0.05, momentum=0.9), model.named_parameters())
hook = ElasticTrainHook(schedule, max_step, model, opt)
g = torch.Generator().manual_seed(5)
x = torch.randn(8, 3, 32, 32, generator=g).to(dev).bfloat16().contiguous(memory_format=torch.channels_last)
y = torch.randint(0, 10, (8, ), generator=g).to(dev)
while True:
    np_ = kf.current_cluster_size()
    opt.zero_grad(set_to_none=True)
    loss = F.cross_entropy(model(x).float(), y)
    loss.backward()
    opt.step()
    torch.cuda.synchronize()
    w = kf.ops.all_gather(model.fc.weight.data.float())
    assert torch.isfinite(loss) and torch.equal(w[0], w[-1]), 'replicas diverged at step %d' % hook.state.step
    with open(log_path, 'a') as f:
        f.write('%d %d %d %d %.4f\n' % (hook.state.step, kf.current_rank(), np_, os.getpid(), float(loss)))
    if os.environ.get('KUNGFU_TEST_MEMLOG') and kf.current_rank() == 0:
        # device memory the driver considers free, as the surviving rank 0 sees it after this step
        with open(os.environ['KUNGFU_TEST_MEMLOG'], 'a') as f:
            f.write('%d %d %d\n' % (hook.state.step, np_, torch.cuda.mem_get_info()[0]))
    if not hook.after_step(8):
        break
print('OK gpu elastic worker pid %d done at step %d np %d' % (os.getpid(), hook.state.step, kf.current_cluster_size()))
