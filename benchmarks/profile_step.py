"""Kernel-time breakdown of one training step (torch.profiler, CUDA activity).
Usage: python benchmarks/profile_step.py [--fused-bn 1] [--batch-size 256]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

import kungfu_b200.torch as kf
from kungfu_b200 import models
from kungfu_b200.optim import FusedSGD

ap = argparse.ArgumentParser()
ap.add_argument('--fused-bn', type=int, default=0)
ap.add_argument('--batch-size', type=int, default=256)
ap.add_argument('--model', default='resnet50')
ap.add_argument('--rows', type=int, default=40)
args = ap.parse_args()
torch.backends.cudnn.benchmark = True
model = models.to_b200_layout(getattr(models, args.model)(fused_bn=bool(args.fused_bn)).cuda())
opt = kf.optimizers.SynchronousSGDOptimizer(FusedSGD(model.parameters(), lr=0.01), model.named_parameters())
x = torch.randn(args.batch_size, 3, 224, 224, device='cuda').bfloat16().contiguous(memory_format=torch.channels_last)
y = torch.randint(0, 1000, (args.batch_size, ), device='cuda')


def step():
    opt.zero_grad(set_to_none=True)
    loss = F.cross_entropy(model(x).float(), y)
    loss.backward()
    opt.step()


for _ in range(5):
    step()
torch.cuda.synchronize()
with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by='cuda_time_total', row_limit=args.rows, max_name_column_width=90))
