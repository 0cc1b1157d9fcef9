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
ap.add_argument('--fused-ln', type=int, default=0)
ap.add_argument('--pad-vocab', type=int, default=1)
ap.add_argument('--fused-grads', type=int, default=0)
ap.add_argument('--fused-linear', type=int, default=0)
args = ap.parse_args()
torch.backends.cudnn.benchmark = True
is_bert = args.model.startswith('bert')
if is_bert:
    B, S, P = (args.batch_size if args.batch_size != 256 else 32), 128, 20
    model = getattr(models, args.model)(fused_ln=bool(args.fused_ln), pad_vocab=args.pad_vocab, fused_grads=bool(args.fused_grads), fused_linear=bool(args.fused_linear)).cuda().to(torch.bfloat16)
    ids = torch.randint(0, 30522, (B, S), device='cuda')
    pos = torch.stack([torch.randperm(S, device='cuda')[:P].sort().values for _ in range(B)])
    labels = torch.gather(ids, 1, pos)
    nsp = torch.randint(0, 2, (B, ), device='cuda')
else:
    model = models.to_b200_layout(getattr(models, args.model)(fused_bn=bool(args.fused_bn)).cuda())
    x = torch.randn(args.batch_size, 3, 224, 224, device='cuda').bfloat16().contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 1000, (args.batch_size, ), device='cuda')
opt = kf.optimizers.SynchronousSGDOptimizer(FusedSGD(model.parameters(), lr=0.01, momentum=0.9), model.named_parameters())


def step():
    opt.zero_grad(set_to_none=True)
    if is_bert:
        loss = model(ids, mlm_labels=labels, nsp_labels=nsp, mlm_positions=pos)
    else:
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
rows = sorted(prof.key_averages(), key=lambda e: -e.self_device_time_total)
total = sum(e.self_device_time_total for e in rows)
print('# total device time per step: %.3f ms' % (total / 3e3))
for e in rows[:args.rows]:
    print('%8.3f ms/step %5d x  %s' % (e.self_device_time_total / 3e3, e.count // 3, e.key[:110]))
