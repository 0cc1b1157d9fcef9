"""BASELINE.json config 5: elastic ResNet-50 whose cluster grows 1 -> 2 -> 4 -> 8
GPUs when the gradient-noise-scale monitor says larger batches still pay off.

    kungfu-run -w -builtin-config-port 9100 -np 1 -H 127.0.0.1:8 -allow-nvlink \
        python examples/elastic_resnet_gns.py --steps 200 --batch-size 128

Surviving workers are never restarted: after each resize the device layer
re-maps peer memory for the new membership, joiners receive the step counter
and the model from rank 0."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))  # run from a source checkout

import argparse
import time

import torch
import torch.nn.functional as F

import kungfu_b200.torch as kf
from kungfu_b200 import models
from kungfu_b200.optim import FusedSGD
from kungfu_b200.torch.elastic import ElasticState, warm_up
from kungfu_b200.torch.policy import NoiseScaleElasticPolicy

ap = argparse.ArgumentParser()
ap.add_argument('--model', default='resnet50')
ap.add_argument('--batch-size', type=int, default=128)
ap.add_argument('--steps', type=int, default=120)
ap.add_argument('--max-size', type=int, default=8)
ap.add_argument('--check-every', type=int, default=10)
ap.add_argument('--image-size', type=int, default=224)
ap.add_argument('--report-every', type=int, default=20)
args = ap.parse_args()

torch.cuda.set_device(kf.get_cuda_index())
dev = torch.device('cuda', torch.cuda.current_device())
torch.backends.cudnn.benchmark = True
model = models.to_b200_layout(getattr(models, args.model)(fused_bn=True).to(dev))
x = torch.randn(args.batch_size, 3, args.image_size, args.image_size, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
y = torch.randint(0, 1000, (args.batch_size, ), device=dev)
# Nothing above has touched the control plane yet (the device index comes from the launcher's environment):
# a worker that joins a running job finishes cuDNN's algorithm search on its own first, so that the first
# steps of the grown cluster do not run at the speed of its autotuning.
warm_up(model, lambda: (x, y), lambda out, tgt: F.cross_entropy(out.float(), tgt), steps=3)
opt = kf.optimizers.MonitorGradientNoiseScaleOptimizer(
    FusedSGD(model.parameters(), lr=0.01, momentum=0.9), model.named_parameters(), device_batch_size=args.batch_size)
state = ElasticState(model, opt, max_step=args.steps)
policy = NoiseScaleElasticPolicy(opt, args.batch_size, max_size=args.max_size, check_every=args.check_every,
                                 min_steps_between=2 * args.check_every)
t0, n0 = time.time(), state.trained_samples
while not state.should_stop():
    opt.zero_grad(set_to_none=True)
    loss = F.cross_entropy(model(x).float(), y)
    loss.backward()
    opt.step()
    state.end_step(args.batch_size)
    np_before = kf.current_cluster_size()
    policy.after_step(state)
    if state.stopped:
        break
    if kf.current_cluster_size() != np_before or state.step % args.report_every == 0:
        torch.cuda.synchronize()
        dt = time.time() - t0
        if kf.current_rank() == 0:
            ns = opt.noise_scale
            print('step %d np %d -> %d  %.0f img/s  noise scale %s  %s' % (
                state.step, np_before, kf.current_cluster_size(), (state.trained_samples - n0) / max(dt, 1e-9),
                'n/a' if ns is None else '%.1f' % float(ns), state.profiler.report()[-1:] or ''), flush=True)
        t0, n0 = time.time(), state.trained_samples
print('worker %x finished at step %d (cluster size %d)' % (kf.uid(), state.step, kf.current_cluster_size()))
