"""Conventional baselines on the same GPUs: stand-ins for
"the reference's own NCCL build", which cannot be built offline.

  ref-nccl-unfused  one ncclAllReduce(sum) per gradient tensor in fixed order on a
                    dedicated stream, each followed by a stream synchronize, then a
                    separate g/np kernel; after backward, no overlap
                    (/root/reference/srcs/cpp/src/nccl/gpu_collective.cpp:118-127,
                     /root/reference/srcs/cpp/src/nccl/scheduler.cpp:9-36,
                     /root/reference/srcs/python/kungfu/tensorflow/optimizers/sync_sgd.py:94-104)
  ref-nccl-fused    same, but one ncclAllReduce on the flattened gradient (nccl_fusion)
                    (sync_sgd.py:87-92)
  ddp               torch.nn.parallel.DistributedDataParallel (bucketed, overlapped)
  ref-host-staged   D2H -> CPU all-reduce -> H2D per tensor: the reference's actual
                    torch path (/root/reference/srcs/cpp/src/torch/ops/cuda/collective.cpp:31-54)

All use stock PyTorch ops for the model (torch BN/ReLU, torch.optim.SGD); `--our-model 1`
swaps in this repo's fused model kernels to isolate the communication difference.

    kungfu_b200/bin/kungfu-run -np 8 -allow-nvlink python benchmarks/baselines.py --mode ddp
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import torch.nn.functional as F

import kungfu_b200.torch as kf
from kungfu_b200 import models

ap = argparse.ArgumentParser()
ap.add_argument('--mode', default='ddp', choices=['ddp', 'ref-nccl-unfused', 'ref-nccl-fused', 'ref-host-staged', 'single'])
ap.add_argument('--model', default='resnet50')
ap.add_argument('--batch-size', type=int, default=256)
ap.add_argument('--steps', type=int, default=20)
ap.add_argument('--warmup', type=int, default=5)
ap.add_argument('--our-model', type=int, default=0)
ap.add_argument('--dtype', default='bf16', choices=['bf16', 'fp32'])
args = ap.parse_args()

rank, np_ = kf.current_rank(), kf.current_cluster_size()
torch.cuda.set_device(kf.get_cuda_index())
dev = torch.device('cuda', torch.cuda.current_device())
torch.backends.cudnn.benchmark = True
if np_ > 1 and args.mode != 'ref-host-staged':
    dist.init_process_group('nccl', init_method='tcp://127.0.0.1:%d' % (29100 + np_), rank=rank, world_size=np_)

torch.manual_seed(1234)
model = getattr(models, args.model)(fused_bn=bool(args.our_model)).to(dev)
if args.dtype == 'bf16':
    model = models.to_b200_layout(model, torch.bfloat16)
    x = torch.randn(args.batch_size, 3, 224, 224, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
else:
    model = model.to(memory_format=torch.channels_last)
    x = torch.randn(args.batch_size, 3, 224, 224, device=dev).contiguous(memory_format=torch.channels_last)
y = torch.randint(0, 1000, (args.batch_size, ), device=dev)
model.train()
opt = torch.optim.SGD(model.parameters(), lr=0.01 * np_)
net = model
if args.mode == 'ddp' and np_ > 1:
    net = torch.nn.parallel.DistributedDataParallel(model, device_ids=[dev.index], gradient_as_bucket_view=True)
comm_stream = torch.cuda.Stream()
params = [p for p in model.parameters() if p.requires_grad]


def sync_grads():
    if np_ == 1 or args.mode in ('ddp', 'single'):
        return
    if args.mode == 'ref-nccl-unfused':
        comm_stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(comm_stream):
            for p in params:
                dist.all_reduce(p.grad)
                comm_stream.synchronize()  # the reference blocks after every NCCL call
        torch.cuda.current_stream().wait_stream(comm_stream)
        for p in params:
            p.grad.div_(np_)
    elif args.mode == 'ref-nccl-fused':
        flat = torch.cat([p.grad.reshape(-1).float() for p in params])
        comm_stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(comm_stream):
            dist.all_reduce(flat)
            comm_stream.synchronize()
        torch.cuda.current_stream().wait_stream(comm_stream)
        flat.div_(np_)
        off = 0
        for p in params:
            n = p.numel()
            p.grad.copy_(flat[off:off + n].view_as(p.grad))
            off += n
    elif args.mode == 'ref-host-staged':
        for i, p in enumerate(params):
            h = p.grad.float().cpu()  # pageable staging buffer, like `new char[size]`
            kf.ops.inplace_all_reduce_op(h.view(-1), 'sum', name='g%d' % i)
            p.grad.copy_(h.to(dev) / np_)


def step():
    opt.zero_grad(set_to_none=True)
    loss = F.cross_entropy(net(x).float(), y)
    loss.backward()
    sync_grads()
    opt.step()
    return loss


for _ in range(args.warmup):
    loss = step()
torch.cuda.synchronize()
kf.run_barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(args.steps):
    step()
e1.record()
torch.cuda.synchronize()
kf.run_barrier()
ms = float(kf.ops.all_reduce_fn(torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64), 'max')[0]) / args.steps
if rank == 0:
    print('RESULT: ' + json.dumps({'mode': args.mode, 'model': args.model, 'n_gpus': np_, 'per_gpu_batch': args.batch_size,
                                   'our_model_kernels': bool(args.our_model), 'dtype': args.dtype, 'ms_per_step': ms,
                                   'value': args.batch_size * np_ / (ms / 1e3), 'unit': 'img/s'}))
kf.run_barrier()
if dist.is_initialized():
    dist.destroy_process_group()
