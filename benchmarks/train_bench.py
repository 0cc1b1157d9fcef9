"""Training-throughput benchmark for any model x optimizer wrapper of the library
(BASELINE.json configs 2-4: ResNet-50 S-SGD, DenseNet-121 SMA, BERT-base
PairAveraging). Same protocol as bench.py: synthetic data, device-timed with
CUDA events, barrier + synchronize on both sides, max over ranks.

    kungfu_b200/bin/kungfu-run -np 8 -allow-nvlink python benchmarks/train_bench.py \
        --model densenet121 --optimizer sma --batch-size 128
    python benchmarks/train_bench.py --model bert_base --optimizer pair --batch-size 32 --seq-len 128

Mirrors /root/reference/benchmarks/system/benchmark_kungfu_torch.py (and the TF
variant's --kf-optimizer {sync-sgd,async-sgd,sma,...} switch,
/root/reference/benchmarks/system/benchmark_kungfu.py).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

import kungfu_b200.torch as kf
from kungfu_b200 import models
from kungfu_b200.optim import FusedAdamW, FusedSGD

ap = argparse.ArgumentParser()
ap.add_argument('--model', default='resnet50')
ap.add_argument('--optimizer', default='ssgd', choices=['ssgd', 'sma', 'pair', 'gns', 'variance', 'ada', 'none'])
ap.add_argument('--base', default='fused', choices=['fused', 'fused-adamw', 'sgd', 'adam', 'adamw'])
ap.add_argument('--overlap', action='store_true')
ap.add_argument('--batch-size', type=int, default=256)
ap.add_argument('--seq-len', type=int, default=128)
ap.add_argument('--image-size', type=int, default=224)
ap.add_argument('--steps', type=int, default=20)
ap.add_argument('--warmup', type=int, default=5)
ap.add_argument('--fused-bn', type=int, default=1)
ap.add_argument('--fused-linear', type=int, default=0, help='BERT: tcgen05 linear layers with fused bias/GELU')
ap.add_argument('--fused-ln', type=int, default=0, help='BERT: fused dropout+residual+LayerNorm kernels')
ap.add_argument('--fused-grads', type=int, default=0, help='BERT: column-sum bias grads + one-pass qkv grad gather')
ap.add_argument('--pad-vocab', type=int, default=1, help='BERT: round the vocabulary up to a multiple of this (8 = aligned decoder GEMMs)')
ap.add_argument('--mlm-positions', type=int, default=0,
                help='BERT: masked positions per sequence fed as masked_lm_positions (original BERT uses 20 at seq 128); '
                     '0 = MLM head on every position with -100 labels')
ap.add_argument('--graph', action='store_true', help='capture the step in a CUDA graph (pair averaging: needs --pair-device-select 1)')
ap.add_argument('--pair-device-select', type=int, default=0, help='pair averaging: the kernel draws its peer on the device (graph-capturable)')
ap.add_argument('--momentum', type=float, default=0.9)
args = ap.parse_args()

rank, np_ = kf.current_rank(), kf.current_cluster_size()
torch.cuda.set_device(kf.get_cuda_index())
dev = torch.device('cuda', torch.cuda.current_device())
torch.backends.cudnn.benchmark = True
torch.manual_seed(1234 + rank)
is_bert = args.model.startswith('bert')
B = args.batch_size

if is_bert:
    from kungfu_b200.ops.fused_ln import DropoutState
    DropoutState.manual_seed(1234, rank)  # every worker draws its own dropout masks
    model = getattr(models, args.model)(fused_linear=bool(args.fused_linear), fused_ln=bool(args.fused_ln), fused_grads=bool(args.fused_grads),
                                       pad_vocab=args.pad_vocab).to(dev).to(torch.bfloat16)
    ids = torch.randint(0, 30522, (B, args.seq_len), device=dev)
    labels = ids.clone()
    labels[torch.rand(B, args.seq_len, device=dev) > 0.15] = -100
    nsp = torch.randint(0, 2, (B, ), device=dev)
    unit = 'seq/s'

    mlm_pos = None
    if args.mlm_positions > 0:
        mlm_pos = torch.stack([torch.randperm(args.seq_len, device=dev)[:args.mlm_positions].sort().values for _ in range(B)])
        labels = torch.gather(ids, 1, mlm_pos)

    def loss_fn():
        return model(ids, mlm_labels=labels, nsp_labels=nsp, mlm_positions=mlm_pos)
else:
    if hasattr(models, args.model):
        kw = {'fused_bn': bool(args.fused_bn)} if not args.model.startswith('vgg') else {}
        net = getattr(models, args.model)(**kw)
    else:
        # any other torchvision architecture by name, as the reference's benchmark accepts
        # (benchmark_kungfu_torch.py: getattr(torchvision.models, args.model)()); random init, stock torch layers
        import torchvision
        net = getattr(torchvision.models, args.model)(weights=None)
    model = models.to_b200_layout(net.to(dev), torch.bfloat16)
    x = torch.randn(B, 3, args.image_size, args.image_size, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 1000, (B, ), device=dev)
    unit = 'img/s'

    def loss_fn():
        return F.cross_entropy(model(x).float(), y)
model.train()

params = list(model.parameters())
if args.base == 'fused':
    base = FusedSGD(params, lr=0.01, momentum=args.momentum)
elif args.base == 'fused-adamw':
    base = FusedAdamW(params, lr=1e-4, weight_decay=0.01)
elif args.base == 'sgd':
    base = torch.optim.SGD(params, lr=0.01, momentum=args.momentum)
elif args.base == 'adamw':
    base = torch.optim.AdamW(params, lr=1e-4, weight_decay=0.01)
else:
    base = torch.optim.Adam(params, lr=1e-4)
named = list(model.named_parameters())
O = kf.optimizers
opt = {
    'ssgd': lambda: O.SynchronousSGDOptimizer(base, named, overlap=args.overlap),
    'sma': lambda: O.SynchronousAveragingOptimizer(base, named),
    'pair': lambda: O.PairAveragingOptimizer(base, named),
    'gns': lambda: O.MonitorGradientNoiseScaleOptimizer(base, named, device_batch_size=B),
    'variance': lambda: O.MonitorGradientVarianceOptimizer(base, named),
    'ada': lambda: O.AdaptiveSGDOptimizer(base, named, change_step=args.warmup + args.steps // 2),
    'none': lambda: base,
}[args.optimizer]()
if args.optimizer == 'pair':
    opt.device_select = bool(args.pair_device_select)
if hasattr(opt, 'arenas'):
    opt.arenas()
kf.broadcast_parameters(model.state_dict())


def step():
    opt.zero_grad(set_to_none=True)
    loss = loss_fn()
    loss.backward()
    opt.step()
    return loss


if args.graph and (args.optimizer != 'pair' or args.pair_device_select):
    # warm up on a side stream: autograd must not tie the capture to the legacy stream
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(args.warmup):
            loss = step()
    torch.cuda.current_stream().wait_stream(side)
else:
    for _ in range(args.warmup):
        loss = step()
torch.cuda.synchronize()
run = step
if args.graph and (args.optimizer != 'pair' or args.pair_device_select):
    g = torch.cuda.CUDAGraph()
    kf.run_barrier()
    with torch.cuda.graph(g):
        step()
    run = g.replay
    for _ in range(3):
        run()
    torch.cuda.synchronize()
kf.run_barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(args.steps):
    run()
e1.record()
torch.cuda.synchronize()
kf.run_barrier()
ms = float(kf.ops.all_reduce_fn(torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64), 'max')[0]) / args.steps
extra = {}
if args.optimizer == 'gns' and getattr(opt, 'noise_scale', None) is not None:
    extra['noise_scale'] = float(opt.noise_scale)
if args.optimizer == 'pair' and getattr(opt, 'torn_reads', None) is not None:
    extra['torn_reads'] = int(opt.torn_reads)
if rank == 0:
    print('RESULT: ' + json.dumps(dict({
        'model': args.model, 'optimizer': args.optimizer, 'base': args.base, 'n_gpus': np_, 'per_gpu_batch': B,
        'ms_per_step': ms, 'value': B * np_ / (ms / 1e3), 'unit': unit, 'cuda_graph': bool(args.graph),
        'loss': float(loss.detach())}, **extra)))
kf.run_barrier()
