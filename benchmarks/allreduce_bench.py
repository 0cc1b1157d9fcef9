"""All-reduce micro-benchmark: peer-memory kernels (one-shot / two-shot / NVLS)
vs NCCL on the same GPUs. Device-timed with CUDA events, max over ranks.

    kungfu_b200/bin/kungfu-run -np 8 -allow-nvlink python benchmarks/allreduce_bench.py [--dtype bf16] [--nccl]

Reports algBW = bytes/time, busBW = 2(n-1)/n * bytes/time (nccl-tests
convention) against 900 GB/s per direction per GPU, and also the reference's
own "equivalent data rate" 4(n-1)*bytes/time
(/root/reference/srcs/python/kungfu/tensorflow/v1/benchmarks/__main__.py:145,166)
so numbers can be compared with its tools. Also runs the ResNet-50 / VGG16 /
BERT fused gradient sizes of the reference's size tables
(/root/reference/srcs/python/kungfu/tensorflow/v1/benchmarks/model_sizes.py).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import kungfu_b200.torch as kf
from kungfu_b200.ops import _cuda as K
from kungfu_b200.ops import collective as coll
from kungfu_b200.parallel import comm as pc

ap = argparse.ArgumentParser()
ap.add_argument('--dtype', default='bf16', choices=['bf16', 'f32'])
ap.add_argument('--min-bytes', type=int, default=4 << 10)
ap.add_argument('--max-bytes', type=int, default=256 << 20)
ap.add_argument('--iters', type=int, default=20)
ap.add_argument('--warmup', type=int, default=5)
ap.add_argument('--nccl', action='store_true', help='also time ncclAllReduce through torch.distributed')
ap.add_argument('--blocks', default='8,16,32,64')
ap.add_argument('--json', default='')
args = ap.parse_args()

rank, np_ = kf.current_rank(), kf.current_cluster_size()
torch.cuda.set_device(kf.get_cuda_index())
dev = torch.device('cuda', torch.cuda.current_device())
comm = pc.default_comm()
dtype = torch.bfloat16 if args.dtype == 'bf16' else torch.float32
es = 2 if args.dtype == 'bf16' else 4
code = K.dtype_code(dtype)

nccl = None
if args.nccl and np_ > 1:
    import torch.distributed as dist
    port = 23456 + (os.getpid() % 1000 if False else 0)
    dist.init_process_group('nccl', init_method='tcp://127.0.0.1:%d' % (29000 + np_), rank=rank, world_size=np_)
    nccl = dist


def time_max(fn, iters, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    kf.run_barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    mx = kf.ops.all_reduce_fn(torch.tensor([ms], dtype=torch.float64), 'max')
    return float(mx[0])


sizes = []
b = args.min_bytes
while b <= args.max_bytes:
    sizes.append(b)
    b *= 4
named = [('resnet50-fused', 25583592 * es), ('vgg16-fused', 138357544 * es), ('bert-fused', 327270150 * es)]
cap = max(sizes + [n for _, n in named])
buf = comm.alloc(cap + 4096, multicast=True)
out = comm.alloc(cap + 4096, multicast=True)
buf.view(dtype)[:cap // es].normal_()
x_nccl = torch.empty(cap // es, dtype=dtype, device=dev).normal_() if nccl else None
blocks_list = [int(x) for x in args.blocks.split(',')]
rows = []
if rank == 0:
    print('# np=%d dtype=%s multicast=%s' % (np_, args.dtype, bool(buf.mc_ptr)))
    print('# busBW = 2(n-1)/n * bytes/time (nccl-tests convention, comparable across algorithms and with NCCL);')
    print('# link GB/s = bytes one GPU really sends per direction / time, by the byte model of SURVEY.md 7.5.8:')
    print('#   one-shot (n-1) S, two-shot 2(n-1)/n S, NVLS S (1 + 1/n); "of 900" is that figure over the nominal link rate')
    print('%-16s %-9s %7s %10s %10s %10s %10s %9s' % ('size', 'algo', 'blocks', 'us', 'algBW GB/s', 'busBW GB/s', 'link GB/s', 'of 900'))
for label, nbytes in [(str(s), s) for s in sizes] + named:
    count = nbytes // es
    best = {}
    algos = [('oneshot', K.ALGO_ONESHOT), ('twoshot', K.ALGO_TWOSHOT)]
    if buf.mc_ptr:
        algos.append(('nvls', K.ALGO_NVLS))
    for name, algo in algos:
        if np_ == 1 and name != 'oneshot':
            continue
        if name == 'oneshot' and nbytes > (32 << 20):
            continue
        for blocks in blocks_list:
            def fn():
                coll.all_reduce(comm, buf, count, code, 'sum', scale=1.0 / np_, algo=algo, out=out, blocks=blocks)
            iters = args.iters if nbytes <= (64 << 20) else max(5, args.iters // 4)
            ms = time_max(fn, iters, args.warmup)
            if name not in best or ms < best[name][0]:
                best[name] = (ms, blocks)
    if nccl:
        xs = x_nccl[:count]

        def fn():
            nccl.all_reduce(xs)
        best['nccl'] = (time_max(fn, args.iters if nbytes <= (64 << 20) else 5, args.warmup), 0)
    for name, (ms, blocks) in best.items():
        alg = nbytes / ms / 1e6
        bus = alg * 2 * (np_ - 1) / max(np_, 1)
        wire = {'oneshot': float(np_ - 1), 'twoshot': 2.0 * (np_ - 1) / max(np_, 1), 'nvls': 1.0 + 1.0 / max(np_, 1)}.get(name)
        link = alg * wire if wire is not None else float('nan')
        rows.append({'size': label, 'bytes': nbytes, 'algo': name, 'blocks': blocks, 'us': ms * 1e3,
                     'algbw_gbs': alg, 'busbw_gbs': bus, 'link_gbs': link, 'link_of_900': link / 900.0,
                     'kungfu_rate_gbs': 4 * (np_ - 1) * nbytes / ms / 1e6})
        if rank == 0:
            print('%-16s %-9s %7d %10.1f %10.1f %10.1f %10.1f %8.1f%%' % (label, name, blocks, ms * 1e3, alg, bus, link,
                                                                     100 * link / 900.0))
if rank == 0 and args.json:
    with open(args.json, 'w') as f:
        json.dump({'np': np_, 'dtype': args.dtype, 'multicast': bool(buf.mc_ptr), 'rows': rows}, f, indent=1)
kf.run_barrier()
if nccl:
    nccl.destroy_process_group()
