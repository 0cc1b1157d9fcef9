"""Built-in all-reduce benchmark over a model's gradient sizes.

    kungfu-run -np 4 python -m kungfu_b200.benchmarks --method CPU --model ResNet50 [--fuse]
    kungfu-run -np 8 -allow-nvlink python -m kungfu_b200.benchmarks --method P2P --model VGG16 --fuse

--method CPU   host control-plane collectives (graph strategies over sockets)
--method P2P   peer-memory kernels on the GPUs (the product path)
--method NCCL  ncclAllReduce through torch.distributed (baseline)
Reference: python -m kungfu.tensorflow.v1.benchmarks
(/root/reference/srcs/python/kungfu/tensorflow/v1/benchmarks/__main__.py:135-171); the
"equivalent data rate" 4(np-1)*bytes/time it prints is reported alongside busBW.
"""
import argparse
import time

import torch

import kungfu_b200.torch as kf
from kungfu_b200.benchmarks.model_sizes import model_sizes


def main():
    ap = argparse.ArgumentParser()
    # the reference's method names are accepted too: NCCL+CPU (its hierarchical NCCL + TCP path) is what the device
    # communicator does by itself when a job spans hosts, i.e. P2P here; HOROVOD is not in this image
    ap.add_argument('--method', default='CPU', choices=['CPU', 'P2P', 'NCCL', 'NCCL+CPU', 'HOROVOD'])
    ap.add_argument('--model', default='ResNet50')
    ap.add_argument('--fuse', action='store_true')
    ap.add_argument('--warmup-steps', type=int, default=3)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--dtype', default='float32', choices=['float32', 'bfloat16'])
    args = ap.parse_args()
    if args.method == 'HOROVOD':
        raise SystemExit('--method HOROVOD: Horovod is not installed in this environment (use NCCL for the library baseline)')
    if args.method == 'NCCL+CPU':
        args.method = 'P2P'
    rank, np_ = kf.current_rank(), kf.current_cluster_size()
    sizes = model_sizes(args.model)
    if args.fuse:
        sizes = [sum(sizes)]
    dtype = getattr(torch, args.dtype)
    on_gpu = args.method in ('P2P', 'NCCL')
    if on_gpu:
        torch.cuda.set_device(kf.get_cuda_index())
    dev = torch.device('cuda', torch.cuda.current_device()) if on_gpu else torch.device('cpu')
    xs = [torch.ones(n, dtype=dtype, device=dev) for n in sizes]
    nccl = None
    if args.method == 'NCCL' and np_ > 1:
        import torch.distributed as dist
        dist.init_process_group('nccl', init_method='tcp://127.0.0.1:%d' % (29200 + np_), rank=rank, world_size=np_)
        nccl = dist

    def step():
        if nccl is not None:
            for x in xs:
                nccl.all_reduce(x)
        else:
            for i, x in enumerate(xs):
                kf.ops.inplace_all_reduce_op(x, 'sum', name='g%d' % i)
        if on_gpu:
            torch.cuda.synchronize()

    for _ in range(args.warmup_steps):
        step()
    kf.run_barrier()
    nbytes = sum(sizes) * xs[0].element_size()
    rates = []
    for _ in range(args.steps):
        for x in xs:
            x.fill_(1)
        t0 = time.time()
        step()
        dt = time.time() - t0
        dt = float(kf.ops.all_reduce_fn(torch.tensor([dt], dtype=torch.float64), 'max')[0])
        rates.append(nbytes / dt)
    if rank == 0:
        mean = sum(rates) / len(rates)
        print('RESULT: model=%s method=%s fuse=%s np=%d tensors=%d bytes=%d algBW=%.3f GB/s busBW=%.3f GB/s '
              'kungfu-rate=%.3f GB/s' % (args.model, args.method, args.fuse, np_, len(sizes), nbytes, mean / 1e9,
                                         mean * 2 * (np_ - 1) / np_ / 1e9, mean * 4 * (np_ - 1) / 1e9))
    kf.run_barrier()


if __name__ == '__main__':
    main()
