#!/usr/bin/env python
"""Smallest multi-GPU sanity check (seconds): bootstrap under torchrun, map peer memory, one all-reduce on the device.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 benchmarks/mini_multi_gpu_check.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import kungfu_b200.torch as kf

torch.cuda.set_device(int(os.environ.get('LOCAL_RANK', '0')))
r, n = kf.current_rank(), kf.current_cluster_size()
x = torch.full((1 << 20, ), float(r + 1), device='cuda', dtype=torch.bfloat16)
y = kf.ops.all_reduce_fn(x, name='mini')
torch.cuda.synchronize()
assert float(y[0]) == n * (n + 1) / 2 and float(y[-1]) == float(y[0]), (float(y[0]), n)
kf.run_barrier()
print('OK mini multi-gpu check rank %d of %d' % (r, n), flush=True)
