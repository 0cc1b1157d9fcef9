#!/usr/bin/env python
"""tcgen05 linear (+bias+GELU) against the library path (cuBLAS GEMM + bias, then
a separate GELU kernel) on the BERT-base / BERT-large FFN and projection shapes.
CUDA-event timed, L2 flushed between iterations, median of --iters."""
import argparse
import json
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kungfu_b200.ops.linear import linear_act, set_tiling  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument('--iters', type=int, default=30)
ap.add_argument('--tiling', default='auto', choices=['auto', 'single', 'pair'])
args = ap.parse_args()
set_tiling(args.tiling)
print('# tiling: %s' % args.tiling)

flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')


def timed(fn):
    for _ in range(5):
        fn()
    ts = []
    for _ in range(args.iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


peaks = {}
try:
    peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'MEASURED_PEAKS.json')))
except Exception:
    pass
peak = peaks.get('bf16_tflops', 1675.8)
print('# M N K act | ours ms (TFLOP/s, %% of measured cuBLAS peak %.0f) | library ms (TFLOP/s) | speedup' % peak)
shapes = [(128, 256, 64, None), (128 * 148, 256, 64, None), (128 * 148, 256, 768, None), (4096, 3072, 768, 'gelu'), (4096, 768, 3072, None), (4096, 2304, 768, None), (4096, 768, 768, None),
          (16384, 3072, 768, 'gelu'), (16384, 768, 3072, None), (8192, 4096, 1024, 'gelu'), (8192, 8192, 8192, None)]
for M, N, K, act in shapes:
    x = torch.randn(M, K, device='cuda').to(torch.bfloat16)
    w = (torch.randn(N, K, device='cuda') / K ** 0.5).to(torch.bfloat16)
    b = torch.randn(N, device='cuda').to(torch.bfloat16)
    with torch.no_grad():
        ours = timed(lambda: linear_act(x, w, b, act))
        lib = timed(lambda: F.gelu(F.linear(x, w, b)) if act == 'gelu' else F.linear(x, w, b))
    fl = 2.0 * M * N * K
    print('%6d %5d %5d %-5s | %.4f ms (%7.1f, %4.1f%%) | %.4f ms (%7.1f) | %.2fx' %
          (M, N, K, act, ours, fl / ours / 1e9, 100 * fl / ours / 1e9 / peak, lib, fl / lib / 1e9, lib / ours))
