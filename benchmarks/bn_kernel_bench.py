"""Microbenchmark of the fused BN(+residual)(+ReLU) kernels at the ResNet-50
(batch 256) shapes: achieved bytes/s vs the measured copy bandwidth of
MEASURED_PEAKS.json. Timed with CUDA events, L2 flushed between iterations."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from kungfu_b200.ops import _cuda as K

ap = argparse.ArgumentParser()
ap.add_argument('--batch', type=int, default=256)
ap.add_argument('--iters', type=int, default=10)
ap.add_argument('--only', default='')
args = ap.parse_args()

peaks = {}
try:
    peaks = json.load(open(os.path.join(os.path.dirname(__file__), '..', 'MEASURED_PEAKS.json')))
except Exception:
    pass
HBM = peaks.get('hbm_gbs', 6650.0)
tag = 'of measured' if 'hbm_gbs' in peaks else 'of fallback'

SHAPES = [(64, 112, True, False), (64, 56, True, False), (256, 56, True, True), (256, 56, False, False),
          (128, 28, True, False), (512, 28, True, True), (256, 14, True, False), (1024, 14, True, True),
          (512, 7, True, False), (2048, 7, True, True)]
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')


def timeit(fn):
    ts = []
    for _ in range(args.iters + 3):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts = sorted(ts[3:])
    return ts[len(ts) // 2]


print('%-28s %10s %10s %8s   %10s %10s %8s' % ('shape C,HW,relu,res', 'fwd ms', 'GB/s', 'frac', 'bwd ms', 'GB/s', 'frac'))
for C, S, relu, res in SHAPES:
    if args.only and args.only != '%dx%d' % (C, S):
        continue
    N = args.batch
    rows = N * S * S
    x = torch.randn(N, C, S, S, device='cuda').bfloat16().contiguous(memory_format=torch.channels_last)
    r = torch.randn_like(x) if res else None
    y = torch.empty_like(x)
    dy = torch.randn_like(x)
    dx = torch.empty_like(x)
    dr = torch.empty_like(x) if res else None
    w = torch.ones(C, device='cuda')
    b = torch.zeros(C, device='cuda')
    rm, rv = torch.zeros(C, device='cuda'), torch.ones(C, device='cuda')
    sm, si = torch.empty(C, device='cuda'), torch.empty(C, device='cuda')
    dw, db = torch.empty(C, device='cuda'), torch.empty(C, device='cuda')
    mask = torch.empty(rows * C // 8, dtype=torch.uint8, device='cuda')
    scratch = torch.empty(K.lib().kfb_bn_scratch_floats(rows, C), device='cuda')

    def fwd():
        K.call('kfb_bn_act_fwd', x.data_ptr(), r.data_ptr() if res else None, y.data_ptr(), w.data_ptr(), b.data_ptr(),
               rm.data_ptr(), rv.data_ptr(), sm.data_ptr(), si.data_ptr(), rows, C, 1e-5, 0.1, int(relu), 1, K.BF16,
               scratch.data_ptr(), mask.data_ptr() if relu else None, K.stream_ptr())

    def bwd():
        K.call('kfb_bn_act_bwd', dy.data_ptr(), x.data_ptr(), mask.data_ptr() if relu else None, w.data_ptr(),
               sm.data_ptr(), si.data_ptr(), dx.data_ptr(), dr.data_ptr() if res else None, dw.data_ptr(),
               db.data_ptr(), rows, C, int(relu), K.BF16, scratch.data_ptr(), K.stream_ptr())

    t = x.numel() * 2
    fb = t * (3 + (1 if res else 0)) + (t // 16 if relu else 0)
    bb = t * (5 + (1 if res else 0)) + (2 * t // 16 if relu else 0)
    tf, tb = timeit(fwd), timeit(bwd)
    print('%-28s %10.3f %10.0f %7.2f%%   %10.3f %10.0f %7.2f%%  %s' %
          ('%d,%d,%s,%s' % (C, S, relu, res), tf, fb / tf / 1e6, 100 * fb / tf / 1e6 / HBM,
           tb, bb / tb / 1e6, 100 * bb / tb / 1e6 / HBM, tag))
