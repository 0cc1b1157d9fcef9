"""Device-timed table of every fused communication / optimizer kernel at one
cluster size (SURVEY.md section 8 deliverable: time, algorithmic bytes per
direction, fraction of the NVLink rate).

    kungfu_b200/bin/kungfu-run -np 8 -allow-nvlink python benchmarks/fused_step_bench.py [--elems 25583616]

Timing: CUDA events on the launching stream, 5 warm-up + `--iters` timed
launches back to back, MAX over ranks. Every launch starts with a peer barrier,
so the per-launch time includes the skew between ranks (what a training step
pays). The working set of one launch (ResNet-50's 25.6 M parameters by
default: 48.8 MiB of bf16 gradients, 48.8 MiB of bf16 parameters, 97.6 MiB of
fp32 masters, plus moments) exceeds the 126 MB L2, and peer / multicast
addresses bypass the local L2 anyway.

Byte model per GPU and direction for S bytes: one-shot
(n-1) S, two-shot 2 (n-1)/n S, NVLS S (1 + 1/n); broadcast: root sends S once
with multimem.st, (n-1) S with the push fan-out; pair pull: S ingress.
Link rate: 900 GB/s nominal per direction, 770 GB/s measured peer copy
(/opt/skills/guides/B200_PROFILING.md).
"""
import argparse
import ctypes as C
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
ap.add_argument('--elems', type=int, default=25583616)  # ResNet-50, padded to 128
ap.add_argument('--iters', type=int, default=20)
ap.add_argument('--out', default='')
args = ap.parse_args()

rank, np_ = kf.current_rank(), kf.current_cluster_size()
torch.cuda.set_device(kf.get_cuda_index())
dev = torch.device('cuda', torch.cuda.current_device())
os.environ.setdefault('KUNGFU_PROBE_VARIANTS', '0')
comm = pc.default_comm()
n = args.elems // 128 * 128
S = n * 2  # bf16 bytes
rows = []


def timed(fn):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    kf.run_barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.iters
    return float(kf.ops.all_reduce_fn(torch.tensor([ms], dtype=torch.float64), 'max')[0]) * 1e3  # us


def row(kernel, us, wire_bytes, note=''):
    gbs = wire_bytes / us / 1e3 if us > 0 else 0.0
    rows.append({'kernel': kernel, 'np': np_, 'us': round(us, 1), 'wire_MB_per_dir': round(wire_bytes / 1e6, 2),
                 'GBps': round(gbs, 1), 'of_900': round(gbs / 900, 3), 'of_770_measured': round(gbs / 770, 3), 'note': note})


def hyper(momentum=0.0, kind=0, step_t=None, v=None):
    h = K.SgdHyper()
    h.lr, h.momentum, h.kind = 1e-4, momentum, kind
    h.beta1, h.beta2, h.eps = 0.9, 0.999, 1e-8
    if step_t is not None:
        h.step = step_t.data_ptr()
    if v is not None:
        h.exp_avg_sq = v.data_ptr()
    return h


grads = comm.alloc(S, multicast=True)
params = comm.alloc(S, multicast=True)
grads.view(torch.bfloat16)[:n].normal_()
params.view(torch.bfloat16)[:n].normal_()
master = params.view(torch.bfloat16)[:n].float()
mom = torch.zeros_like(master)
mom2 = torch.zeros_like(master)
step_t = torch.zeros(1, dtype=torch.int64, device=dev)
blocks, threads = comm.default_blocks, comm.default_threads
me = max(rank, 0)
f = (np_ - 1) / np_

if np_ > 1:
    # ---- plain all-reduce, three variants ----
    out = comm.alloc(S, multicast=True)
    for algo, name, wire in ((K.ALGO_TWOSHOT, 'k_allreduce_twoshot', 2 * f * S), (K.ALGO_NVLS, 'k_allreduce_nvls', S * (1 + 1 / np_))):
        if algo == K.ALGO_NVLS and not grads.mc_ptr:
            continue
        us = timed(lambda: coll.all_reduce(comm, grads, n, K.BF16, 'sum', scale=1.0 / np_, algo=algo, out=out))
        row(name, us, wire, 'bf16 48.8 MiB; nccl-tests busBW = %.1f GB/s' % (2 * f * S / us / 1e3))
    small = 64 << 10
    us = timed(lambda: coll.all_reduce(comm, grads, small // 2, K.BF16, 'sum', algo=K.ALGO_ONESHOT, out=out))
    row('k_allreduce_oneshot', us, (np_ - 1) * small, 'bf16 64 KiB (latency-bound size)')
    # ---- fused S-SGD step ----
    for label, h in (('sgd', hyper()), ('sgd+momentum', hyper(0.9)), ('adamw', hyper(kind=1, step_t=step_t, v=mom2))):
        if grads.mc_ptr:
            us = timed(lambda: K.call('kfb_ssgd_fused_step_nvls', comm.cstruct, grads.mc_ptr, K.BF16, params.mc_ptr, K.BF16,
                                      master.data_ptr(), mom.data_ptr(), n, 1.0 / np_, C.byref(h), None,
                                      min(blocks, 32), threads, K.stream_ptr()))
            row('k_ssgd_fused_nvls[%s]' % label, us, S * (1 + 1 / np_) / 2 + S / 2 * (1 + 1 / np_),
                'reduce-scatter + update of the owned shard + all-gather of the new parameters')
        ga, _1 = grads.ptr_array()
        pa, _2 = params.ptr_array()
        us = timed(lambda: K.call('kfb_ssgd_fused_step', comm.cstruct, ga, K.BF16, pa, K.BF16, master.data_ptr(),
                                  mom.data_ptr(), n, 1.0 / np_, C.byref(h), None, blocks, threads, K.stream_ptr()))
        row('k_ssgd_fused[%s]' % label, us, 2 * f * S, 'peer loads / stores')
    # ---- SMA: average only (phase 1), both variants; and the whole step ----
    stage = comm.alloc(n * 4, multicast=True)
    pa, _2 = params.ptr_array()
    sa, _3 = stage.ptr_array()
    h = hyper()
    for nv in ((False, True) if params.mc_ptr else (False, )):
        us = timed(lambda: K.call('kfb_sma_fused_step_ex', comm.cstruct, pa, K.BF16, sa, params.mc_ptr if nv else None,
                                  stage.mc_ptr if nv else None, grads.ptrs[me], K.BF16, master.data_ptr(), mom.data_ptr(), n,
                                  0.1, 1.0, C.byref(h), 1, blocks, threads, K.stream_ptr()))
        # reads bf16 (S), delivers the fp32 average (2 S)
        wire = (S / np_ + 2 * S) if nv else (f * S + f * 2 * S)
        row('k_sma_fused[average%s]' % (', nvls' if nv else ''), us, wire, 'bf16 parameters in, fp32 average out')
    us = timed(lambda: K.call('kfb_sma_fused_step_ex', comm.cstruct, pa, K.BF16, sa, None, None, grads.ptrs[me], K.BF16,
                              master.data_ptr(), mom.data_ptr(), n, 0.1, 1.0, C.byref(h), 0, blocks, threads, K.stream_ptr()))
    row('k_sma_fused[whole step]', us, f * S + f * 2 * S, 'average + blend + SGD in one launch')
    # ---- broadcast ----
    for mc in ((False, True) if params.mc_ptr else (False, )):
        us = timed(lambda: coll.broadcast(comm, params, S, root=0, use_multicast=mc))
        row('k_broadcast[%s]' % ('multimem.st' if mc else 'push fan-out'), us, S if mc else (np_ - 1) * S, 'root egress')
    # ---- PairAveraging: one-sided pull + blend + SGD + publish ----
    snap = comm.alloc(2 * S)
    K.call('kfb_pair_publish', comm.cstruct, params.ptrs[me], snap.ptrs[me], S, S, blocks, threads, K.stream_ptr())
    torch.cuda.synchronize()
    kf.run_barrier()
    torn = torch.zeros(1, dtype=torch.int32, device=dev)
    target = (rank + 1) % np_
    us = timed(lambda: K.call('kfb_pair_avg_step', comm.cstruct, target, snap.ptrs[target], snap.ptrs[me], S, params.ptrs[me],
                              K.BF16, grads.ptrs[me], K.BF16, master.data_ptr(), mom.data_ptr(), n, 1.0, C.byref(h),
                              torn.data_ptr(), K.MAX_BLOCKS, threads, K.stream_ptr()))
    row('k_pair_avg', us, S, 'ingress: the peer\'s snapshot; no barrier; torn reads counted: %d' % int(torn))
else:
    ga, _1 = K.ptr_array([grads.ptrs[0]])
    pa, _2 = K.ptr_array([params.ptrs[0]])
    for label, h in (('sgd', hyper()), ('sgd+momentum', hyper(0.9)), ('adamw', hyper(kind=1, step_t=step_t, v=mom2))):
        us = timed(lambda: K.call('kfb_ssgd_fused_step', comm.cstruct, ga, K.BF16, pa, K.BF16, master.data_ptr(),
                                  mom.data_ptr(), n, 1.0, C.byref(h), None, 128, threads, K.stream_ptr()))
        hbm = n * (2 + 4 + 4 + 2 + (8 if label != 'sgd' else 0) + (8 if label == 'adamw' else 0))
        rows.append({'kernel': 'k_ssgd_fused[%s, local]' % label, 'np': 1, 'us': round(us, 1), 'hbm_MB': round(hbm / 1e6, 1),
                     'GBps': round(hbm / us / 1e3, 1), 'note': 'HBM-bound local update'})

if rank == 0:
    for r in rows:
        print('ROW: ' + json.dumps(r), flush=True)
    if args.out:
        with open(args.out, 'a') as fh:
            for r in rows:
                fh.write(json.dumps(r) + '\n')
kf.run_barrier()
