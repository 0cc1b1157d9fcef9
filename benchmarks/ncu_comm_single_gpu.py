"""The communication / optimizer kernels on ONE GPU, for Nsight Compute.

ncu replays every kernel many times, which a kernel that waits for other ranks
cannot survive, so multi-rank runs are never profiled. Here the same kernels
run under single-GPU rank emulation (one cooperative launch whose blockIdx.y
plays the rank: every "peer" pointer is local memory), at ResNet-50 size.
What the capture shows is the kernel's own instruction mix, memory pipeline
behaviour and occupancy; NVLink is of course not in it (and the NVLS variants,
which need a real multicast object, are not emulated at all: for those see the
device-timed tables of benchmarks/fused_step_bench.py and the SASS listings).

    python benchmarks/ncu_comm_single_gpu.py [--np 8] [--elems 25583616]
    ncu --set full --clock-control none --import-source on -k regex:'k_allreduce_twoshot|k_ssgd_fused|k_sma_fused|k_pair_avg|k_broadcast|k_grad_variance' \
        -s 12 -c 6 -o gpurun_out/comm_emulated python benchmarks/ncu_comm_single_gpu.py
"""
import argparse
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from kungfu_b200.ops import _cuda as K
from kungfu_b200.ops import collective as coll
from kungfu_b200.parallel.comm import EmulatedComm, LocalComm

ap = argparse.ArgumentParser()
ap.add_argument('--np', type=int, default=8)
ap.add_argument('--elems', type=int, default=25583616)
ap.add_argument('--reps', type=int, default=3)
args = ap.parse_args()
torch.cuda.set_device(0)
np_, n = args.np, args.elems // 128 * 128
comm = EmulatedComm(np_)
S = n * 2
grads, params, out = comm.alloc(S), comm.alloc(S), comm.alloc(S)
stage = comm.alloc(n * 4)
for r in range(np_):
    grads.view(torch.bfloat16, r)[:n].normal_()
    params.view(torch.bfloat16, r)[:n].normal_()
master = torch.cat([params.view(torch.bfloat16, r)[:n].float() for r in range(np_)])
mom = torch.zeros_like(master)
h = K.SgdHyper()
h.lr, h.momentum = 1e-4, 0.9
blocks, threads = comm.default_blocks, comm.default_threads
ga, _1 = grads.ptr_array()
pa, _2 = params.ptr_array()
sa, _3 = stage.ptr_array()
ends = torch.tensor([n // 3 // 16 * 16, n // 2 // 16 * 16, n], dtype=torch.int64, device='cuda')
seg = torch.zeros(np_ * 3, device='cuda')
work = torch.zeros(K.VAR_WORK_WORDS * np_, device='cuda')
# the pair-averaging kernel has no barrier: it runs as a plain launch against a local "peer" snapshot
lc = LocalComm()
snap = torch.zeros(2 * S, dtype=torch.uint8, device='cuda')
one_p = params.view(torch.bfloat16, 0)[:n]
one_g = grads.view(torch.bfloat16, 0)[:n]
one_master = one_p.float()
one_mom = torch.zeros_like(one_master)
torn = torch.zeros(1, dtype=torch.int32, device='cuda')
K.call('kfb_pair_publish', lc.cstruct, one_p.data_ptr(), snap.data_ptr(), S, S, 32, 512, K.stream_ptr())
for _ in range(args.reps):
    coll.all_reduce(comm, grads, n, K.BF16, 'sum', scale=1.0 / np_, algo=K.ALGO_TWOSHOT, out=out)
    K.call('kfb_ssgd_fused_step', comm.cstruct, ga, K.BF16, pa, K.BF16, master.data_ptr(), mom.data_ptr(), n,
           1.0 / np_, C.byref(h), None, blocks, threads, K.stream_ptr())
    garr = C.cast(ga, C.c_void_p)
    K.call('kfb_sma_fused_step_ex', comm.cstruct, pa, K.BF16, sa, None, None, garr, K.BF16, master.data_ptr(),
           mom.data_ptr(), n, 0.1, 1.0, C.byref(h), 0, blocks, threads, K.stream_ptr())
    # (target -1: blend skipped, update + publish only; the peer pull itself needs a second real rank)
    K.call('kfb_pair_avg_step', lc.cstruct, -1, snap.data_ptr(), snap.data_ptr(), S, one_p.data_ptr(), K.BF16,
           one_g.data_ptr(), K.BF16, one_master.data_ptr(), one_mom.data_ptr(), n, 1.0, C.byref(h), torn.data_ptr(),
           K.MAX_BLOCKS, 512, K.stream_ptr())
    coll.broadcast(comm, params, S, root=0, use_multicast=False)
    K.call('kfb_grad_variance', comm.cstruct, ga, K.BF16, n, ends.data_ptr(), 3, seg.data_ptr(), work.data_ptr(),
           blocks, threads, K.stream_ptr())
torch.cuda.synchronize()
print('OK ncu_comm_single_gpu np=%d elems=%d blocks=%d' % (np_, n, blocks))
