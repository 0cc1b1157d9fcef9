"""Adaptive strategy switching demo: monitored all-reduce, per-step throughput
stats, and after a few steps a cluster-wide vote on switching the tree.

    kungfu-run -np 4 -strategy BINARY_TREE python examples/adapt_strategy.py

Reference: /root/reference/experimental/adapt_strategy/adapt_strategy.py:26-28,179-216
(monitor=True S-SGD, calc_stats/log_stats each step, check_interference after
step 4, on consensus set_tree([1]*np) = star rooted at rank 1).

    kungfu-run -np 8 -allow-nvlink python examples/adapt_strategy.py --cuda [--inject-slow]

On GPUs the gradients never touch the host runtime: the statistics come from
the device kernels of the TRAINING step (bytes and CUDA-event time per kernel
variant), and what the vote replaces is the kernel variant (one-shot /
two-shot / NVLS). --inject-slow makes the active variant run on a single CTA
from step 6 on, which the vote then detects."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))  # run from a source checkout

import torch

import kungfu_b200.python as kfpy
import kungfu_b200.torch as kf

np_ = kf.current_cluster_size()
if '--cuda' in sys.argv:
    import torch.nn.functional as F

    from kungfu_b200.ops import collective as coll
    from kungfu_b200.optim import FusedSGD
    from kungfu_b200.parallel.devstats import VARIANTS
    torch.cuda.set_device(kf.get_cuda_index())
    dev = torch.device('cuda', torch.cuda.current_device())
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(2048, 4096), torch.nn.ReLU(), torch.nn.Linear(4096, 2048),
                                torch.nn.ReLU(), torch.nn.Linear(2048, 10)).to(dev).to(torch.bfloat16)
    opt = kf.optimizers.SynchronousSGDOptimizer(FusedSGD(model.parameters(), lr=0.01), model.named_parameters())
    x = torch.randn(64, 2048, device=dev, dtype=torch.bfloat16)
    y = torch.randint(0, 10, (64, ), device=dev)
    switched_at = None
    for step in range(16):
        opt.zero_grad(set_to_none=True)
        F.cross_entropy(model(x).float(), y).backward()
        opt.step()
        kfpy.calc_stats()
        if step == 6 and '--inject-slow' in sys.argv:
            # (job-wide: a collective kernel needs the same grid on every rank)
            coll.SLOW_VARIANT = opt.comm.stats.active
        if step > 4 and step % 2 == 1 and kfpy.check_interference():
            coll.SLOW_VARIANT = None
            switched_at = switched_at or step
            if kf.current_rank() == 0:
                print('step %d: cluster voted the active kernel variant out; now using %s' %
                      (step, VARIANTS.get(coll.PREFERRED_LARGE)), flush=True)
    if kf.current_rank() == 0:
        kfpy.print_strategy_stats()
        print('OK adapt_strategy cuda: switched at step %s, fused step ran %s' % (switched_at, opt.step_algos))
    kf.run_barrier()
    sys.exit(0)
grad = torch.ones(4 << 20)
for step in range(10):
    out = kf.ops.monitored_all_reduce(grad, name='grad')
    assert float(out[0]) == np_
    kfpy.calc_stats()
    kfpy.log_stats()
    if step > 4 and kfpy.check_interference():
        if kf.current_rank() == 0:
            print('step %d: cluster voted to switch strategy' % step)
        kfpy.set_tree([min(1, np_ - 1)] * np_)
if kf.current_rank() == 0:
    kfpy.print_strategy_stats()
