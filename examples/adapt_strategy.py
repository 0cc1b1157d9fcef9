"""Adaptive strategy switching demo: monitored all-reduce, per-step throughput
stats, and after a few steps a cluster-wide vote on switching the tree.

    kungfu-run -np 4 -strategy BINARY_TREE python examples/adapt_strategy.py

Reference: /root/reference/experimental/adapt_strategy/adapt_strategy.py:26-28,179-216
(monitor=True S-SGD, calc_stats/log_stats each step, check_interference after
step 4, on consensus set_tree([1]*np) = star rooted at rank 1). On GPUs the same
loop can drive kungfu_b200.parallel.planner.VariantChooser.check_interference."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))  # run from a source checkout

import torch

import kungfu_b200.python as kfpy
import kungfu_b200.torch as kf

np_ = kf.current_cluster_size()
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
