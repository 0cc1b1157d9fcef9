"""kungfu-run -np 4 python examples/torch_simple_example.py
Reference: /root/reference/examples/torch_simple_example.py:1-10."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))  # run from a source checkout

import torch

import kungfu_b200.torch as kf

x = torch.ones([2, 3])
y = kf.ops.all_reduce_fn(x)
print('rank %d/%d' % (kf.current_rank(), kf.current_cluster_size()), y)
