#!/usr/bin/env python
"""Launches the tcgen05 linear kernel on two shapes (for ncu: -k regex:k_linear_act --launch-skip 2 -c 2)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kungfu_b200.ops.linear import linear_act  # noqa: E402

shapes = [(4096, 3072, 768, 'gelu'), (8192, 8192, 8192, None)]
ts = []
for M, N, K, act in shapes:
    x = torch.randn(M, K, device='cuda').to(torch.bfloat16)
    w = (torch.randn(N, K, device='cuda') / K ** 0.5).to(torch.bfloat16)
    b = torch.randn(N, device='cuda').to(torch.bfloat16)
    ts.append((x, w, b, act))
with torch.no_grad():
    for rep in range(2):
        for x, w, b, act in ts:
            linear_act(x, w, b, act)
        torch.cuda.synchronize()
print('ok')
