"""S-SGD on f(x) = x^2 from x0 = 1: after k steps x = (1 - 2 lr)^k for ANY number of workers, because the
averaged gradient of identical replicas is the single-worker gradient (the reference asserts the same in
/root/reference/srcs/python/kungfu/tensorflow/v1/examples/__main__.py:26-46, tolerance 1e-6)."""
import torch

import kungfu_b200.torch as kf

lr, steps = 0.1, 12
x = torch.nn.Parameter(torch.ones(4))
opt = kf.optimizers.SynchronousSGDOptimizer(torch.optim.SGD([x], lr=lr), named_parameters=[('x', x)])
kf.broadcast_parameters({'x': x.data})
for k in range(1, steps + 1):
    opt.zero_grad()
    (x * x).sum().backward()
    opt.step()
    want = (1 - 2 * lr) ** k
    assert float((x.data - want).abs().max()) < 1e-6, (k, x.data.tolist(), want)
# SMA with alpha = 1 on identical replicas is a no-op on the model, so it follows the same curve
y = torch.nn.Parameter(torch.ones(4))
sma = kf.optimizers.SynchronousAveragingOptimizer(torch.optim.SGD([y], lr=lr), named_parameters=[('y', y)], alpha=1.0)
for k in range(1, 6):
    sma.zero_grad()
    (y * y).sum().backward()
    sma.step()
    assert float((y.data - (1 - 2 * lr) ** k).abs().max()) < 1e-6
print('OK quadratic rank %d of %d' % (kf.current_rank(), kf.current_cluster_size()))
