"""np-invariance of S-SGD: a single-layer perceptron on MNIST-shaped synthetic
data, zero-initialised, fixed sharded order, must end at the SAME weights for
every split of the global batch over np workers. Mirrors
/root/reference/tests/python/integration/test_mnist_slp.py:153-176 (there is no
network for the real MNIST files, so data is synthetic of the same shape)."""
import sys

import torch

import kungfu_b200.torch as kf

rank, np_ = kf.current_rank(), kf.current_cluster_size()
global_batch, steps = int(sys.argv[1]), int(sys.argv[2])
out_path = sys.argv[3]
assert global_batch % np_ == 0
local = global_batch // np_


def train(dtype):
    g = torch.Generator().manual_seed(2020)
    X = torch.rand(steps * global_batch, 784, generator=g).to(dtype)
    W_true = torch.randn(784, 10, generator=g).to(dtype)
    Y = (X @ W_true).argmax(1)

    model = torch.nn.Linear(784, 10).to(dtype)
    torch.nn.init.zeros_(model.weight)
    torch.nn.init.zeros_(model.bias)
    opt = torch.optim.SGD(model.parameters(), lr=0.1)
    opt = kf.optimizers.SynchronousSGDOptimizer(opt, model.named_parameters())
    kf.broadcast_parameters(model.state_dict())
    for s in range(steps):
        lo = s * global_batch + rank * local
        x, y = X[lo:lo + local], Y[lo:lo + local]
        opt.zero_grad()
        # sum over the shard / global batch == mean over the global batch after summing workers
        loss = torch.nn.functional.cross_entropy(model(x), y, reduction='sum') / global_batch * np_
        loss.backward()
        opt.step()
    acc = float((model(X[:2000]).argmax(1) == Y[:2000]).float().mean())
    return {'w': model.weight.detach(), 'b': model.bias.detach(), 'acc': acc}


# one or several precisions in one launch: "f64", "f32" or "f64,f32" (result file per precision: <out>.<dtype>)
names = (sys.argv[4] if len(sys.argv) > 4 else 'f32').split(',')
for name in names:
    res = train(torch.float64 if name == 'f64' else torch.float32)
    if rank == 0:
        torch.save(res, out_path if len(names) == 1 else '%s.%s' % (out_path, name))
    print('OK rank %d %s acc %.4f' % (rank, name, res['acc']))
