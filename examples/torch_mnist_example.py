"""MNIST with SynchronousSGDOptimizer (or --kf-optimizer sma|async-sgd|ada-sgd).

    kungfu-run -np 4 python examples/torch_mnist_example.py --epochs 1 --no-cuda
    kungfu-run -np 4 -allow-nvlink python examples/torch_mnist_example.py

User-side conventions of the reference's examples
(/root/reference/examples/torch_mnist_example.py:162-174,
/root/reference/examples/tf1_mnist_session.py:69-84,137): scale the learning rate by
the cluster size, shard the data by (rank, size), broadcast the initial model,
checkpoint on rank 0 only. Data: real MNIST idx files under --data-dir if present,
otherwise a synthetic MNIST-shaped set (no network here)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))  # run from a source checkout

import argparse

import torch
import torch.nn.functional as F

import kungfu_b200.torch as kf
from kungfu_b200.datasets import load_mnist
from kungfu_b200.models import MnistCNN

ap = argparse.ArgumentParser()
ap.add_argument('--batch-size', type=int, default=64)
ap.add_argument('--epochs', type=int, default=1)
ap.add_argument('--lr', type=float, default=0.01)
ap.add_argument('--momentum', type=float, default=0.5)
ap.add_argument('--no-cuda', action='store_true')
ap.add_argument('--data-dir', default='')
ap.add_argument('--kf-optimizer', default='sync-sgd', choices=['sync-sgd', 'async-sgd', 'sma', 'ada-sgd'])
ap.add_argument('--n-train', type=int, default=6000)
ap.add_argument('--save', default='')
args = ap.parse_args()

use_cuda = not args.no_cuda and torch.cuda.is_available()
if use_cuda:
    torch.cuda.set_device(kf.get_cuda_index())
device = torch.device('cuda' if use_cuda else 'cpu')
rank, size = kf.current_rank(), kf.current_cluster_size()
torch.manual_seed(1)
x_train, y_train, x_test, y_test = load_mnist(args.data_dir, n_train=args.n_train, n_test=1000)
x_train, y_train = x_train[rank::size], y_train[rank::size]  # shard by (rank, size)

model = MnistCNN().to(device)
opt = torch.optim.SGD(model.parameters(), lr=args.lr * size, momentum=args.momentum)
named = list(model.named_parameters())
opt = {
    'sync-sgd': lambda: kf.optimizers.SynchronousSGDOptimizer(opt, named),
    'async-sgd': lambda: kf.optimizers.PairAveragingOptimizer(opt, named),
    'sma': lambda: kf.optimizers.SynchronousAveragingOptimizer(opt, named),
    'ada-sgd': lambda: kf.optimizers.AdaptiveSGDOptimizer(opt, named, change_step=50),
}[args.kf_optimizer]()
kf.broadcast_parameters(model.state_dict())

for epoch in range(args.epochs):
    model.train()
    for i in range(0, len(x_train) - args.batch_size + 1, args.batch_size):
        x, y = x_train[i:i + args.batch_size].to(device), y_train[i:i + args.batch_size].to(device)
        opt.zero_grad()
        loss = F.nll_loss(model(x), y)
        loss.backward()
        opt.step()
    model.eval()
    with torch.no_grad():
        acc = (model(x_test.to(device)).argmax(1).cpu() == y_test).float().mean()
    acc = kf.ops.all_reduce_fn(acc.reshape(1)) / size
    if rank == 0:
        print('epoch %d loss %.4f test accuracy %.4f' % (epoch, float(loss.detach()), float(acc)))
if rank == 0 and args.save:
    torch.save(model.state_dict(), args.save)
