"""Hierarchical (multi-host) GPU path: two 'hosts' on loopback aliases, each
with its own GPUs. NVLink kernels inside a host, control plane between hosts."""
import os

import torch

import kungfu_b200.python as kfpy
import kungfu_b200.torch as kf
from kungfu_b200.optim import FusedSGD
from kungfu_b200.parallel import comm as pc

rank, np_ = kf.current_rank(), kf.current_cluster_size()
torch.cuda.set_device(kf.get_cuda_index() + int(os.environ.get('KUNGFU_TEST_GPU_OFFSET', '0')))
dev = torch.device('cuda', torch.cuda.current_device())
comm = pc.default_comm()
assert comm.cross and comm.global_np == np_ and comm.np == kfpy.current_local_size()

for dtype, tol in ((torch.float32, 1e-5), (torch.bfloat16, 2e-2)):
    xs = [(torch.randn(200000, generator=torch.Generator().manual_seed(50 + r)) * 2).to(dtype) for r in range(np_)]
    got = kf.ops.all_reduce_fn(xs[rank].to(dev)).float().cpu()
    ref = torch.stack([x.float() for x in xs]).sum(0)
    assert torch.allclose(got, ref, rtol=tol, atol=tol * 4), (dtype, (got - ref).abs().max())
x = torch.full((1000, ), float(rank), device=dev)
assert (kf.ops.all_reduce_fn(x, 'max') == np_ - 1).all()
assert kf.ops.all_gather(x)[:, 0].tolist() == [float(r) for r in range(np_)]
sd = {'w': torch.full((77, ), float(rank + 1), device=dev)}
kf.broadcast_parameters(sd)
assert (sd['w'] == 1).all()


def make():
    torch.manual_seed(0)
    return torch.nn.Sequential(torch.nn.Linear(32, 64), torch.nn.ReLU(), torch.nn.Linear(64, 8)).to(dev)


def batch(r, step):
    g = torch.Generator().manual_seed(7 + 13 * r + step)
    return torch.randn(16, 32, generator=g).to(dev), torch.randint(0, 8, (16, ), generator=g).to(dev)


for mk in (lambda m: FusedSGD(m.parameters(), lr=0.1, momentum=0.9), lambda m: torch.optim.SGD(m.parameters(), lr=0.1, momentum=0.9)):
    m, ref = make(), make()
    opt = kf.optimizers.SynchronousSGDOptimizer(mk(m), m.named_parameters())
    ropt = torch.optim.SGD(ref.parameters(), lr=0.1, momentum=0.9)
    for step in range(3):
        xb, yb = batch(rank, step)
        opt.zero_grad(set_to_none=True)
        torch.nn.functional.cross_entropy(m(xb), yb).backward()
        opt.step()
        ropt.zero_grad()
        xs, ys = zip(*[batch(r, step) for r in range(np_)])
        torch.nn.functional.cross_entropy(ref(torch.cat(xs)), torch.cat(ys)).backward()
        ropt.step()
    torch.cuda.synchronize()
    for p, q in zip(m.parameters(), ref.parameters()):
        assert torch.allclose(p, q, rtol=1e-4, atol=1e-5), (p - q).abs().max()
kf.run_barrier()
print('OK hier rank %d/%d local np %d hosts %d' % (rank, np_, comm.np, kfpy.host_count()))
