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


# S-SGD on a job that spans hosts, every flavour: fused and plain optimizer, with and without overlap=True (which must
# not silently average per host: the overlapped paths keep the post-backward hierarchical exchange on such jobs)
for mk, overlap in ((lambda m: FusedSGD(m.parameters(), lr=0.1, momentum=0.9), False),
                    (lambda m: FusedSGD(m.parameters(), lr=0.1, momentum=0.9), True),
                    (lambda m: torch.optim.SGD(m.parameters(), lr=0.1, momentum=0.9), False),
                    (lambda m: torch.optim.SGD(m.parameters(), lr=0.1, momentum=0.9), True)):
    m, ref = make(), make()
    opt = kf.optimizers.SynchronousSGDOptimizer(mk(m), m.named_parameters(), overlap=overlap)
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

# SMA across hosts (fused and plain optimizer): the average is over ALL workers of the job, not over this host's
def sma_oracle(alpha, lr, steps):
    ms = [make() for _ in range(np_)]
    for r, mm in enumerate(ms):
        for p in mm.parameters():
            p.data += 0.1 * r
    opts = [torch.optim.SGD(mm.parameters(), lr=lr) for mm in ms]
    for step in range(steps):
        for r, (mm, oo) in enumerate(zip(ms, opts)):
            xb, yb = batch(r, step)
            oo.zero_grad()
            torch.nn.functional.cross_entropy(mm(xb), yb).backward()
        with torch.no_grad():
            for ps in zip(*[list(mm.parameters()) for mm in ms]):
                avg = torch.stack([p.data for p in ps]).mean(0)
                for p in ps:
                    p.data.add_(avg - p.data, alpha=alpha)
        for oo in opts:
            oo.step()
    return [p.detach().clone() for p in ms[rank].parameters()]


want = sma_oracle(0.5, 0.05, 3)
for mk in (lambda m: FusedSGD(m.parameters(), lr=0.05), lambda m: torch.optim.SGD(m.parameters(), lr=0.05)):
    m = make()
    for p in m.parameters():
        p.data += 0.1 * rank
    opt = kf.optimizers.SynchronousAveragingOptimizer(mk(m), m.named_parameters(), alpha=0.5)
    for step in range(3):
        xb, yb = batch(rank, step)
        opt.zero_grad(set_to_none=True)
        torch.nn.functional.cross_entropy(m(xb), yb).backward()
        opt.step()
    torch.cuda.synchronize()
    for p, q in zip(m.parameters(), want):
        assert torch.allclose(p, q, rtol=1e-3, atol=1e-4), ('sma across hosts', (p - q).abs().max())

# PairAveraging across hosts: a plain optimizer pairs over the control plane's store with ANY worker of the job;
# the fused (peer-memory) flavour says clearly that it cannot reach the other host
m = make()
for p in m.parameters():
    p.data += 0.1 * rank
opt = kf.optimizers.PairAveragingOptimizer(torch.optim.SGD(m.parameters(), lr=0.0), m.named_parameters())
first = next(m.parameters())
w0 = kf.ops.all_gather(first.data.clone())
for step in range(10):
    opt.step()
kf.run_barrier()
w1 = kf.ops.all_gather(first.data.clone())
assert opt._kf_gen is None, 'a job that spans hosts must not take the peer-memory path'
s0 = float((w0.max(0).values - w0.min(0).values).max())
s1 = float((w1.max(0).values - w1.min(0).values).max())
assert s1 < 0.7 * s0, (s0, s1)
m = make()
opt = kf.optimizers.PairAveragingOptimizer(FusedSGD(m.parameters(), lr=0.0), m.named_parameters())
try:
    opt.step()
    raise AssertionError('fused PairAveraging accepted a job that spans hosts')
except RuntimeError as e:
    assert 'spans' in str(e), e
kf.run_barrier()
print('OK hier rank %d/%d local np %d hosts %d' % (rank, np_, comm.np, kfpy.host_count()))
