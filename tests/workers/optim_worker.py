"""Every optimizer wrapper on CPU tensors with value checks against a
single-process oracle (the reference only ran them: 'FIXME: check values',
/root/reference/tests/python/integration/test_optimizers.py:9-54)."""
import sys

import torch

import kungfu_b200.torch as kf

rank, np_ = kf.current_rank(), kf.current_cluster_size()
kinds = sys.argv[1].split(',')  # one or several checks in one launch: ssgd,sma,monitors


def model():
    torch.manual_seed(0)
    return torch.nn.Linear(4, 3)


def data(r):
    g = torch.Generator().manual_seed(100 + r)
    return torch.randn(8, 4, generator=g), torch.randn(8, 3, generator=g)


def grads_of(m, r):
    x, y = data(r)
    m.zero_grad()
    torch.nn.functional.mse_loss(m(x), y).backward()
    return [p.grad.clone() for p in m.parameters()]


for kind in kinds:
    m = model()
    base = torch.optim.SGD(m.parameters(), lr=0.1)

    if kind == 'ssgd':
        # x^2 example of the reference: S-SGD on f(x)=x^2 gives x_k = (1-2lr)^k x_0
        x = torch.nn.Parameter(torch.tensor([1.0 + 0 * rank]))
        o = kf.optimizers.SynchronousSGDOptimizer(torch.optim.SGD([x], lr=0.1), [('x', x)])
        for k in range(1, 6):
            o.zero_grad()
            (x * x).sum().backward()
            o.step()
            assert abs(float(x) - (1 - 0.2) ** k) < 1e-6
        # averaged gradients == oracle
        opt = kf.optimizers.SynchronousSGDOptimizer(base, m.named_parameters())
        ref = model()
        for step in range(3):
            gs = [grads_of(ref, r) for r in range(np_)]
            avg = [sum(g[i] for g in gs) / np_ for i in range(2)]
            for p, g in zip(ref.parameters(), avg):
                p.data -= 0.1 * g
            grads_of(m, rank)
            opt.step()
            for p, q in zip(m.parameters(), ref.parameters()):
                assert torch.allclose(p, q, atol=1e-6), (step, p, q)
        # op='sum' reproduces the reference torch stub
        m2 = model()
        o2 = kf.optimizers.SynchronousSGDOptimizer(torch.optim.SGD(m2.parameters(), lr=0.1), m2.named_parameters(), op='sum')
        gs = [grads_of(model(), r) for r in range(np_)]
        grads_of(m2, rank)
        o2.step()
        exp = [p - 0.1 * sum(g[i] for g in gs) for i, p in enumerate(model().parameters())]
        for p, q in zip(m2.parameters(), exp):
            assert torch.allclose(p, q, atol=1e-6)
    elif kind == 'sma':
        alpha = 0.1
        for i, p in enumerate(m.parameters()):
            p.data += rank  # diverged replicas
        opt = kf.optimizers.SynchronousAveragingOptimizer(base, m.named_parameters(), alpha=alpha)
        replicas = []
        for r in range(np_):
            mr = model()
            for p in mr.parameters():
                p.data += r
            replicas.append(mr)
        for step in range(3):
            avg = [sum(list(mr.parameters())[i].data for mr in replicas) / np_ for i in range(2)]
            gs = [grads_of(mr, r) for r, mr in enumerate(replicas)]  # gradients at the PRE-blend weights
            for mr, g in zip(replicas, gs):
                for i, p in enumerate(mr.parameters()):
                    p.data = (1 - alpha) * p.data + alpha * avg[i] - 0.1 * g[i]
            grads_of(m, rank)
            opt.step()
            for p, q in zip(m.parameters(), replicas[rank].parameters()):
                assert torch.allclose(p, q, atol=1e-5), (step, p, q)
    elif kind == 'pair':
        opt = kf.optimizers.PairAveragingOptimizer(base, m.named_parameters(), seed=7)
        for step in range(4):
            grads_of(m, rank)
            opt.step()
        for p in m.parameters():
            assert torch.isfinite(p).all()
        kf.run_barrier()
    elif kind == 'monitors':
        opt = kf.optimizers.MonitorGradientVarianceOptimizer(base, m.named_parameters())
        gs = [grads_of(model(), r) for r in range(np_)]
        exp = 0.0
        for i in range(2):
            st = torch.stack([g[i] for g in gs])
            exp += float(torch.norm((st * st).mean(0) - st.mean(0) ** 2))
        grads_of(m, rank)
        opt.step()
        assert abs(float(opt.variance) - exp) < 1e-5 * max(1, exp), (float(opt.variance), exp)
        m3 = model()
        o3 = kf.optimizers.MonitorGradientNoiseScaleOptimizer(torch.optim.SGD(m3.parameters(), lr=0.1), m3.named_parameters(), device_batch_size=8)
        grads_of(m3, rank)
        o3.step()
        o4 = kf.optimizers.AdaptiveSGDOptimizer(torch.optim.SGD(m3.parameters(), lr=0.1), m3.named_parameters(), change_step=2)
        for _ in range(4):
            grads_of(m3, rank)
            o4.step()
        ws = kf.ops.all_gather(next(m3.parameters()).data)
        assert torch.allclose(ws[0], ws[-1], atol=1e-6)  # S-SGD phase keeps replicas identical
    kf.run_barrier()
    print('OK', kind, rank, np_)
