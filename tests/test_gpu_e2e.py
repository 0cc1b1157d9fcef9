"""End-to-end GPU checks through the public API on one GPU."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_smoke_entry():
    import __graft_entry__ as g
    g.smoke()


def test_resnet_fused_sgd_tracks_torch_sgd():
    """A small ResNet trained with SynchronousSGDOptimizer(FusedSGD) follows
    torch.optim.SGD on an identical fp32 model step by step."""
    import torch.nn.functional as F

    import kungfu_b200.torch as kf
    from kungfu_b200 import models
    from kungfu_b200.optim import FusedSGD

    def make():
        torch.manual_seed(0)
        return models.resnet18(num_classes=10).cuda().to(memory_format=torch.channels_last)

    a, b = make(), make()
    oa = torch.optim.SGD(a.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
    ob = kf.optimizers.SynchronousSGDOptimizer(FusedSGD(b.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4),
                                                b.named_parameters())
    x = torch.randn(8, 3, 32, 32, device='cuda').contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 10, (8, ), device='cuda')
    la, lb = [], []
    for step in range(4):
        for m, o, l in ((a, oa, la), (b, ob, lb)):
            o.zero_grad(set_to_none=True)
            loss = F.cross_entropy(m(x), y)
            loss.backward()
            o.step()
            l.append(float(loss.detach()))
        if step == 0:
            # one update from identical weights and (numerically) identical gradients
            for (n, pa), pb in zip(a.named_parameters(), b.parameters()):
                assert torch.allclose(pa, pb, rtol=1e-4, atol=1e-5), (n, (pa - pb).abs().max())
    # training is chaotic (batch 8, BN): only require both to learn and to start together
    assert abs(la[1] - lb[1]) < 2e-2 * abs(la[1]), (la, lb)
    assert lb[-1] < lb[0] and la[-1] < la[0]


def test_generic_optimizer_grad_sync_single_gpu():
    """Non-fused optimizers go through GradSync (bucket gather + views)."""
    import kungfu_b200.torch as kf
    torch.manual_seed(0)
    m = torch.nn.Sequential(torch.nn.Linear(32, 64), torch.nn.Tanh(), torch.nn.Linear(64, 4)).cuda()
    ref = [p.detach().clone() for p in m.parameters()]
    opt = kf.optimizers.SynchronousSGDOptimizer(torch.optim.Adam(m.parameters(), lr=1e-2), m.named_parameters())
    x = torch.randn(16, 32, device='cuda')
    m(x).pow(2).mean().backward()
    g = [p.grad.detach().clone() for p in m.parameters()]
    opt.step()
    torch.cuda.synchronize()
    for p, g0 in zip(m.parameters(), g):
        assert torch.allclose(p.grad, g0)  # np == 1: averaging is the identity
    assert any(not torch.equal(p, r) for p, r in zip(m.parameters(), ref))
