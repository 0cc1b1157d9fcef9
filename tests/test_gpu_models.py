"""Model zoo on one GPU: forward/backward with the fused kernels and every
optimizer wrapper's single-GPU path (np == 1 degenerates to local updates)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _train(model, opt, x, y, steps=3, loss_fn=None):
    losses = []
    for _ in range(steps):
        opt.zero_grad(set_to_none=True)
        loss = loss_fn() if loss_fn else F.cross_entropy(model(x).float(), y)
        loss.backward()
        opt.step()
        losses.append(float(loss.detach()))
    torch.cuda.synchronize()
    assert all(l == l for l in losses), losses
    return losses


@pytest.mark.parametrize('name,kw', [('densenet121', {'fused_bn': True}), ('resnet18', {'fused_bn': True}), ('vgg11', {})])
@pytest.mark.parametrize('wrapper', ['ssgd', 'sma', 'pair', 'gns'])
def test_models_with_fused_optimizers(name, kw, wrapper):
    import kungfu_b200.torch as kf
    from kungfu_b200 import models
    from kungfu_b200.optim import FusedSGD
    torch.manual_seed(0)
    model = models.to_b200_layout(getattr(models, name)(num_classes=10, **kw).cuda())
    if wrapper == 'pair':  # fused pair averaging keeps one snapshot sequence: one dtype
        model = model.to(torch.bfloat16) if name.startswith('vgg') else model
        if not name.startswith('vgg'):
            pytest.skip('mixed bf16 / fp32-BN model: PairAveraging fused path needs one dtype')
    base = FusedSGD(model.parameters(), lr=0.01, momentum=0.9)
    named = list(model.named_parameters())
    opt = {'ssgd': lambda: kf.optimizers.SynchronousSGDOptimizer(base, named),
           'sma': lambda: kf.optimizers.SynchronousAveragingOptimizer(base, named),
           'pair': lambda: kf.optimizers.PairAveragingOptimizer(base, named),
           'gns': lambda: kf.optimizers.MonitorGradientNoiseScaleOptimizer(base, named, device_batch_size=4)}[wrapper]()
    size = 64 if not name.startswith('vgg') else 32
    x = torch.randn(4, 3, size, size, device='cuda').bfloat16().contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 10, (4, ), device='cuda')
    w0 = next(model.parameters()).detach().clone()
    _train(model, opt, x, y)
    assert not torch.equal(w0, next(model.parameters()))


def test_bert_tiny_pair_averaging_and_state_dict_roundtrip(tmp_path):
    import kungfu_b200.torch as kf
    from kungfu_b200 import models
    from kungfu_b200.optim import FusedSGD
    torch.manual_seed(0)
    model = models.bert_base(layers=2, hidden=128, heads=4, intermediate=256, vocab_size=1000).cuda().to(torch.bfloat16)
    opt = kf.optimizers.PairAveragingOptimizer(FusedSGD(model.parameters(), lr=0.01), model.named_parameters())
    ids = torch.randint(0, 1000, (4, 32), device='cuda')
    labels = ids.clone()
    nsp = torch.randint(0, 2, (4, ), device='cuda')
    losses = _train(model, opt, None, None, steps=4, loss_fn=lambda: model(ids, mlm_labels=labels, nsp_labels=nsp))
    assert losses[-1] < losses[0]
    # parameters live in the flat arena but checkpoints stay plain state_dicts
    path = str(tmp_path / 'ckpt.pt')
    torch.save(model.state_dict(), path)
    fresh = models.bert_base(layers=2, hidden=128, heads=4, intermediate=256, vocab_size=1000).to(torch.bfloat16)
    fresh.load_state_dict(torch.load(path))
    for (n, a), (_, b) in zip(model.state_dict().items(), fresh.state_dict().items()):
        assert torch.equal(a.cpu(), b), n


def test_cuda_graph_capture_of_a_full_step():
    import kungfu_b200.torch as kf
    from kungfu_b200 import models
    from kungfu_b200.optim import FusedSGD
    torch.manual_seed(0)
    model = models.to_b200_layout(models.resnet18(num_classes=10, fused_bn=True).cuda())
    opt = kf.optimizers.SynchronousSGDOptimizer(FusedSGD(model.parameters(), lr=0.05, momentum=0.9), model.named_parameters())
    x = torch.randn(8, 3, 64, 64, device='cuda').bfloat16().contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 10, (8, ), device='cuda')
    loss_out = torch.zeros((), device='cuda')

    def step():
        opt.zero_grad(set_to_none=True)
        loss = F.cross_entropy(model(x).float(), y)
        loss.backward()
        opt.step()
        loss_out.copy_(loss.detach())

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            step()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        step()
    first = None
    for i in range(10):
        g.replay()
        torch.cuda.synchronize()
        first = float(loss_out) if first is None else first
    assert float(loss_out) < first  # replays keep training (epoch flags, momentum, BN stats all advance)


def test_inception_v3_trains_on_the_fused_bn_path():
    import kungfu_b200.torch as kf
    from kungfu_b200 import models
    from kungfu_b200.optim import FusedSGD
    torch.manual_seed(0)
    m = models.to_b200_layout(models.inception_v3(num_classes=10, fused_bn=True).cuda())
    x = torch.randn(4, 3, 299, 299, device='cuda').bfloat16().contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 10, (4, ), device='cuda')
    m.train()
    assert m(x).shape == (4, 10)
    opt = kf.optimizers.SynchronousSGDOptimizer(FusedSGD(m.parameters(), lr=0.001, momentum=0.9), list(m.named_parameters()))
    w0 = next(m.parameters()).detach().clone()
    for _ in range(2):
        opt.zero_grad(set_to_none=True)
        loss = torch.nn.functional.cross_entropy(m(x).float(), y)
        loss.backward()
        opt.step()
    torch.cuda.synchronize()
    assert bool(torch.isfinite(loss)), float(loss)
    assert not torch.equal(w0, next(m.parameters()))
