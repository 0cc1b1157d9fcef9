"""CPU checks of the model zoo (architecture parity, equivalent formulations)."""
import pytest
import torch


def test_torchvision_parity():
    torchvision = pytest.importorskip('torchvision')
    from kungfu_b200 import models
    for name in ('resnet50', 'densenet121', 'vgg16', 'inception_v3'):
        kw = dict(aux_logits=True, init_weights=False) if name == 'inception_v3' else {}
        ours = getattr(models, name)(**({'aux_logits': True} if name == 'inception_v3' else {}))
        theirs = getattr(torchvision.models, name)(**kw)
        assert set(ours.state_dict().keys()) == set(theirs.state_dict().keys()), name
        theirs.load_state_dict(ours.state_dict())  # checkpoints interchange
        assert sum(p.numel() for p in ours.parameters()) == sum(p.numel() for p in theirs.parameters())


def test_space_to_depth_stem_is_the_same_convolution():
    from kungfu_b200 import models
    torch.manual_seed(0)
    a = models.resnet18(num_classes=10)
    b = models.resnet18(num_classes=10, s2d_stem=True)
    b.load_state_dict(a.state_dict())
    a.eval(), b.eval()
    x = torch.randn(2, 3, 64, 64).contiguous(memory_format=torch.channels_last)
    with torch.no_grad():
        assert torch.allclose(a(x), b(x), rtol=1e-4, atol=1e-4)
    # gradient w.r.t. the shared [64,3,7,7] parameter agrees too
    a.train(), b.train()
    ga = torch.autograd.grad(a._stem(x).pow(2).sum(), a.conv1.weight)[0]
    gb = torch.autograd.grad(b._stem(x).pow(2).sum(), b.conv1.weight)[0]
    assert torch.allclose(ga, gb, rtol=1e-3, atol=1e-3)


def test_bert_and_simple_models_forward():
    from kungfu_b200 import models
    m = models.bert_base(layers=1, hidden=64, heads=4, intermediate=128, vocab_size=100)
    ids = torch.randint(0, 100, (2, 8))
    logits, nsp = m(ids)
    assert logits.shape == (2, 8, 100) and nsp.shape == (2, 2)
    assert models.SLP()(torch.randn(3, 1, 28, 28)).shape == (3, 10)
    assert models.MnistCNN()(torch.randn(3, 1, 28, 28)).shape == (3, 10)


def test_bert_masked_positions_give_the_same_loss():
    """mlm_positions (masked_lm_positions format) == full-sequence head with -100 labels."""
    import torch
    from kungfu_b200 import models
    torch.manual_seed(0)
    m = models.bert_tiny(dropout=0.0)
    m.eval()
    ids = torch.randint(0, 1000, (3, 16))
    pos = torch.stack([torch.randperm(16)[:4].sort().values for _ in range(3)])
    lab_p = torch.gather(ids, 1, pos)
    lab_full = torch.full_like(ids, -100)
    lab_full.scatter_(1, pos, lab_p)
    nsp = torch.tensor([0, 1, 0])
    a = m(ids, mlm_labels=lab_full, nsp_labels=nsp)
    b = m(ids, mlm_labels=lab_p, nsp_labels=nsp, mlm_positions=pos)
    assert abs(float(a) - float(b)) < 1e-5


def test_inception_v3_shapes_and_size():
    """The third model of the reference's published plots: 23.8 M parameters (27.2 M with the auxiliary head)."""
    from kungfu_b200 import models
    m = models.inception_v3()
    assert sum(p.numel() for p in m.parameters()) == 23834568
    assert sum(p.numel() for p in models.inception_v3(aux_logits=True).parameters()) == 27161264
    m.eval()
    with torch.no_grad():
        assert m(torch.randn(1, 3, 299, 299)).shape == (1, 1000)
