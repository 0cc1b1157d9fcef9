"""Fused BN(+residual)(+ReLU) kernels vs a plain PyTorch fp32 reference."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from kungfu_b200.ops.bn_act import bn_act


def _ref(x, res, w, b, relu, eps=1e-5):
    y = F.batch_norm(x, None, None, w, b, True, 0.1, eps)
    if res is not None:
        y = y + res
    return torch.relu(y) if relu else y


@pytest.mark.parametrize('shape', [(8, 64, 14, 14), (4, 256, 7, 9), (2, 2048, 3, 3), (16, 96, 5, 5), (3, 8, 33, 17)])
@pytest.mark.parametrize('relu', [True, False])
@pytest.mark.parametrize('use_res', [True, False])
@pytest.mark.parametrize('dtype', [torch.bfloat16, torch.float16])
def test_bn_act_forward_backward(shape, relu, use_res, dtype):
    torch.manual_seed(0)
    N, C, H, W = shape
    x = (torch.randn(shape, device='cuda') * 2 + 0.5).to(dtype).contiguous(memory_format=torch.channels_last)
    res = torch.randn(shape, device='cuda').to(dtype).contiguous(memory_format=torch.channels_last) if use_res else None
    w = (torch.rand(C, device='cuda') + 0.5).requires_grad_()
    b = torch.randn(C, device='cuda').requires_grad_()
    rm, rv = torch.zeros(C, device='cuda'), torch.ones(C, device='cuda')
    xa = x.clone().requires_grad_()
    ra = res.clone().requires_grad_() if use_res else None
    y = bn_act(xa, ra, w, b, rm, rv, True, 0.1, 1e-5, relu)
    dy = torch.randn_like(y)
    y.backward(dy)
    # fp32 reference on the same (rounded) inputs
    xr = x.float().requires_grad_()
    rr = res.float().requires_grad_() if use_res else None
    wr, br = w.detach().clone().requires_grad_(), b.detach().clone().requires_grad_()
    yr = _ref(xr, rr, wr, br, relu)
    yr.backward(dy.float())
    tol = 2e-2 if dtype == torch.bfloat16 else 4e-3
    assert y.is_contiguous(memory_format=torch.channels_last)
    assert torch.allclose(y.float(), yr, rtol=tol, atol=tol), (y.float() - yr).abs().max()
    # gradient wrt x: compare where the ReLU mask agrees (rounding can flip y ~ 0)
    mask = torch.ones_like(yr, dtype=torch.bool) if not relu else ((yr.abs() > 3e-2) | (yr == 0))
    gx, gxr = xa.grad.float(), xr.grad
    err = ((gx - gxr).abs() * mask).max()
    assert err < 6 * tol * max(1.0, float(gxr.abs().max())), err
    assert torch.allclose(w.grad, wr.grad, rtol=5e-2, atol=5e-2 * float(wr.grad.abs().max()))
    assert torch.allclose(b.grad, br.grad, rtol=5e-2, atol=5e-2 * float(br.grad.abs().max()))
    if use_res:
        assert (((ra.grad.float() - rr.grad).abs() * mask).max()) < 6 * tol
    # running statistics follow nn.BatchNorm2d (unbiased variance in the running estimate)
    m = x.float().mean((0, 2, 3))
    v = x.float().var((0, 2, 3), unbiased=True)
    assert torch.allclose(rm, 0.1 * m, rtol=1e-3, atol=1e-3)
    assert torch.allclose(rv, 0.9 + 0.1 * v, rtol=1e-3, atol=1e-3)


def test_bn_act_eval_mode_and_module():
    from kungfu_b200.models.resnet import BNAct
    torch.manual_seed(1)
    m = BNAct(64, relu=True, fused=True).cuda()
    ref = torch.nn.BatchNorm2d(64).cuda()
    x = torch.randn(8, 64, 10, 10, device='cuda').bfloat16().contiguous(memory_format=torch.channels_last)
    for _ in range(3):
        m(x)
        ref(x.float())
    assert torch.allclose(m.running_mean, ref.running_mean, rtol=1e-2, atol=1e-3)
    assert torch.allclose(m.running_var, ref.running_var, rtol=1e-2, atol=1e-3)
    assert int(m.num_batches_tracked) == 3
    m.eval()
    ref.eval()
    with torch.no_grad():
        y = m(x)
        yr = torch.relu(ref(x.float()))
    assert torch.allclose(y.float(), yr, rtol=2e-2, atol=2e-2)


def test_resnet_fused_matches_unfused():
    """Whole-model check. bf16 rounding differences are amplified by depth, so
    compare in relative L2 rather than element-wise."""
    from kungfu_b200 import models
    torch.manual_seed(0)
    a = models.to_b200_layout(models.resnet18(fused_bn=False).cuda())
    b = models.to_b200_layout(models.resnet18(fused_bn=True).cuda())
    b.load_state_dict(a.state_dict())
    x = torch.randn(32, 3, 64, 64, device='cuda').bfloat16().contiguous(memory_format=torch.channels_last)
    ya, yb = a(x).float(), b(x).float()
    rel = float((ya - yb).norm() / ya.norm())
    assert rel < 0.05, rel
    ya.pow(2).mean().backward()
    yb.pow(2).mean().backward()
    for n in ('fc.weight', ):  # early-layer gradients amplify bf16 rounding through 17 BN layers
        ga = dict(a.named_parameters())[n].grad.float()
        gb = dict(b.named_parameters())[n].grad.float()
        assert float((ga - gb).norm() / ga.norm()) < 0.15, n


@pytest.mark.parametrize('shape', [(4, 64, 56, 56), (2, 8, 7, 9), (3, 16, 112, 112), (2, 32, 5, 4)])
def test_maxpool3x3s2(shape):
    from kungfu_b200.ops.pool import max_pool_3x3s2
    torch.manual_seed(0)
    x = torch.randn(shape, device='cuda').bfloat16().contiguous(memory_format=torch.channels_last)
    xa = x.clone().requires_grad_()
    xr = x.float().requires_grad_()
    y = max_pool_3x3s2(xa)
    yr = F.max_pool2d(xr, 3, 2, 1)
    assert y.shape == yr.shape and torch.equal(y.float(), yr)
    dy = torch.randn_like(yr)
    y.backward(dy.bfloat16())
    yr.backward(dy.bfloat16().float())
    assert torch.allclose(xa.grad.float(), xr.grad, rtol=1e-2, atol=1e-2)


def test_global_avg_pool():
    from kungfu_b200.ops.pool import global_avg_pool
    torch.manual_seed(0)
    x = torch.randn(6, 2048, 7, 7, device='cuda').bfloat16().contiguous(memory_format=torch.channels_last)
    xa = x.clone().requires_grad_()
    xr = x.float().requires_grad_()
    y = global_avg_pool(xa)
    yr = xr.mean((2, 3))
    assert torch.allclose(y.float(), yr, rtol=1e-2, atol=1e-2)
    dy = torch.randn_like(yr)
    y.backward(dy.bfloat16())
    yr.backward(dy.bfloat16().float())
    assert xa.grad.is_contiguous(memory_format=torch.channels_last)
    assert torch.allclose(xa.grad.float(), xr.grad, rtol=1e-2, atol=1e-3)
