"""tcgen05 linear layer (csrc/kernels/linear_tc.cu) against a plain fp32 PyTorch
reference of the same op."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(params=['single', 'pair'], autouse=True)
def tiling(request):
    """Every test runs on both tilings of the kernel (cta_group::1 and cta_group::2)."""
    from kungfu_b200.ops.linear import set_tiling
    set_tiling(request.param)
    yield request.param
    set_tiling('auto')


def _ref(x, w, b, act):
    z = x.float() @ w.float().t()
    if b is not None:
        z = z + b.float()
    if act == 'gelu':
        return F.gelu(z)
    if act == 'relu':
        return F.relu(z)
    return z


def _mk(M, N, K, bias_dtype, seed=0):
    g = torch.Generator(device='cuda').manual_seed(seed)
    x = torch.randn(M, K, device='cuda', generator=g).to(torch.bfloat16)
    w = (torch.randn(N, K, device='cuda', generator=g) / K ** 0.5).to(torch.bfloat16)
    b = None if bias_dtype is None else torch.randn(N, device='cuda', generator=g).to(bias_dtype)
    return x, w, b


@pytest.mark.parametrize('shape', [(128, 256, 64), (256, 512, 256), (4096, 3072, 768), (4096, 768, 3072),
                                   (200, 136, 72), (1, 8, 8), (1000, 1000, 2048), (37 * 128, 256, 64)])
@pytest.mark.parametrize('act', [None, 'gelu', 'relu'])
def test_forward_matches_fp32_reference(shape, act):
    from kungfu_b200.ops.linear import linear_act, supported
    M, N, K = shape
    x, w, b = _mk(M, N, K, torch.float32)
    assert supported(x, w, b)
    y = linear_act(x, w, b, act)
    torch.cuda.synchronize()
    ref = _ref(x, w, b, act)
    assert y.shape == ref.shape and y.dtype == torch.bfloat16
    err = (y.float() - ref).abs()
    tol = 1e-2 * ref.abs() + 2e-2
    assert bool((err <= tol).all()), 'max err %g at %s' % (float(err.max()), shape)
    # tighter: relative L2 at bf16 output resolution
    assert float((y.float() - ref).norm() / ref.norm().clamp_min(1e-6)) < 4e-3


@pytest.mark.parametrize('bias_dtype', [None, torch.bfloat16, torch.float32])
def test_bias_variants_and_batch_dims(bias_dtype):
    from kungfu_b200.ops.linear import linear_act
    x, w, b = _mk(6 * 50, 264, 128, bias_dtype, seed=3)
    x3 = x.view(6, 50, 128)
    y = linear_act(x3, w, b, 'gelu')
    assert y.shape == (6, 50, 264)
    ref = _ref(x, w, b, 'gelu').view(6, 50, 264)
    assert float((y.float() - ref).norm() / ref.norm()) < 4e-3


@pytest.mark.parametrize('act', [None, 'gelu', 'relu'])
def test_backward_matches_fp32_autograd(act):
    from kungfu_b200.ops.linear import linear_act
    x, w, b = _mk(512, 384, 256, torch.bfloat16, seed=5)
    x.requires_grad_(True), w.requires_grad_(True), b.requires_grad_(True)
    y = linear_act(x, w, b, act)
    dy = torch.randn_like(y)
    y.backward(dy)
    xf, wf, bf = [t.detach().float().requires_grad_(True) for t in (x, w, b)]
    _ref(xf, wf, bf, act).backward(dy.float())
    for got, want, name in ((x.grad, xf.grad, 'dx'), (w.grad, wf.grad, 'dw'), (b.grad, bf.grad, 'db')):
        rel = float((got.float() - want).norm() / want.norm())
        assert rel < 2e-2, '%s rel err %g' % (name, rel)


def test_repeated_calls_are_deterministic_and_graph_capturable():
    from kungfu_b200.ops.linear import linear_act
    x, w, b = _mk(1024, 1024, 512, torch.float32, seed=7)
    y0 = linear_act(x, w, b, 'gelu')
    for _ in range(3):
        assert torch.equal(linear_act(x, w, b, 'gelu'), y0)
    out = torch.empty_like(y0)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        out.copy_(linear_act(x, w, b, 'gelu'))
    torch.cuda.current_stream().wait_stream(s)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        out.copy_(linear_act(x, w, b, 'gelu'))
    out.zero_()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, y0)


def test_bert_layer_fused_linear_matches_unfused():
    from kungfu_b200 import models
    torch.manual_seed(0)
    a = models.bert_tiny(fused_linear=False).cuda().to(torch.bfloat16)
    b = models.bert_tiny(fused_linear=True).cuda().to(torch.bfloat16)
    b.load_state_dict(a.state_dict())
    a.eval(), b.eval()
    ids = torch.randint(0, 1000, (4, 64), device='cuda')
    with torch.no_grad():
        ya, yb = a(ids), b(ids)
    ya = ya[0] if isinstance(ya, (tuple, list)) else ya
    yb = yb[0] if isinstance(yb, (tuple, list)) else yb
    assert float((ya.float() - yb.float()).norm() / ya.float().norm()) < 3e-2
