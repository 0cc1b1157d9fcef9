"""Fused add + dropout + LayerNorm (csrc/kernels/ln_fused.cu) against a plain fp32 PyTorch reference."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _mk(M, H, seed=0):
    g = torch.Generator(device='cuda').manual_seed(seed)
    h = torch.randn(M, H, device='cuda', generator=g).to(torch.bfloat16)
    x = torch.randn(M, H, device='cuda', generator=g).to(torch.bfloat16)
    gamma = (1 + 0.1 * torch.randn(H, device='cuda', generator=g)).to(torch.bfloat16)
    beta = (0.1 * torch.randn(H, device='cuda', generator=g)).to(torch.bfloat16)
    return h, x, gamma, beta


def _ref(h, x, gamma, beta, eps):
    s = (x.float() + h.float()) if x is not None else h.float()
    s = s.to(torch.bfloat16).float()  # the kernel normalises the bf16-rounded sum (what it saves for backward)
    return F.layer_norm(s, (h.shape[-1], ), gamma.float(), beta.float(), eps)


@pytest.mark.parametrize('M,H', [(4096, 768), (1000, 1024), (37, 256), (1, 512), (5000, 768)])
@pytest.mark.parametrize('with_x', [True, False])
def test_forward_without_dropout_matches_fp32_reference(M, H, with_x):
    from kungfu_b200.ops.fused_ln import add_dropout_layer_norm, supported
    h, x, gamma, beta = _mk(M, H)
    x = x if with_x else None
    assert supported(h, x, gamma, beta)
    y = add_dropout_layer_norm(h, x, gamma, beta, p=0.1, training=False, eps=1e-12)
    ref = _ref(h, x, gamma, beta, 1e-12)
    assert y.dtype == torch.bfloat16 and y.shape == h.shape
    assert float((y.float() - ref).abs().max()) < 3e-2
    assert float((y.float() - ref).norm() / ref.norm()) < 4e-3


@pytest.mark.parametrize('M,H', [(2048, 768), (300, 1024), (64, 256)])
def test_backward_without_dropout_matches_fp32_autograd(M, H):
    from kungfu_b200.ops.fused_ln import add_dropout_layer_norm
    h, x, gamma, beta = _mk(M, H, seed=3)
    for t in (h, x, gamma, beta):
        t.requires_grad_(True)
    y = add_dropout_layer_norm(h.view(4, M // 4, H), x.view(4, M // 4, H), gamma, beta, p=0.0, training=True, eps=1e-5)
    dy = torch.randn_like(y)
    y.backward(dy)
    hf, xf, gf, bf = [t.detach().float().requires_grad_(True) for t in (h, x, gamma, beta)]
    F.layer_norm(xf + hf, (H, ), gf, bf, 1e-5).backward(dy.float().view(M, H))
    for got, want, name, tol in ((h.grad, hf.grad, 'dh', 1e-2), (x.grad, xf.grad, 'dx', 1e-2),
                                 (gamma.grad, gf.grad, 'dgamma', 2e-2), (beta.grad, bf.grad, 'dbeta', 2e-2)):
        rel = float((got.float() - want).norm() / want.norm())
        assert rel < tol, '%s rel err %g' % (name, rel)


def test_dropout_mask_is_consistent_between_forward_and_backward_and_changes_with_the_step():
    from kungfu_b200.ops.fused_ln import DropoutState, add_dropout_layer_norm
    M, H, p = 1024, 768, 0.25
    h, x, gamma, beta = _mk(M, H, seed=5)
    h = h.abs() + 1  # no zeros, so dropped positions are recognisable
    x = torch.zeros_like(x)
    gamma, beta = torch.ones_like(gamma), torch.zeros_like(beta)
    h.requires_grad_(True)
    DropoutState.manual_seed(1234)
    y = add_dropout_layer_norm(h, x, gamma, beta, p=p, training=True, eps=1e-5, stream_id=7)
    y.backward(torch.randn_like(y))
    # the same call again: same step, same stream -> identical mask and output
    y2 = add_dropout_layer_norm(h.detach(), x, gamma, beta, p=p, training=True, eps=1e-5, stream_id=7)
    assert torch.equal(y, y2)
    # dropped elements are the row minimum of s = dropout(h) (= 0 < every kept element); recover the mask from y:
    # kept elements of one row share the same (s - mean) * rstd scale, dropped ones all equal the row minimum
    dropped = y.detach() == y.detach().min(dim=1, keepdim=True).values
    frac = float(dropped.float().mean())
    assert abs(frac - p) < 0.02, frac
    # backward zeroes exactly the dropped positions
    assert bool((h.grad[dropped] == 0).all())
    assert float((h.grad[~dropped] != 0).float().mean()) > 0.99
    # another call site (stream id) or another step draws another mask
    y3 = add_dropout_layer_norm(h.detach(), x, gamma, beta, p=p, training=True, eps=1e-5, stream_id=8)
    assert not torch.equal(y, y3)
    DropoutState.advance(h.device)
    y4 = add_dropout_layer_norm(h.detach(), x, gamma, beta, p=p, training=True, eps=1e-5, stream_id=7)
    assert not torch.equal(y, y4)


def test_bert_fused_ln_matches_unfused_in_eval_and_trains():
    from kungfu_b200 import models
    torch.manual_seed(0)
    kw = dict(vocab_size=1000, hidden=256, layers=2, heads=4, intermediate=1024, max_pos=128)
    a = models.bert_base(fused_ln=False, **kw).cuda().to(torch.bfloat16)
    b = models.bert_base(fused_ln=True, **kw).cuda().to(torch.bfloat16)
    b.load_state_dict(a.state_dict())
    ids = torch.randint(0, 1000, (4, 64), device='cuda')
    a.eval(), b.eval()
    with torch.no_grad():
        ya, yb = a(ids)[0], b(ids)[0]
    assert float((ya.float() - yb.float()).norm() / ya.float().norm()) < 3e-2
    b.train()
    labels = torch.randint(0, 1000, (4, 64), device='cuda')
    loss = b(ids, mlm_labels=labels, nsp_labels=torch.randint(0, 2, (4, ), device='cuda'))
    loss.backward()
    assert torch.isfinite(loss)
    assert all(p.grad is not None and bool(torch.isfinite(p.grad).all()) for p in b.parameters())


@pytest.mark.parametrize('M,N', [(4096, 768), (4096, 3072), (640, 2304), (3, 64), (5000, 1024)])
def test_col_sum_matches_fp32_and_is_deterministic(M, N):
    from kungfu_b200.ops.fused_ln import col_sum
    g = torch.Generator(device='cuda').manual_seed(M + N)
    a = torch.randn(M, N, device='cuda', generator=g).to(torch.bfloat16)
    out = col_sum(a)
    ref = a.float().sum(0)
    assert out.shape == (N, ) and out.dtype == torch.bfloat16
    assert float((out.float() - ref).abs().max()) <= 0.01 * float(ref.abs().max()) + 0.05
    for _ in range(3):  # tickets reset themselves: repeated calls on the same scratch agree bit for bit
        assert torch.equal(col_sum(a), out)


def test_split_qkv_forward_views_and_backward_gather():
    from kungfu_b200.ops.fused_ln import split_qkv
    B, S, heads, hd = 3, 40, 4, 64
    qkv = torch.randn(B, S, 3 * heads * hd, device='cuda').to(torch.bfloat16).requires_grad_(True)
    q, k, v = split_qkv(qkv, heads)
    qr, kr, vr = qkv.detach().view(B, S, 3, heads, hd).permute(2, 0, 3, 1, 4)
    assert torch.equal(q, qr) and torch.equal(k, kr) and torch.equal(v, vr)
    gq, gk, gv = [torch.randn(B, heads, S, hd, device='cuda').to(torch.bfloat16) for _ in range(3)]
    # one gradient arrives transposed in memory (strided), like attention backward may produce
    gk_strided = gk.transpose(1, 2).contiguous().transpose(1, 2)
    torch.autograd.backward([q, k, v], [gq, gk_strided, gv])
    ref = torch.stack([gq, gk, gv], 0).permute(1, 3, 0, 2, 4).reshape(B, S, 3 * heads * hd)
    assert torch.equal(qkv.grad, ref)


def test_linear_fast_bias_grad_matches_autograd():
    from kungfu_b200.ops.fused_ln import linear_fast_bias_grad
    x = torch.randn(6, 100, 256, device='cuda').to(torch.bfloat16).requires_grad_(True)
    w = (torch.randn(512, 256, device='cuda') / 16).to(torch.bfloat16).requires_grad_(True)
    b = torch.randn(512, device='cuda').to(torch.bfloat16).requires_grad_(True)
    y = linear_fast_bias_grad(x, w, b)
    dy = torch.randn_like(y)
    y.backward(dy)
    x2, w2, b2 = [t.detach().clone().requires_grad_(True) for t in (x, w, b)]
    F.linear(x2, w2, b2).backward(dy)
    assert torch.equal(y, F.linear(x2, w2, b2))
    assert float((x.grad.float() - x2.grad.float()).norm() / x2.grad.float().norm()) < 1e-2
    assert float((w.grad.float() - w2.grad.float()).norm() / w2.grad.float().norm()) < 1e-2
    assert float((b.grad.float() - b2.grad.float()).norm() / b2.grad.float().norm()) < 1e-2


def test_bert_fused_grads_match_unfused_grads():
    from kungfu_b200 import models
    torch.manual_seed(0)
    kw = dict(vocab_size=1000, hidden=256, layers=2, heads=4, intermediate=1024, max_pos=128, dropout=0.0)
    a = models.bert_base(**kw).cuda().to(torch.bfloat16)
    b = models.bert_base(fused_grads=True, fused_ln=True, **kw).cuda().to(torch.bfloat16)
    b.load_state_dict(a.state_dict())
    ids = torch.randint(0, 1000, (4, 64), device='cuda')
    labels = torch.randint(0, 1000, (4, 64), device='cuda')
    nsp = torch.randint(0, 2, (4, ), device='cuda')
    la = a(ids, mlm_labels=labels, nsp_labels=nsp)
    lb = b(ids, mlm_labels=labels, nsp_labels=nsp)
    la.backward(), lb.backward()
    assert abs(float(la) - float(lb)) < 5e-2
    for (n, p), q in zip(a.named_parameters(), b.parameters()):
        rel = float((p.grad.float() - q.grad.float()).norm() / p.grad.float().norm().clamp_min(1e-6))
        assert rel < 0.15, '%s: %g' % (n, rel)
