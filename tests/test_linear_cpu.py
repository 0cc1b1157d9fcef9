"""CPU-side contract of kungfu_b200.ops.linear (the CUDA kernel itself is covered by tests/test_gpu_linear_tc.py)."""
import torch
import torch.nn as nn
import torch.nn.functional as F


def test_linear_act_falls_back_to_torch_ops_off_gpu():
    from kungfu_b200.ops.linear import linear_act, supported
    x, w, b = torch.randn(5, 7, 16), torch.randn(24, 16), torch.randn(24)
    assert not supported(x, w, b)
    assert torch.allclose(linear_act(x, w, b, 'gelu'), F.gelu(F.linear(x, w, b)))
    assert torch.allclose(linear_act(x, w, b, 'relu'), F.relu(F.linear(x, w, b)))
    assert torch.allclose(linear_act(x, w, None, None), F.linear(x, w))


def test_linear_act_module_is_state_dict_compatible_with_nn_linear():
    from kungfu_b200.ops.linear import LinearAct
    a, b = nn.Linear(16, 24), LinearAct(16, 24, act='gelu')
    assert list(a.state_dict().keys()) == list(b.state_dict().keys())
    b.load_state_dict(a.state_dict())
    x = torch.randn(3, 16)
    assert torch.allclose(b(x), F.gelu(a(x)))
    b.fused = False
    assert torch.allclose(b(x), F.gelu(a(x)))


def test_bert_fused_linear_flag_keeps_parameters_and_cpu_results():
    from kungfu_b200 import models
    torch.manual_seed(0)
    a, b = models.bert_tiny(fused_linear=False), models.bert_tiny(fused_linear=True)
    assert [k for k, _ in a.named_parameters()] == [k for k, _ in b.named_parameters()]
    b.load_state_dict(a.state_dict())
    a.eval(), b.eval()
    ids = torch.randint(0, 1000, (2, 12))
    assert torch.equal(a(ids)[0], b(ids)[0])
