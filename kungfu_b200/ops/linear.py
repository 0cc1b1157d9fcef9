"""Linear layer with a fused bias + activation epilogue on the tcgen05 tensor
cores (csrc/kernels/linear_tc.cu): ``y = act(x @ w.T + b)`` for bf16 CUDA
tensors, fp32 accumulation in tensor memory.

Forward is the hand-written TMA + tcgen05 kernel (no separate bias / GELU pass
over the [M, N] activation). Backward: ``dz = dy * act'(z)`` is one kernel
(kfb_act_bwd); the two weight/input gradient GEMMs are plain library GEMMs
(cuBLAS through torch.matmul).

The reference has no model kernels of its own (its models are Keras /
torchvision, /root/reference/benchmarks/system/benchmark_kungfu.py:84); this
serves the BERT benchmark model of this repo.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _cuda as K

ACT_NONE, ACT_GELU, ACT_RELU = 0, 1, 2
_ACT = {None: ACT_NONE, 'none': ACT_NONE, 'gelu': ACT_GELU, 'relu': ACT_RELU}


def set_tiling(mode):
    """'auto' (default: CTA pairs whenever M > 128), 'single' (one CTA per 128x256
    tile, cta_group::1) or 'pair' (two CTAs per 256x256 tile, cta_group::2)."""
    K.lib().kfb_linear_act_set_mode({'auto': -1, 'single': 0, 'pair': 1}[mode])


def supported(x, weight, bias=None):
    """bf16 CUDA operands, K and N multiples of 8, dense row-major."""
    if not (x.is_cuda and weight.is_cuda and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16):
        return False
    if bias is not None and bias.dtype not in (torch.bfloat16, torch.float32):
        return False
    N, Kd = weight.shape
    return x.shape[-1] == Kd and Kd % 8 == 0 and N % 8 == 0 and weight.is_contiguous() and x.numel() > 0


def _launch(x2, weight, bias, act, save_z):
    M, Kd = x2.shape
    N = weight.shape[0]
    y = torch.empty((M, N), dtype=torch.bfloat16, device=x2.device)
    z = torch.empty_like(y) if save_z else None
    K.call('kfb_linear_act', x2.data_ptr(), weight.data_ptr(),
           bias.data_ptr() if bias is not None else None,
           K.dtype_code(bias.dtype) if bias is not None else 0,
           y.data_ptr(), z.data_ptr() if z is not None else None, M, N, Kd, act, K.stream_ptr())
    return y, z


class _LinearAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, act):
        x2 = x.reshape(-1, x.shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        need_grad = any(ctx.needs_input_grad[:3])
        y, z = _launch(x2, weight, bias, act, save_z=need_grad and act != ACT_NONE)
        ctx.act = act
        ctx.has_bias = bias is not None
        ctx.bias_dtype = bias.dtype if bias is not None else None
        ctx.xshape = x.shape
        if need_grad:
            ctx.save_for_backward(x2, weight, z if z is not None else y.new_empty(0))
        return y.view(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, weight, z = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        if ctx.act != ACT_NONE:
            dz = torch.empty_like(dy2)
            K.call('kfb_act_bwd', dy2.data_ptr(), z.data_ptr(), dz.data_ptr(), dz.numel(), ctx.act, K.stream_ptr())
        else:
            dz = dy2
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = (dz @ weight).view(ctx.xshape)
        if ctx.needs_input_grad[1]:
            dw = dz.t() @ x2
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = dz.sum(0, dtype=torch.float32).to(ctx.bias_dtype)
        return dx, dw, db, None


def linear_act(x, weight, bias=None, act=None):
    """act(x @ weight.T + bias); act in {None, 'gelu', 'relu'}. Falls back to
    PyTorch ops for anything the kernel does not take (CPU, fp32, odd shapes)."""
    code = _ACT[act]
    if supported(x, weight, bias):
        return _LinearAct.apply(x, weight, bias, code)
    y = F.linear(x, weight, bias)
    if code == ACT_GELU:
        return F.gelu(y)
    if code == ACT_RELU:
        return F.relu(y)
    return y


class LinearAct(nn.Linear):
    """nn.Linear (same parameters / state_dict) with an optional fused activation."""

    def __init__(self, in_features, out_features, bias=True, act=None, fused=True, **kw):
        super().__init__(in_features, out_features, bias=bias, **kw)
        self.act = act
        self.fused = fused

    def forward(self, x):
        if self.fused:
            return linear_act(x, self.weight, self.bias, self.act)
        y = F.linear(x, self.weight, self.bias)
        return F.gelu(y) if self.act == 'gelu' else F.relu(y) if self.act == 'relu' else y
