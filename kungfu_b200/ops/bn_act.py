"""Fused BatchNorm (+ residual add) (+ ReLU) for channels-last bf16/fp16
activations (csrc/kernels/bn_act.cu), as a torch.autograd.Function.

Semantics are those of ``relu(batch_norm(x) + residual)`` with fp32 statistics,
biased variance for normalisation and unbiased variance in the running
estimate (torch.nn.BatchNorm2d).
"""
import torch

from . import _cuda as K

_scratch = {}


def _get_scratch(dev, floats):
    key = (dev, torch.cuda.current_stream(dev).cuda_stream if not torch.cuda.is_current_stream_capturing() else 'cap')
    buf = _scratch.get(key)
    if buf is None or buf.numel() < floats:
        buf = torch.empty(max(floats, 1 << 20), dtype=torch.float32, device=dev)
        _scratch[key] = buf
    return buf


def supported(x):
    return (x.is_cuda and x.dim() == 4 and x.dtype in (torch.bfloat16, torch.float16) and x.shape[1] % 8 == 0
            and x.shape[1] <= 2048 and x.is_contiguous(memory_format=torch.channels_last))


class _BNAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, residual, weight, bias, running_mean, running_var, training, momentum, eps, relu):
        N, Cc, H, W = x.shape
        rows = N * H * W
        y = torch.empty_like(x, memory_format=torch.channels_last)
        save_mean = torch.empty(Cc, dtype=torch.float32, device=x.device)
        save_invstd = torch.empty(Cc, dtype=torch.float32, device=x.device)
        floats = K.lib().kfb_bn_scratch_floats(rows, Cc)
        # scratch is private to this call: the backward of another layer may run concurrently on another stream
        scratch = torch.empty(floats, dtype=torch.float32, device=x.device)
        # one bit per element (y > 0): the backward reads this instead of y
        mask = torch.empty(rows * (Cc // 8), dtype=torch.uint8, device=x.device) if relu else None
        if residual is not None and not residual.is_contiguous(memory_format=torch.channels_last):
            residual = residual.contiguous(memory_format=torch.channels_last)
        K.call('kfb_bn_act_fwd', x.data_ptr(), residual.data_ptr() if residual is not None else None,
               y.data_ptr(), weight.data_ptr() if weight is not None else None,
               bias.data_ptr() if bias is not None else None,
               running_mean.data_ptr() if (running_mean is not None and training) else
               (running_mean.data_ptr() if running_mean is not None else None),
               running_var.data_ptr() if running_var is not None else None,
               save_mean.data_ptr(), save_invstd.data_ptr(), rows, Cc, float(eps), float(momentum),
               1 if relu else 0, 1 if training else 0, K.dtype_code(x.dtype), scratch.data_ptr(),
               mask.data_ptr() if mask is not None else None, K.stream_ptr())
        ctx.save_for_backward(x, mask, weight, save_mean, save_invstd)
        ctx.relu = relu
        ctx.has_res = residual is not None
        ctx.training = training
        return y

    @staticmethod
    def backward(ctx, dy):
        x, mask, weight, save_mean, save_invstd = ctx.saved_tensors
        N, Cc, H, W = x.shape
        rows = N * H * W
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dx = torch.empty_like(x, memory_format=torch.channels_last)
        dres = torch.empty_like(x, memory_format=torch.channels_last) if ctx.has_res else None
        dweight = torch.empty(Cc, dtype=torch.float32, device=x.device)
        dbias = torch.empty(Cc, dtype=torch.float32, device=x.device)
        scratch = torch.empty(K.lib().kfb_bn_scratch_floats(rows, Cc), dtype=torch.float32, device=x.device)
        K.call('kfb_bn_act_bwd', dy.data_ptr(), x.data_ptr(), mask.data_ptr() if mask is not None else None,
               weight.data_ptr() if weight is not None else None, save_mean.data_ptr(), save_invstd.data_ptr(),
               dx.data_ptr(), dres.data_ptr() if dres is not None else None, dweight.data_ptr(), dbias.data_ptr(),
               rows, Cc, 1 if ctx.relu else 0, K.dtype_code(x.dtype), scratch.data_ptr(), K.stream_ptr())
        if weight is not None and weight.dtype != torch.float32:
            dweight, dbias = dweight.to(weight.dtype), dbias.to(weight.dtype)
        return dx, dres, dweight if weight is not None else None, dbias if weight is not None else None, \
            None, None, None, None, None, None


def bn_act(x, residual, weight, bias, running_mean, running_var, training, momentum, eps, relu,
           num_batches_tracked=None):
    use_batch_stats = training or running_mean is None
    if not supported(x) or (weight is not None and weight.dtype != torch.float32) or \
            (not use_batch_stats and torch.is_grad_enabled() and x.requires_grad):
        y = torch.nn.functional.batch_norm(x, running_mean, running_var, weight, bias, use_batch_stats,
                                           momentum if momentum is not None else 0.1, eps)
        if residual is not None:
            y = y + residual
        return torch.relu(y) if relu else y
    if training and num_batches_tracked is not None:
        num_batches_tracked.add_(1)
    if momentum is None:  # cumulative moving average
        momentum = 1.0 / float(num_batches_tracked) if num_batches_tracked is not None else 0.1
    return _BNAct.apply(x, residual, weight, bias, running_mean, running_var, use_batch_stats, momentum, eps, relu)
