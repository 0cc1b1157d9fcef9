"""3x3 / stride 2 / pad 1 max pooling for channels-last bf16/fp16
(csrc/kernels/pool.cu) as a torch.autograd.Function."""
import torch
import torch.nn.functional as F

from . import _cuda as K


def supported(x):
    return (x.is_cuda and x.dim() == 4 and x.dtype in (torch.bfloat16, torch.float16) and x.shape[1] % 8 == 0
            and x.is_contiguous(memory_format=torch.channels_last))


class _MaxPool3x3s2(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        N, Cc, H, W = x.shape
        OH, OW = (H - 1) // 2 + 1, (W - 1) // 2 + 1
        y = torch.empty((N, Cc, OH, OW), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
        idx = torch.empty(N * OH * OW * Cc, dtype=torch.uint8, device=x.device)
        K.call('kfb_maxpool3x3s2_fwd', x.data_ptr(), y.data_ptr(), idx.data_ptr(), N, H, W, Cc,
               K.dtype_code(x.dtype), K.stream_ptr())
        ctx.save_for_backward(idx)
        ctx.shape = (N, Cc, H, W)
        return y

    @staticmethod
    def backward(ctx, dy):
        idx, = ctx.saved_tensors
        N, Cc, H, W = ctx.shape
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dx = torch.empty((N, Cc, H, W), dtype=dy.dtype, device=dy.device, memory_format=torch.channels_last)
        K.call('kfb_maxpool3x3s2_bwd', dy.data_ptr(), idx.data_ptr(), dx.data_ptr(), N, H, W, Cc,
               K.dtype_code(dy.dtype), K.stream_ptr())
        return dx


def max_pool_3x3s2(x):
    if supported(x):
        return _MaxPool3x3s2.apply(x)
    return F.max_pool2d(x, 3, 2, 1)


class MaxPool3x3s2(torch.nn.Module):
    """Drop-in for nn.MaxPool2d(3, stride=2, padding=1) (no parameters)."""

    def __init__(self, fused=True):
        super().__init__()
        self.fused = fused

    def forward(self, x):
        return max_pool_3x3s2(x) if self.fused else F.max_pool2d(x, 3, 2, 1)


class _GlobalAvgPool(torch.autograd.Function):
    """(N,C,H,W) channels-last -> (N,C); the backward writes a channels-last
    gradient directly (PyTorch's produces an NCHW expand that then has to be
    re-laid out by a slow strided copy)."""

    @staticmethod
    def forward(ctx, x):
        N, Cc, H, W = x.shape
        y = torch.empty((N, Cc), dtype=x.dtype, device=x.device)
        K.call('kfb_gap_fwd', x.data_ptr(), y.data_ptr(), N, H * W, Cc, K.dtype_code(x.dtype), K.stream_ptr())
        ctx.shape = (N, Cc, H, W)
        return y

    @staticmethod
    def backward(ctx, dy):
        N, Cc, H, W = ctx.shape
        dy = dy.contiguous()
        dx = torch.empty((N, Cc, H, W), dtype=dy.dtype, device=dy.device, memory_format=torch.channels_last)
        K.call('kfb_gap_bwd', dy.data_ptr(), dx.data_ptr(), N, H * W, Cc, K.dtype_code(dy.dtype), K.stream_ptr())
        return dx


def global_avg_pool(x):
    """Flattened global average pool: (N,C,H,W) -> (N,C)."""
    if supported(x):
        return _GlobalAvgPool.apply(x)
    return torch.flatten(F.adaptive_avg_pool2d(x, (1, 1)), 1)
