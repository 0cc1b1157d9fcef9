"""``LayerNorm(x + dropout(h))`` - the transformer block's "add & norm" - as one
kernel forward and one (plus a tiny finalize) backward (csrc/kernels/ln_fused.cu).

The dropout mask is not stored: both directions regenerate it from a
counter-based generator keyed by ``(seed, stream_id, step, element)``.
``DropoutState`` holds the seed and the per-device step counter (a device
scalar, so a captured CUDA graph sees fresh masks on every replay as long as
``advance()`` - itself one tiny device op - is part of the captured step).

No counterpart in the reference (its models are TensorFlow's); used by
kungfu_b200/models/bert.py when ``fused_ln=True``.
"""
import torch
import torch.nn.functional as F

from . import _cuda as K

_SUPPORTED_H = (256, 512, 768, 1024)


class DropoutState(object):
    """Seed + device-resident step counter shared by every fused dropout of a model."""
    seed = 0x5DEECE66D
    _steps = {}

    @classmethod
    def manual_seed(cls, seed):
        cls.seed = int(seed) & 0xFFFFFFFFFFFFFFFF

    @classmethod
    def step_tensor(cls, device):
        key = (device.type, device.index)
        t = cls._steps.get(key)
        if t is None:
            t = torch.zeros(1, dtype=torch.int64, device=device)
            cls._steps[key] = t
        return t

    @classmethod
    def advance(cls, device):
        """Call once per training step (before the forward)."""
        cls.step_tensor(device).add_(1)


def supported(h, x, gamma, beta):
    if not (h.is_cuda and h.dtype == torch.bfloat16 and gamma.dtype == torch.bfloat16 and beta.dtype == torch.bfloat16):
        return False
    if h.shape[-1] not in _SUPPORTED_H or h.numel() == 0:
        return False
    return x is None or (x.shape == h.shape and x.dtype == h.dtype)


class _AddDropoutLN(torch.autograd.Function):
    @staticmethod
    def forward(ctx, h, x, gamma, beta, p, eps, stream_id):
        H = h.shape[-1]
        h2 = h.reshape(-1, H)
        h2 = h2 if h2.is_contiguous() else h2.contiguous()
        x2 = None
        if x is not None:
            x2 = x.reshape(-1, H)
            x2 = x2 if x2.is_contiguous() else x2.contiguous()
        M = h2.shape[0]
        need_grad = any(ctx.needs_input_grad[:4])
        y = torch.empty_like(h2)
        s = torch.empty_like(h2) if need_grad else None
        stats = torch.empty(2, M, dtype=torch.float32, device=h.device) if need_grad else None
        step = DropoutState.step_tensor(h.device)
        K.call('kfb_add_dropout_ln_fwd', h2.data_ptr(), x2.data_ptr() if x2 is not None else None, gamma.data_ptr(),
               beta.data_ptr(), y.data_ptr(), s.data_ptr() if s is not None else None,
               stats[0].data_ptr() if stats is not None else None, stats[1].data_ptr() if stats is not None else None,
               M, H, float(eps), float(p), DropoutState.seed, step.data_ptr(), int(stream_id), K.stream_ptr())
        if need_grad:
            ctx.save_for_backward(s, stats, gamma)
        ctx.p, ctx.stream_id, ctx.shape, ctx.has_x = float(p), int(stream_id), h.shape, x is not None
        return y.view(h.shape)

    @staticmethod
    def backward(ctx, dy):
        s, stats, gamma = ctx.saved_tensors
        M, H = s.shape
        dy2 = dy.reshape(M, H)
        dy2 = dy2 if dy2.is_contiguous() else dy2.contiguous()
        want_h, want_x = ctx.needs_input_grad[0], ctx.has_x and ctx.needs_input_grad[1]
        dx = torch.empty_like(s) if (want_x or (want_h and ctx.p == 0.0)) else None
        # without dropout dh == dx: write it once
        dh = torch.empty_like(s) if (want_h and ctx.p > 0.0) else None
        dgamma, dbeta = torch.empty_like(gamma), torch.empty_like(gamma)
        partial = torch.empty(K.lib().kfb_ln_partial_floats(M, H), dtype=torch.float32, device=s.device)
        step = DropoutState.step_tensor(s.device)
        K.call('kfb_add_dropout_ln_bwd', dy2.data_ptr(), s.data_ptr(), stats[0].data_ptr(), stats[1].data_ptr(),
               gamma.data_ptr(), dx.data_ptr() if dx is not None else None, dh.data_ptr() if dh is not None else None,
               dgamma.data_ptr(), dbeta.data_ptr(), partial.data_ptr(), M, H, ctx.p, DropoutState.seed,
               step.data_ptr(), ctx.stream_id, K.stream_ptr())
        gh = (dh if dh is not None else dx) if want_h else None
        gx = dx if want_x else None
        return (gh.view(ctx.shape) if gh is not None else None, gx.view(ctx.shape) if gx is not None else None,
                dgamma, dbeta, None, None, None)


def add_dropout_layer_norm(h, x, gamma, beta, p=0.0, training=True, eps=1e-5, stream_id=0):
    """``layer_norm(x + dropout(h, p))``; ``x`` may be None. Falls back to PyTorch ops for anything
    the kernel does not take (CPU, fp32, hidden sizes other than 256/512/768/1024)."""
    p_eff = float(p) if training else 0.0
    if supported(h, x, gamma, beta):
        return _AddDropoutLN.apply(h, x, gamma, beta, p_eff, eps, stream_id)
    s = F.dropout(h, p_eff, True) if p_eff > 0 else h
    if x is not None:
        s = x + s
    return F.layer_norm(s, (h.shape[-1], ), gamma, beta, eps)
