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
    def manual_seed(cls, seed, rank=0):
        """Workers of one job should pass their rank: same seed + different rank = decorrelated masks."""
        cls.seed = (int(seed) + 0x9E3779B97F4A7C15 * (int(rank) + 1)) & 0xFFFFFFFFFFFFFFFF

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


# ---- gradient-side helpers for transformer blocks ----------------------------------------------

_colsum_scratch = {}


def col_sum(a2):
    """Column sums of a 2-D bf16 CUDA tensor (a bias gradient) with the deterministic single-launch
    kernel (kfb_colsum); anything else goes to torch.sum."""
    M, N = a2.shape
    if not (a2.is_cuda and a2.dtype == torch.bfloat16 and N % 64 == 0 and M > 0 and a2.is_contiguous()):
        return a2.sum(0)
    key = (a2.device.index, N)
    scratch = _colsum_scratch.get(key)
    if scratch is None:
        scratch = torch.zeros(K.lib().kfb_colsum_scratch_floats(N), dtype=torch.float32, device=a2.device)
        _colsum_scratch[key] = scratch
    out = torch.empty(N, dtype=a2.dtype, device=a2.device)
    K.call('kfb_colsum', a2.data_ptr(), M, N, out.data_ptr(), scratch.data_ptr(), K.stream_ptr())
    return out


class _LinearFastBiasGrad(torch.autograd.Function):
    """F.linear whose backward takes the bias gradient from col_sum instead of a generic reduction
    (all three GEMMs stay library GEMMs)."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        ctx.save_for_backward(x, weight)
        ctx.has_bias = bias is not None
        return F.linear(x, weight, bias)

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        dy2 = dy2 if dy2.is_contiguous() else dy2.contiguous()
        x2 = x.reshape(-1, x.shape[-1])
        dx = (dy2 @ weight).view(x.shape) if ctx.needs_input_grad[0] else None
        dw = dy2.t() @ x2 if ctx.needs_input_grad[1] else None
        db = col_sum(dy2) if (ctx.has_bias and ctx.needs_input_grad[2]) else None
        return dx, dw, db


def linear_fast_bias_grad(x, weight, bias):
    if x.is_cuda and x.dtype == torch.bfloat16 and bias is not None:
        return _LinearFastBiasGrad.apply(x, weight, bias)
    return F.linear(x, weight, bias)


class _SplitQKV(torch.autograd.Function):
    """[B, S, 3*H] -> q, k, v views of shape [B, heads, S, hd]. The backward gathers dq, dk, dv
    straight into the packed [B, S, 3*H] layout with one kernel (autograd would stack them and then
    run a permuting copy: two extra passes)."""

    @staticmethod
    def forward(ctx, qkv, heads):
        B, S, H3 = qkv.shape
        hd = H3 // 3 // heads
        ctx.dims = (B, S, heads, hd)
        t = qkv.view(B, S, 3, heads, hd)
        return t[:, :, 0].transpose(1, 2), t[:, :, 1].transpose(1, 2), t[:, :, 2].transpose(1, 2)

    @staticmethod
    def backward(ctx, dq, dk, dv):
        B, S, heads, hd = ctx.dims
        ref = next(g for g in (dq, dk, dv) if g is not None)
        gs = []
        for g in (dq, dk, dv):
            if g is None:
                g = torch.zeros((B, heads, S, hd), dtype=ref.dtype, device=ref.device)
            if g.stride(-1) != 1 or any(s % 8 for s in g.stride()[:3]) or g.data_ptr() % 16:
                g = g.contiguous()
            gs.append(g)
        out = torch.empty((B, S, 3 * heads * hd), dtype=ref.dtype, device=ref.device)
        strides = []
        for g in gs:
            strides += [g.stride(0), g.stride(1), g.stride(2)]
        arr = (K.C.c_longlong * 9)(*strides)
        K.call('kfb_pack_qkv_grad', gs[0].data_ptr(), gs[1].data_ptr(), gs[2].data_ptr(), arr, B, heads, S, hd,
               out.data_ptr(), K.stream_ptr())
        return out, None


def split_qkv(qkv, heads):
    """q, k, v [B, heads, S, hd] from a packed projection [B, S, 3*H]."""
    B, S, H3 = qkv.shape
    hd = H3 // 3 // heads
    if qkv.is_cuda and qkv.dtype in (torch.bfloat16, torch.float16) and hd % 8 == 0 and qkv.is_contiguous():
        return _SplitQKV.apply(qkv, heads)
    return qkv.view(B, S, 3, heads, hd).permute(2, 0, 3, 1, 4)
