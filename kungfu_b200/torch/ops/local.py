"""Local store ops. Reference: /root/reference/srcs/python/kungfu/tensorflow/ops/local.py,
/root/reference/srcs/python/kungfu/tensorflow/ops/p2p.py."""
import torch

from .collective import request_variable, save_variable


def save_variables(tensors, names=None, version=None):
    """Publish several tensors in this worker's store under ``names`` (default v0, v1, ...), optionally as ``version``."""
    for i, t in enumerate(tensors):
        save_variable(t, names[i] if names else 'v%d' % i, version)


def save_model(named_tensors, name='kungfu::fused-model', version=None):
    """Publish all tensors as ONE fused fp32 blob (what PairAveraging peers pull); returns its element count."""
    flat = torch.cat([t.detach().reshape(-1).float().cpu() for _, t in named_tensors])
    save_variable(flat, name, version)
    return flat.numel()


def request_variable_with_template(target, template, name, version=None):
    """request_variable with the shape / dtype taken from ``template``."""
    return request_variable(target, name, template, version)


def request_model(target, named_tensors, name='kungfu::fused-model', version=None):
    """Pull a peer's fused model and split it back into tensors shaped like ``named_tensors``: (ok, [tensors])."""
    total = sum(t.numel() for _, t in named_tensors)
    ok, flat = request_variable(target, name, torch.empty(total), version)
    if not ok:
        return False, None
    out, off = [], 0
    for _, t in named_tensors:
        out.append(flat[off:off + t.numel()].view(t.shape).to(t.dtype))
        off += t.numel()
    return True, out
