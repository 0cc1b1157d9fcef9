"""Local store ops. Reference: /root/reference/srcs/python/kungfu/tensorflow/ops/local.py,
/root/reference/srcs/python/kungfu/tensorflow/ops/p2p.py."""
import torch

from .collective import request_variable, save_variable


def save_variables(tensors, names=None, version=None):
    for i, t in enumerate(tensors):
        save_variable(t, names[i] if names else 'v%d' % i, version)


def save_model(named_tensors, name='kungfu::fused-model', version=None):
    flat = torch.cat([t.detach().reshape(-1).float().cpu() for _, t in named_tensors])
    save_variable(flat, name, version)
    return flat.numel()


def request_variable_with_template(target, template, name, version=None):
    return request_variable(target, name, template, version)


def request_model(target, named_tensors, name='kungfu::fused-model', version=None):
    total = sum(t.numel() for _, t in named_tensors)
    ok, flat = request_variable(target, name, torch.empty(total), version)
    if not ok:
        return False, None
    out, off = [], 0
    for _, t in named_tensors:
        out.append(flat[off:off + t.numel()].view(t.shape).to(t.dtype))
        off += t.numel()
    return True, out
