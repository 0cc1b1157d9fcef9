"""Flat parameter arenas (kungfu_b200.parallel.flat) on the host: a fake communicator hands out CPU buffers, so the
layout logic - offsets, padding, channels-last aware re-pointing of parameters into the flat buffer, gradient views - is
checked without a GPU."""
import torch
import torch.nn as nn


class _Buf(object):
    def __init__(self, nbytes):
        self.t = torch.zeros(nbytes, dtype=torch.uint8)
        self.mc_ptr = 0

    def view(self, dtype, r=0):
        return self.t.view(dtype)


class _Comm(object):
    np, rank, n_local_ranks, emulated = 1, 0, 1, False
    device = torch.device('cpu')

    def __init__(self):
        self.allocs = []

    def alloc(self, nbytes, multicast=True):
        self.allocs.append(nbytes)
        return _Buf(nbytes)


def _model():
    torch.manual_seed(0)
    m = nn.Sequential(nn.Conv2d(3, 8, 3, padding=1, bias=False), nn.BatchNorm2d(8), nn.ReLU(), nn.Conv2d(8, 5, 1),
                      nn.AdaptiveAvgPool2d(1), nn.Flatten(), nn.Linear(5, 7))
    return m.to(memory_format=torch.channels_last)


def test_arena_repoints_parameters_without_changing_the_model():
    from kungfu_b200.parallel.flat import ALIGN_ELEMS, build_arenas
    m = _model()
    x = torch.randn(2, 3, 6, 6).contiguous(memory_format=torch.channels_last)
    before = m(x).detach().clone()
    values = {n: p.detach().clone() for n, p in m.named_parameters()}
    comm = _Comm()
    arenas = build_arenas(comm, list(m.named_parameters()))
    assert len(arenas) == 1  # one dtype
    a = arenas[0]
    assert a.count % (ALIGN_ELEMS * 8) == 0 and a.nbytes == a.count * 4 and comm.allocs == [a.nbytes, a.nbytes]
    assert all(off % ALIGN_ELEMS == 0 for off in a.offsets)
    assert a.offsets == sorted(a.offsets) and a.offsets[-1] + a.params[-1].numel() <= a.count
    base = a.flat_params.data_ptr()
    for (n, p), off in zip(m.named_parameters(), a.offsets):
        assert torch.equal(p.detach(), values[n]), n            # values survive
        assert p.data_ptr() == base + off * 4, n                 # and now live inside the flat buffer
        if p.dim() == 4:                                         # conv weights keep their channels-last strides
            assert p.is_contiguous(memory_format=torch.channels_last), n
    assert torch.equal(m(x), before)
    # writing the flat buffer IS writing the parameters (what the fused optimizer step relies on)
    a.flat_params.zero_()
    assert all(float(p.detach().abs().sum()) == 0 for p in m.parameters())


def test_gradient_views_match_parameter_layout_and_segments_cover_the_arena():
    from kungfu_b200.parallel.flat import build_arenas
    m = _model()
    a = build_arenas(_Comm(), list(m.named_parameters()))[0]
    for i, p in enumerate(a.params):
        g = a.grad_view(i)
        assert g.shape == p.shape and g.stride() == p.stride()
        g.fill_(float(i + 1))
    for i, (p, off) in enumerate(zip(a.params, a.offsets)):
        assert bool((a.flat_grads[off:off + p.numel()] == float(i + 1)).all())
    ends = a.segment_ends().tolist()
    assert ends == sorted(ends) and ends[-1] == a.count and len(ends) == len(a.params)
    assert all(e > off for e, off in zip(ends, a.offsets))
