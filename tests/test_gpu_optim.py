"""Fused optimizer kernels vs plain PyTorch fp32 references."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

from kungfu_b200.ops import _cuda as K
from kungfu_b200.optim import FusedSGD
from kungfu_b200.parallel.comm import EmulatedComm, LocalComm


def _hyper(lr=0.1, momentum=0.9, wd=1e-4, nesterov=False, first=False, dampening=0.0):
    h = K.SgdHyper()
    h.lr, h.momentum, h.dampening, h.weight_decay = lr, momentum, dampening, wd
    h.nesterov, h.first_step = int(nesterov), int(first)
    return h


def _ref_sgd(p, m, g, h):
    d = g + h.weight_decay * p
    if h.momentum != 0:
        m = d.clone() if h.first_step else h.momentum * m + (1 - h.dampening) * d
        d = d + h.momentum * m if h.nesterov else m
    return p - h.lr * d, m


@pytest.mark.parametrize('np_', [1, 2, 4, 8, 3])
@pytest.mark.parametrize('gdt,pdt', [(torch.bfloat16, torch.bfloat16), (torch.float32, torch.float32)])
def test_ssgd_fused_step(np_, gdt, pdt):
    torch.manual_seed(np_)
    comm = EmulatedComm(np_) if np_ > 1 else LocalComm()
    n = 8 * 3000
    ges, pes = torch.empty(0, dtype=gdt).element_size(), torch.empty(0, dtype=pdt).element_size()
    grads = comm.alloc(n * ges)
    params = comm.alloc(n * pes)
    R = max(np_, 1)
    gs = [torch.randn(n, device='cuda').to(gdt) for _ in range(R)]
    p0 = torch.randn(n, device='cuda')
    m0 = torch.randn(n, device='cuda')
    for r in range(R):
        grads.view(gdt, r if np_ > 1 else None)[:n].copy_(gs[r])
    master = p0.repeat(R).contiguous()
    mom = m0.repeat(R).contiguous()
    ss = torch.zeros(R, device='cuda')
    for nesterov, first in [(False, False), (True, False), (False, True)]:
        h = _hyper(nesterov=nesterov, first=first)
        master.copy_(p0.repeat(R))
        mom.copy_(m0.repeat(R))
        ga, _1 = grads.ptr_array()
        pa, _2 = params.ptr_array()
        K.call('kfb_ssgd_fused_step', comm.cstruct, ga, K.dtype_code(gdt), pa, K.dtype_code(pdt),
               master.data_ptr(), mom.data_ptr(), n, 1.0 / R, C.byref(h), ss.data_ptr(),
               comm.default_blocks, 256, K.stream_ptr())
        torch.cuda.synchronize()
        gavg = torch.stack([g.float() for g in gs]).sum(0) / R
        pref, mref = _ref_sgd(p0, m0, gavg, h)
        tol = 1e-5 if pdt == torch.float32 else 2e-2
        for r in range(R):
            got = params.view(pdt, r if np_ > 1 else None)[:n].float()
            assert torch.allclose(got, pref, rtol=tol, atol=tol), (r, (got - pref).abs().max())
        # each rank updated only its own shard of master/momentum
        npk = n // 8
        per = (npk + R - 1) // R * 8
        for r in range(R):
            lo, hi = min(r * per, n), min((r + 1) * per, n)
            assert torch.allclose(master[r * n + lo:r * n + hi], pref[lo:hi], rtol=1e-5, atol=1e-6)
            assert torch.allclose(mom[r * n + lo:r * n + hi], mref[lo:hi], rtol=1e-5, atol=1e-6)
        assert torch.allclose(ss, (gavg * gavg).sum().expand(R), rtol=1e-3)


@pytest.mark.parametrize('np_', [2, 4])
def test_sma_fused_step(np_):
    torch.manual_seed(7)
    comm = EmulatedComm(np_)
    n = 8 * 2500
    alpha = 0.1
    params = comm.alloc(n * 2)
    stage = comm.alloc(n * 4)
    grads = comm.alloc(n * 2)
    ps = [torch.randn(n, device='cuda').bfloat16() for _ in range(np_)]
    gs = [torch.randn(n, device='cuda').bfloat16() for _ in range(np_)]
    for r in range(np_):
        params.view(torch.bfloat16, r)[:n].copy_(ps[r])
        grads.view(torch.bfloat16, r)[:n].copy_(gs[r])
    master = torch.cat([p.float() for p in ps]).contiguous()
    mom = torch.zeros_like(master)
    h = _hyper(momentum=0.9, wd=0.0, first=True)
    pa, _1 = params.ptr_array()
    sa, _2 = stage.ptr_array()
    ga, _3 = grads.ptr_array()
    K.call('kfb_sma_fused_step', comm.cstruct, pa, K.BF16, sa, C.cast(ga, C.c_void_p), K.BF16,
           master.data_ptr(), mom.data_ptr(), n, alpha, 1.0, C.byref(h), comm.default_blocks, 256,
           K.stream_ptr())
    torch.cuda.synchronize()
    avg = torch.stack([p.float() for p in ps]).mean(0)
    for r in range(np_):
        v = (1 - alpha) * ps[r].float() + alpha * avg
        ref, _ = _ref_sgd(v, torch.zeros_like(v), gs[r].float(), h)
        got = params.view(torch.bfloat16, r)[:n].float()
        assert torch.allclose(got, ref, rtol=2e-2, atol=2e-2), (got - ref).abs().max()
        assert torch.allclose(master[r * n:(r + 1) * n], ref, rtol=1e-4, atol=1e-4)


def test_pair_avg_step_and_publish():
    torch.manual_seed(3)
    comm = LocalComm()
    n = 8 * 4000
    nbytes = n * 2
    # a fake "peer": second pad + snapshot area on the same device
    peer_pad = torch.zeros(K.PAD_BYTES, dtype=torch.uint8, device='cuda')
    comm._struct.np = 2
    comm._struct.pads[1] = peer_pad.data_ptr()
    snap_self = torch.zeros(2 * nbytes, dtype=torch.uint8, device='cuda')
    snap_peer = torch.zeros(2 * nbytes, dtype=torch.uint8, device='cuda')
    vt = torch.randn(n, device='cuda').bfloat16()
    # peer published version 1 into slot 1
    snap_peer.view(torch.bfloat16)[n:2 * n].copy_(vt)
    peer_pad.view(torch.int32)[K.PAD_MISC_WORD] = 1
    p = torch.randn(n, device='cuda').bfloat16()
    g = torch.randn(n, device='cuda').bfloat16()
    p0 = p.float().clone()
    master = p0.clone()
    mom = torch.zeros(n, device='cuda')
    torn = torch.zeros(1, dtype=torch.int32, device='cuda')
    h = _hyper(momentum=0.0, wd=0.0)
    K.call('kfb_pair_avg_step', comm.cstruct, 1, snap_peer.data_ptr(), snap_self.data_ptr(), nbytes,
           p.data_ptr(), K.BF16, g.data_ptr(), K.BF16, master.data_ptr(), mom.data_ptr(), n, 1.0,
           C.byref(h), torn.data_ptr(), 16, 256, K.stream_ptr())
    torch.cuda.synchronize()
    assert torch.allclose(master, 0.5 * (p0 + vt.float()) - 0.1 * g.float(), rtol=1e-5, atol=1e-5)
    assert torch.allclose(p.float(), master, rtol=2e-2, atol=2e-2)
    # own snapshot published as version 1 in slot 1
    own_pad = K.tensor_from_ptr(comm._struct.pads[0], K.PAD_BYTES, 'cuda:0')
    assert int(own_pad.view(torch.int32)[K.PAD_MISC_WORD]) == 1
    assert torch.equal(snap_self.view(torch.bfloat16)[n:2 * n], p)
    assert int(torn) == 0


def test_fuse_defuse_and_norm():
    comm = LocalComm()
    shapes = [(64, 3, 7, 7), (64, ), (128, 64, 1, 1), (1000, 512), (3, ), (17, 5)] * 15  # > 64 tensors
    ts = [torch.randn(s, device='cuda').bfloat16() for s in shapes]
    offs, off = [], 0
    for t in ts:
        offs.append(off)
        off += (t.numel() + 15) // 16 * 16
    flat = torch.zeros(off, dtype=torch.bfloat16, device='cuda')
    ss = torch.zeros(1, device='cuda')
    pa, _1 = K.ptr_array([t.data_ptr() for t in ts])
    ca, _2 = K.i64_array([t.numel() for t in ts])
    oa, _3 = K.i64_array(offs)
    K.call('kfb_fuse', pa, ca, oa, len(ts), K.BF16, flat.data_ptr(), K.BF16, ss.data_ptr(),
           comm._fuse_scratch.data_ptr(), K.stream_ptr())
    torch.cuda.synchronize()
    for t, o in zip(ts, offs):
        assert torch.equal(flat[o:o + t.numel()], t.reshape(-1))
    ref = sum((t.float() ** 2).sum() for t in ts)
    assert torch.allclose(ss[0], ref, rtol=1e-4)
    # cast variants + defuse
    flat32 = torch.zeros(off, dtype=torch.float32, device='cuda')
    K.call('kfb_fuse', pa, ca, oa, len(ts), K.BF16, flat32.data_ptr(), K.F32, None, None, K.stream_ptr())
    outs = [torch.zeros_like(t) for t in ts]
    da, _4 = K.ptr_array([t.data_ptr() for t in outs])
    K.call('kfb_defuse', flat32.data_ptr(), K.F32, da, ca, oa, len(ts), K.BF16, K.stream_ptr())
    torch.cuda.synchronize()
    for t, o in zip(ts, outs):
        assert torch.equal(t, o)


def test_fused_sgd_matches_torch_sgd():
    torch.manual_seed(0)
    def make():
        torch.manual_seed(0)
        return torch.nn.Sequential(torch.nn.Linear(64, 128), torch.nn.ReLU(), torch.nn.Linear(128, 10)).cuda()
    a, b = make(), make()
    oa = torch.optim.SGD(a.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4, nesterov=True)
    ob = FusedSGD(b.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4, nesterov=True, comm=LocalComm())
    x = torch.randn(32, 64, device='cuda')
    y = torch.randint(0, 10, (32, ), device='cuda')
    for _ in range(5):
        for m, o in ((a, oa), (b, ob)):
            o.zero_grad(set_to_none=True)
            torch.nn.functional.cross_entropy(m(x), y).backward()
            o.step()
    for pa, pb in zip(a.parameters(), b.parameters()):
        assert torch.allclose(pa, pb, rtol=1e-5, atol=1e-6), (pa - pb).abs().max()


@pytest.mark.parametrize('np_', [2, 4])
def test_grad_variance_and_gns(np_):
    comm = EmulatedComm(np_)
    n = 8 * 1000
    grads = comm.alloc(n * 4)
    gs = [torch.randn(n, device='cuda') for _ in range(np_)]
    for r in range(np_):
        grads.view(torch.float32, r)[:n].copy_(gs[r])
    ends = torch.tensor([800, 4000, n], dtype=torch.int64, device='cuda')
    seg = torch.zeros(np_ * 3, device='cuda')
    ga, _1 = grads.ptr_array()
    work = torch.zeros(K.VAR_WORK_WORDS * np_, device='cuda')
    K.call('kfb_grad_variance', comm.cstruct, ga, K.F32, n, ends.data_ptr(), 3, seg.data_ptr(), work.data_ptr(),
           comm.default_blocks, 256, K.stream_ptr())
    torch.cuda.synchronize()
    first = seg.clone()
    seg.zero_()
    K.call('kfb_grad_variance', comm.cstruct, ga, K.F32, n, ends.data_ptr(), 3, seg.data_ptr(), work.data_ptr(),
           comm.default_blocks, 256, K.stream_ptr())
    torch.cuda.synchronize()
    assert torch.equal(first, seg), 'the variance sums are formed in a fixed order: bit-identical run to run'
    torch.cuda.synchronize()
    st = torch.stack(gs)
    v = (st * st).mean(0) - st.mean(0) ** 2
    ref = torch.stack([(v[:800] ** 2).sum(), (v[800:4000] ** 2).sum(), (v[4000:] ** 2).sum()])
    got = seg.view(np_, 3).sum(0)
    assert torch.allclose(got, ref, rtol=1e-3)
    # noise-scale finalize vs the reference formulas + EMA
    state = torch.zeros(4, device='cuda')
    exp_g = exp_s = None
    for gs_, gb_ in [(4.0, 1.5), (3.0, 1.0), (5.0, 2.5)]:
        a, b = torch.tensor([gs_], device='cuda'), torch.tensor([gb_], device='cuda')
        K.call('kfb_gns_finalize', a.data_ptr(), b.data_ptr(), 32.0, 128.0, 0.9, state.data_ptr(), K.stream_ptr())
        G = (128.0 * gb_ - 32.0 * gs_) / (128.0 - 32.0)
        S = (gs_ - gb_) / (1 / 32.0 - 1 / 128.0)
        exp_g = G if exp_g is None else 0.9 * exp_g + 0.1 * G
        exp_s = S if exp_s is None else 0.9 * exp_s + 0.1 * S
    torch.cuda.synchronize()
    assert abs(float(state[3]) - exp_s / exp_g) < 1e-3 * abs(exp_s / exp_g)


@pytest.mark.parametrize('np_', [1, 2, 4, 8])
@pytest.mark.parametrize('kind', [1, 2])
def test_ssgd_fused_step_adam_rules(np_, kind):
    """The fused step with the AdamW (kind 1) / Adam (kind 2) update against torch.optim on the averaged
    gradient, three consecutive steps with the step number read from device memory (what keeps a captured
    CUDA graph valid), under single-GPU rank emulation."""
    torch.manual_seed(10 * np_ + kind)
    comm = EmulatedComm(np_) if np_ > 1 else LocalComm()
    n = 8 * 2000
    grads = comm.alloc(n * 4)
    params = comm.alloc(n * 4)
    R = max(np_, 1)
    p0 = torch.randn(n, device='cuda')
    master = p0.repeat(R).contiguous()
    m = torch.zeros_like(master)
    v = torch.zeros_like(master)
    step_t = torch.zeros(1, dtype=torch.int64, device='cuda')
    ref = torch.nn.Parameter(p0.clone())
    cls = torch.optim.AdamW if kind == 1 else torch.optim.Adam
    ropt = cls([ref], lr=1e-2, betas=(0.9, 0.99), eps=1e-8, weight_decay=0.05)
    h = K.SgdHyper()
    h.lr, h.weight_decay, h.kind = 1e-2, 0.05, kind
    h.beta1, h.beta2, h.eps = 0.9, 0.99, 1e-8
    h.step = step_t.data_ptr()
    h.exp_avg_sq = v.data_ptr()
    ga, _1 = grads.ptr_array()
    pa, _2 = params.ptr_array()
    npk = n // 8
    per = (npk + R - 1) // R * 8
    for it in range(3):
        gs = [torch.randn(n, device='cuda') for _ in range(R)]
        for r in range(R):
            grads.view(torch.float32, r if np_ > 1 else None)[:n].copy_(gs[r])
        K.call('kfb_ssgd_fused_step', comm.cstruct, ga, K.F32, pa, K.F32, master.data_ptr(), m.data_ptr(), n,
               1.0 / R, C.byref(h), None, comm.default_blocks, 256, K.stream_ptr())
        step_t.add_(1)
        torch.cuda.synchronize()
        ref.grad = torch.stack(gs).sum(0) / R
        ropt.step()
        for r in range(R):
            got = params.view(torch.float32, r if np_ > 1 else None)[:n]
            assert torch.allclose(got, ref.data, rtol=1e-5, atol=1e-6), (it, r, (got - ref.data).abs().max())
    st = ropt.state[ref]
    for r in range(R):
        lo, hi = min(r * per, n), min((r + 1) * per, n)
        assert torch.allclose(m[r * n + lo:r * n + hi], st['exp_avg'][lo:hi], rtol=1e-5, atol=1e-7)
        assert torch.allclose(v[r * n + lo:r * n + hi], st['exp_avg_sq'][lo:hi], rtol=1e-5, atol=1e-9)


def test_fused_adamw_optimizer_single_gpu_with_overlap_and_checkpoint():
    """FusedAdamW through the public wrapper on one GPU: bucketed launches from autograd hooks (local update, no
    exchange), equality with torch.optim.AdamW, and a state_dict round trip into fresh objects."""
    from kungfu_b200.optim import FusedAdamW
    import kungfu_b200.torch as kf

    def make():
        torch.manual_seed(3)
        return torch.nn.Sequential(torch.nn.Linear(64, 256), torch.nn.ReLU(), torch.nn.Linear(256, 16)).cuda()

    def batch(step):
        g = torch.Generator().manual_seed(100 + step)
        return torch.randn(32, 64, generator=g).cuda(), torch.randint(0, 16, (32, ), generator=g).cuda()

    m, ref = make(), make()
    opt = kf.optimizers.SynchronousSGDOptimizer(FusedAdamW(m.parameters(), lr=1e-2, weight_decay=0.05, bucket_bytes=8000),
                                                m.named_parameters(), overlap=True)
    ropt = torch.optim.AdamW(ref.parameters(), lr=1e-2, weight_decay=0.05)
    for step in range(5):
        if step == 3:
            sd = {'model': m.state_dict(), 'opt': opt.state_dict()}
            m = make()
            for p in m.parameters():
                p.data.zero_()
            opt = kf.optimizers.SynchronousSGDOptimizer(FusedAdamW(m.parameters(), lr=1.0, bucket_bytes=8000),
                                                        m.named_parameters(), overlap=True)
            m.load_state_dict(sd['model'])
            opt.load_state_dict(sd['opt'])
        x, y = batch(step)
        opt.zero_grad(set_to_none=True)
        torch.nn.functional.cross_entropy(m(x), y).backward()
        opt.step()
        ropt.zero_grad()
        torch.nn.functional.cross_entropy(ref(x), y).backward()
        ropt.step()
    torch.cuda.synchronize()
    assert opt._ov is not None and len(opt._ov['buckets']) > 1
    for p, q in zip(m.parameters(), ref.parameters()):
        assert torch.allclose(p, q, rtol=1e-4, atol=1e-5), (p - q).abs().max()
