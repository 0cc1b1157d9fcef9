"""GPU numerics of the peer-memory collectives, checked against a plain fp32
PyTorch reference of the same op. Ranks are emulated on one GPU: each
collective is ONE cooperative launch whose blockIdx.y plays the rank (the
multi-GPU run of the same kernels is exercised by benchmarks/allreduce_bench.py
and tests marked gpu + multi-GPU)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from kungfu_b200.ops import _cuda as K
from kungfu_b200.ops import collective as coll
from kungfu_b200.parallel.comm import EmulatedComm, LocalComm

DT = {torch.float32: 1e-6, torch.bfloat16: 1e-2, torch.float16: 2e-3}


def _fill(comm, buf, dtype, n, gen):
    xs = []
    for r in range(comm.np):
        x = (torch.randn(n, generator=gen, device='cuda') * 2).to(dtype)
        buf.view(dtype, r)[:n].copy_(x)
        xs.append(x)
    return xs


@pytest.mark.parametrize('np_', [2, 3, 4, 8])
@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize('algo', [K.ALGO_ONESHOT, K.ALGO_TWOSHOT])
def test_allreduce_sum_scaled(np_, dtype, algo):
    gen = torch.Generator(device='cuda').manual_seed(1234 + np_)
    comm = EmulatedComm(np_)
    n = 8 * 1024 * 3 + 40  # not a multiple of the shard or the CTA tile
    esize = torch.empty(0, dtype=dtype).element_size()
    buf = comm.alloc(n * esize + 64)
    out = comm.alloc(n * esize + 64)
    xs = _fill(comm, buf, dtype, n, gen)
    ss = torch.zeros(np_, device='cuda')
    coll.all_reduce(comm, buf, n, dtype, 'sum', scale=1.0 / np_, algo=algo, out=out, sumsq_out=ss)
    torch.cuda.synchronize()
    ref = torch.stack([x.float() for x in xs]).sum(0) / np_
    for r in range(np_):
        got = out.view(dtype, r)[:n].float()
        if algo == K.ALGO_ONESHOT and False:
            continue
        assert torch.allclose(got, ref, rtol=DT[dtype], atol=DT[dtype]), (r, (got - ref).abs().max())
    # every rank must produce identical bits
    for r in range(1, np_):
        assert torch.equal(out.view(dtype, r)[:n], out.view(dtype, 0)[:n])
    # ||out||^2 (fp32 pre-rounding values), identical on every rank
    assert torch.allclose(ss, (ref * ref).sum().expand(np_), rtol=2e-3)
    assert (ss == ss[0]).all()


@pytest.mark.parametrize('op', ['min', 'max', 'prod', 'sum'])
@pytest.mark.parametrize('dtype', [torch.int32, torch.float32, torch.int64, torch.float64, torch.uint8])
def test_allreduce_ops_inplace_twoshot(op, dtype):
    if dtype == torch.uint8 and op == 'prod':
        pytest.skip('uint8 product overflows by design')
    comm = EmulatedComm(4)
    n = 5000
    esize = torch.empty(0, dtype=dtype).element_size()
    buf = comm.alloc(n * esize + 64)
    xs = []
    for r in range(4):
        x = torch.randint(1, 4, (n, ), device='cuda').to(dtype)
        buf.view(dtype, r)[:n].copy_(x)
        xs.append(x)
    coll.all_reduce(comm, buf, n, dtype, op, algo=K.ALGO_TWOSHOT)
    torch.cuda.synchronize()
    st = torch.stack([x.double() for x in xs])
    ref = {'min': st.min(0).values, 'max': st.max(0).values, 'prod': st.prod(0), 'sum': st.sum(0)}[op]
    for r in range(4):
        assert torch.equal(buf.view(dtype, r)[:n].double(), ref)


def test_allreduce_repeated_launches_and_epochs():
    """Flags are epoch counters that are never cleared: many back-to-back launches must work."""
    comm = EmulatedComm(4)
    n = 4096
    buf = comm.alloc(n * 4)
    for r in range(4):
        buf.view(torch.float32, r)[:n].fill_(1.0)
    for it in range(50):
        coll.all_reduce(comm, buf, n, torch.float32, 'sum', scale=0.25, algo=K.ALGO_TWOSHOT)
    torch.cuda.synchronize()
    for r in range(4):
        assert torch.equal(buf.view(torch.float32, r)[:n], torch.ones(n, device='cuda'))


@pytest.mark.parametrize('np_', [2, 5, 8])
def test_broadcast_allgather_gather_reduce(np_):
    comm = EmulatedComm(np_)
    n = 3000
    buf = comm.alloc(n * 4)
    for r in range(np_):
        buf.view(torch.float32, r)[:n].copy_(torch.arange(n, device='cuda') + 1000.0 * r)
    root = np_ - 1
    coll.broadcast(comm, buf, n * 4, root=root)
    torch.cuda.synchronize()
    for r in range(np_):
        assert torch.equal(buf.view(torch.float32, r)[:n], torch.arange(n, device='cuda') + 1000.0 * root)
    # all-gather
    slot = 1024
    src = comm.alloc(slot)
    dst = comm.alloc(slot * np_)
    for r in range(np_):
        src.view(torch.uint8, r)[:slot].fill_(r + 1)
    coll.all_gather(comm, src, dst, slot)
    torch.cuda.synchronize()
    for r in range(np_):
        got = dst.view(torch.uint8, r)[:slot * np_].view(np_, slot)
        for k in range(np_):
            assert (got[k] == k + 1).all()
    # gather / reduce to root 0
    out = torch.zeros(slot * np_, dtype=torch.uint8, device='cuda')
    coll.gather(comm, src, out, slot, root=0)
    torch.cuda.synchronize()
    assert torch.equal(out.view(np_, slot)[:, 0].cpu(), torch.arange(1, np_ + 1, dtype=torch.uint8))
    red = torch.zeros(n, device='cuda')
    for r in range(np_):
        buf.view(torch.float32, r)[:n].fill_(float(r))
    coll.reduce(comm, buf, red, n, torch.float32, 'sum', root=0)
    torch.cuda.synchronize()
    assert torch.equal(red, torch.full((n, ), float(sum(range(np_))), device='cuda'))


def test_consensus_and_barrier():
    comm = EmulatedComm(4)
    buf = comm.alloc(4096)
    agree = torch.zeros(4, dtype=torch.int32, device='cuda')
    coll.consensus(comm, buf, 4096, agree)
    comm.barrier()
    torch.cuda.synchronize()
    assert agree.tolist() == [1, 1, 1, 1]
    buf.view(torch.uint8, 2)[100] = 7
    coll.consensus(comm, buf, 4096, agree)
    torch.cuda.synchronize()
    assert agree.tolist() == [0, 0, 0, 0]


def test_edge_kernels():
    a = torch.randn(10000, device='cuda')
    b = torch.randn(10000, device='cuda')
    ref = a + b
    coll.edge_reduce(a.data_ptr(), b.data_ptr(), a.numel(), torch.float32, 'sum')
    c = torch.empty_like(a)
    coll.edge_copy(c.data_ptr(), a.data_ptr(), a.numel() * 4)
    torch.cuda.synchronize()
    assert torch.equal(a, ref) and torch.equal(c, ref)


def test_single_rank_comm():
    comm = LocalComm()
    buf = comm.alloc(4096)
    buf.view(torch.float32)[:1024].fill_(3.0)
    out = comm.alloc(4096)
    coll.all_reduce(comm, buf, 1024, torch.float32, 'sum', scale=0.5, algo=K.ALGO_ONESHOT, out=out)
    torch.cuda.synchronize()
    assert (out.view(torch.float32)[:1024] == 1.5).all()
