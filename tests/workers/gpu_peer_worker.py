"""Multi-GPU checks of the peer-memory path: one process per GPU, launched by
kungfu-run -allow-nvlink. Every result is compared with a local fp32 oracle
built from the other ranks' (deterministically seeded) inputs."""
import ctypes as C
import sys

import torch

import kungfu_b200.python as kfpy
import kungfu_b200.torch as kf
from kungfu_b200.ops import _cuda as K
from kungfu_b200.ops import collective as coll
from kungfu_b200.parallel import comm as pc

rank, np_ = kf.current_rank(), kf.current_cluster_size()
torch.cuda.set_device(kf.get_cuda_index())
dev = torch.device('cuda', torch.cuda.current_device())
comm = pc.default_comm()
assert comm.np == np_ and comm.rank == rank


def inputs(n, dtype, tag):
    out = []
    for r in range(np_):
        g = torch.Generator().manual_seed(1000 * tag + r)
        out.append((torch.randn(n, generator=g) * 2).to(dtype))
    return out


def check_allreduce():
    algos = [K.ALGO_ONESHOT, K.ALGO_TWOSHOT]
    n = 1 << 20
    for dtype, tol in ((torch.float32, 1e-5), (torch.bfloat16, 2e-2)):
        es = torch.empty(0, dtype=dtype).element_size()
        buf = comm.alloc(n * es, multicast=True)
        out = comm.alloc(n * es, multicast=True)
        if pc.require_nvls() and np_ > 1:
            assert buf.mc_ptr and out.mc_ptr, 'KUNGFU_REQUIRE_NVLS=1: the buffers must have a multicast alias'
        these = list(algos) + ([K.ALGO_NVLS] if buf.mc_ptr else [])
        xs = inputs(n, dtype, 1)
        ref = torch.stack([x.float() for x in xs]).sum(0) / np_
        for algo in these:
            buf.view(dtype)[:n].copy_(xs[rank])
            ss = torch.zeros(1, device=dev)
            coll.all_reduce(comm, buf, n, dtype, 'sum', scale=1.0 / np_, algo=algo, out=out, sumsq_out=ss)
            torch.cuda.synchronize()
            got = out.view(dtype)[:n].float().cpu()
            assert torch.allclose(got, ref, rtol=tol, atol=tol), (algo, dtype, (got - ref).abs().max())
            assert abs(float(ss) - float((ref * ref).sum())) < 2e-2 * float((ref * ref).sum()), (algo, float(ss))
            # every rank holds identical bits / norms
            sums = kf.ops.all_gather(torch.tensor([float(ss), float(got.double().sum())], dtype=torch.float64))
            assert (sums == sums[0]).all(), (algo, sums)
        if rank == 0:
            print('allreduce ok', dtype, 'algos', these, flush=True)
    # in place two-shot, repeated (epoch flags), min/max
    buf = comm.alloc(4096 * 4)
    for it in range(20):
        buf.view(torch.float32)[:4096].fill_(float(rank + it))
        coll.all_reduce(comm, buf, 4096, torch.float32, 'max', algo=K.ALGO_TWOSHOT)
        torch.cuda.synchronize()
        assert (buf.view(torch.float32)[:4096] == np_ - 1 + it).all()


def check_movers():
    n = 100000
    buf = comm.alloc(n * 4, multicast=True)
    for use_mc in (False, True):
        buf.view(torch.float32)[:n].fill_(float(rank + 1))
        coll.broadcast(comm, buf, n * 4, root=np_ - 1, use_multicast=use_mc)
        torch.cuda.synchronize()
        assert (buf.view(torch.float32)[:n] == np_).all(), use_mc
    src, dst = comm.alloc(1024), comm.alloc(1024 * np_)
    src.local[:1024].fill_(rank + 1)
    coll.all_gather(comm, src, dst, 1024)
    torch.cuda.synchronize()
    got = dst.local[:1024 * np_].view(np_, 1024)
    for r in range(np_):
        assert (got[r] == r + 1).all()
    agree = torch.zeros(1, dtype=torch.int32, device=dev)
    coll.consensus(comm, src, 1024, agree)
    torch.cuda.synchronize()
    assert int(agree) == (1 if np_ == 1 else 0)
    # tensor API on CUDA tensors
    x = torch.full((1000, ), float(rank), device=dev)
    assert (kf.ops.all_reduce_fn(x) == sum(range(np_))).all()
    assert kf.ops.all_gather(x).shape == (np_, 1000)
    h = kf.ops.inplace_all_reduce_async_op(x, 'x', 'max')
    kf.ops.wait_handle(h)
    torch.cuda.synchronize()
    assert (x == np_ - 1).all()
    sd = {'a': torch.full((33, ), float(rank), device=dev)}
    kf.broadcast_parameters(sd)
    torch.cuda.synchronize()
    assert (sd['a'] == 0).all()


def check_fused_optimizers():
    from kungfu_b200.optim import FusedAdamW, FusedSGD
    torch.manual_seed(0)

    def make():
        torch.manual_seed(0)
        return torch.nn.Sequential(torch.nn.Linear(64, 256), torch.nn.ReLU(), torch.nn.Linear(256, 16)).to(dev)

    def batch(r, step):
        g = torch.Generator().manual_seed(77 + 13 * r + step)
        return torch.randn(32, 64, generator=g).to(dev), torch.randint(0, 16, (32, ), generator=g).to(dev)

    # S-SGD fused == single process on the concatenated batch (P2P kernel, then the NVLS kernel)
    def fused_ssgd_matches_oracle(nvls_min, overlap=False, adam=False):
        m = make()
        if adam:
            inner = FusedAdamW(m.parameters(), lr=1e-2, weight_decay=0.05, bucket_bytes=8000)
        else:
            inner = FusedSGD(m.parameters(), lr=0.1, momentum=0.9, bucket_bytes=8000)
        opt = kf.optimizers.SynchronousSGDOptimizer(inner, m.named_parameters(), overlap=overlap)
        opt.nvls_min_bytes = nvls_min
        ref = make()
        if adam:
            ropt = torch.optim.AdamW(ref.parameters(), lr=1e-2, weight_decay=0.05)
        else:
            ropt = torch.optim.SGD(ref.parameters(), lr=0.1, momentum=0.9)
        for step in range(4):
            x, y = batch(rank, step)
            opt.zero_grad(set_to_none=True)
            torch.nn.functional.cross_entropy(m(x), y).backward()
            opt.step()
            ropt.zero_grad()
            xs, ys = zip(*[batch(r, step) for r in range(np_)])
            torch.nn.functional.cross_entropy(ref(torch.cat(xs)), torch.cat(ys)).backward()
            ropt.step()
        torch.cuda.synchronize()
        if overlap and np_ > 1:
            assert len(opt._ov['buckets']) > 1, 'the test model must span several buckets'
        want = 'nvls' if (nvls_min == 0 and comm.multicast_ok) else ('p2p' if np_ > 1 else 'local')
        assert opt.last_algo == want, (opt.last_algo, want)
        if nvls_min == 0 and pc.require_nvls() and np_ > 1:
            assert opt.last_algo == 'nvls'
        tol = dict(rtol=2e-3, atol=2e-4) if adam else dict(rtol=1e-4, atol=1e-5)
        for p, q in zip(m.parameters(), ref.parameters()):
            assert torch.allclose(p, q, **tol), (nvls_min, overlap, adam, (p - q).abs().max())
        # the sharded fp32 state, made whole, equals the single-process optimizer's
        sd = opt.state_dict()
        for i, q in enumerate(ref.parameters()):
            assert torch.allclose(sd['state'][i]['master'], q, **tol), (i, 'master')
            rs = ropt.state[q]
            if adam:
                assert torch.allclose(sd['state'][i]['exp_avg'], rs['exp_avg'], rtol=2e-3, atol=1e-5)
                assert torch.allclose(sd['state'][i]['exp_avg_sq'], rs['exp_avg_sq'], rtol=2e-3, atol=1e-7)
            else:
                assert torch.allclose(sd['state'][i]['momentum_buffer'], rs['momentum_buffer'], rtol=1e-4, atol=1e-5)
        return [p.detach().clone() for p in m.parameters()]

    for adam in (False, True):
        a = fused_ssgd_matches_oracle(1 << 40, adam=adam)  # peer load/store kernel
        b = fused_ssgd_matches_oracle(0, adam=adam)        # NVLS kernel (when multicast is available)
        c = fused_ssgd_matches_oracle(1 << 40, overlap=True, adam=adam)  # buckets launched from autograd hooks
        d = fused_ssgd_matches_oracle(0, overlap=True, adam=adam)
        # cutting the arena into buckets changes no element's arithmetic
        for x, y in zip(a, c):
            assert torch.equal(x, y), 'bucketed p2p step differs from the single launch'
        for x, y in zip(b, d):
            assert torch.equal(x, y), 'bucketed NVLS step differs from the single launch'
    if rank == 0:
        print('fused S-SGD / AdamW ok (p2p + nvls, single launch + overlapped buckets)', flush=True)

    # save -> new objects -> load -> continue == uninterrupted run
    def run_ckpt(split):
        m = make()
        opt = kf.optimizers.SynchronousSGDOptimizer(FusedSGD(m.parameters(), lr=0.1, momentum=0.9), m.named_parameters())
        for step in range(6):
            if step == split:
                import io
                f = io.BytesIO()
                torch.save({'model': m.state_dict(), 'opt': opt.state_dict()}, f)
                f.seek(0)
                ck = torch.load(f)
                m = make()
                for p in m.parameters():
                    p.data.zero_()
                opt = kf.optimizers.SynchronousSGDOptimizer(FusedSGD(m.parameters(), lr=0.5, momentum=0.1), m.named_parameters())
                m.load_state_dict(ck['model'])
                opt.load_state_dict(ck['opt'])
            x, y = batch(rank, step)
            opt.zero_grad(set_to_none=True)
            torch.nn.functional.cross_entropy(m(x), y).backward()
            opt.step()
        torch.cuda.synchronize()
        return [p.detach().clone() for p in m.parameters()]

    for x, y in zip(run_ckpt(-1), run_ckpt(3)):
        assert torch.allclose(x, y, rtol=1e-6, atol=1e-7), (x - y).abs().max()
    if rank == 0:
        print('checkpoint round trip ok', flush=True)

    # generic optimizer (Adam) through GradSync, plain and overlapped
    for overlap in (False, True):
        m = make()
        opt = kf.optimizers.SynchronousSGDOptimizer(torch.optim.Adam(m.parameters(), lr=1e-2), m.named_parameters(), overlap=overlap)
        ref = make()
        ropt = torch.optim.Adam(ref.parameters(), lr=1e-2)
        for step in range(3):
            x, y = batch(rank, step)
            opt.zero_grad(set_to_none=True)
            torch.nn.functional.cross_entropy(m(x), y).backward()
            opt.step()
            ropt.zero_grad()
            xs, ys = zip(*[batch(r, step) for r in range(np_)])
            torch.nn.functional.cross_entropy(ref(torch.cat(xs)), torch.cat(ys)).backward()
            ropt.step()
        torch.cuda.synchronize()
        for p, q in zip(m.parameters(), ref.parameters()):
            assert torch.allclose(p, q, rtol=5e-3, atol=1e-3), (overlap, (p - q).abs().max())  # Adam amplifies rounding
    if rank == 0:
        print('GradSync ok', flush=True)

    # SMA fused: replicas stay finite and get closer
    m = make()
    for p in m.parameters():
        p.data += 0.1 * rank
    opt = kf.optimizers.SynchronousAveragingOptimizer(FusedSGD(m.parameters(), lr=0.05), m.named_parameters(), alpha=0.5)
    w0 = kf.ops.all_gather(next(m.parameters()).data.clone())
    for step in range(5):
        x, y = batch(rank, step)
        opt.zero_grad(set_to_none=True)
        torch.nn.functional.cross_entropy(m(x), y).backward()
        opt.step()
    w1 = kf.ops.all_gather(next(m.parameters()).data.clone())
    torch.cuda.synchronize()
    if np_ > 1:
        assert float((w1[0] - w1[-1]).abs().max()) < 0.2 * float((w0[0] - w0[-1]).abs().max())
    # SMA: fused (average on the switch, started by zero_grad under forward/backward), fused AdamW and a
    # plain torch optimizer on the device path all follow the same recurrence as a host-side oracle
    def sma_oracle(alpha, lr, steps, adam):
        ms = [make() for _ in range(np_)]
        for r, mm in enumerate(ms):
            for p in mm.parameters():
                p.data += 0.1 * r
        mk = (lambda ps: torch.optim.AdamW(ps, lr=lr, weight_decay=0.0)) if adam else (lambda ps: torch.optim.SGD(ps, lr=lr))
        opts = [mk(mm.parameters()) for mm in ms]
        for step in range(steps):
            for r, (mm, oo) in enumerate(zip(ms, opts)):
                x, y = batch(r, step)
                oo.zero_grad()
                torch.nn.functional.cross_entropy(mm(x), y).backward()
            with torch.no_grad():
                for ps in zip(*[list(mm.parameters()) for mm in ms]):
                    avg = torch.stack([p.data for p in ps]).mean(0)
                    for p in ps:
                        p.data.add_(avg - p.data, alpha=alpha)
            for oo in opts:
                oo.step()
        return [p.detach().clone() for p in ms[rank].parameters()]

    for kind in ('fused-sgd', 'fused-adamw', 'torch-sgd', 'torch-adamw'):
        adam = kind.endswith('adamw')
        m = make()
        for p in m.parameters():
            p.data += 0.1 * rank
        if kind == 'fused-sgd':
            inner = FusedSGD(m.parameters(), lr=0.05)
        elif kind == 'fused-adamw':
            inner = FusedAdamW(m.parameters(), lr=0.01, weight_decay=0.0)
        elif kind == 'torch-sgd':
            inner = torch.optim.SGD(m.parameters(), lr=0.05)
        else:
            inner = torch.optim.AdamW(m.parameters(), lr=0.01, weight_decay=0.0)
        opt = kf.optimizers.SynchronousAveragingOptimizer(inner, m.named_parameters(), alpha=0.5)
        for step in range(4):
            x, y = batch(rank, step)
            opt.zero_grad(set_to_none=True)
            torch.nn.functional.cross_entropy(m(x), y).backward()
            opt.step()
        torch.cuda.synchronize()
        want = sma_oracle(0.5, 0.01 if adam else 0.05, 4, adam)
        for p, q in zip(m.parameters(), want):
            assert torch.allclose(p, q, rtol=2e-3, atol=2e-4), (kind, (p - q).abs().max())
    if rank == 0:
        print('SMA ok: fused SGD / fused AdamW / torch SGD / torch AdamW vs oracle', flush=True)
    # PairAveraging around a plain torch optimizer stays on the device (pull + blend kernel, publish kernel);
    # lr = 0 makes it pure gossip: replicas contract
    if np_ > 1:
        for mixed in (False, True):
            m = make()
            if mixed:
                m[2] = m[2].to(torch.bfloat16)  # two dtypes -> two arenas, one snapshot
            for p in m.parameters():
                p.data += 0.1 * rank
            opt = kf.optimizers.PairAveragingOptimizer(torch.optim.SGD(m.parameters(), lr=0.0), m.named_parameters())
            first = next(m.parameters())
            w0 = kf.ops.all_gather(first.data.clone())
            for step in range(12):
                opt.step()
            torch.cuda.synchronize()
            kf.run_barrier()
            assert opt._kf_gen is not None, 'the device path must have been taken'
            w1 = kf.ops.all_gather(first.data.clone())
            spread0 = float((w0.max(0).values - w0.min(0).values).max())
            spread1 = float((w1.max(0).values - w1.min(0).values).max())
            assert spread1 < 0.5 * spread0, (mixed, spread0, spread1)
        if rank == 0:
            print('PairAveraging (torch optimizer, device path) ok', flush=True)
    # PairAveraging fused: one-sided pulls, no barrier
    m = make()
    opt = kf.optimizers.PairAveragingOptimizer(FusedSGD(m.parameters(), lr=0.05), m.named_parameters())
    for step in range(6):
        x, y = batch(rank, step)
        opt.zero_grad(set_to_none=True)
        torch.nn.functional.cross_entropy(m(x), y).backward()
        opt.step()
    torch.cuda.synchronize()
    assert all(torch.isfinite(p).all() for p in m.parameters())
    kf.run_barrier()
    torn = kf.ops.all_reduce_fn(opt.torn_reads.to(torch.int64).cpu(), 'sum')
    if rank == 0:
        print('PairAveraging fused ok; torn reads (all ranks)', int(torn), flush=True)
    # PairAveraging with the peer drawn ON THE DEVICE: the whole step (forward, backward, pull + blend +
    # update + publish) is one CUDA graph; lr = 0 makes it pure gossip averaging, so replicas must contract
    if np_ > 1:
        m = make()
        for p in m.parameters():
            p.data += 0.1 * rank
        opt = kf.optimizers.PairAveragingOptimizer(FusedSGD(m.parameters(), lr=0.0), m.named_parameters())
        opt.device_select = True
        opt.arenas()
        first = next(m.parameters())
        w0 = kf.ops.all_gather(first.data.clone())
        x, y = batch(rank, 0)

        def pair_step():
            opt.zero_grad(set_to_none=True)
            torch.nn.functional.cross_entropy(m(x), y).backward()
            opt.step()

        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                pair_step()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        kf.run_barrier()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            pair_step()
        for _ in range(8):
            graph.replay()
        torch.cuda.synchronize()
        kf.run_barrier()
        assert int(opt._step_t) == 3 + 8, int(opt._step_t)
        w1 = kf.ops.all_gather(first.data.clone())
        spread0 = float((w0.max(0).values - w0.min(0).values).max())
        spread1 = float((w1.max(0).values - w1.min(0).values).max())
        assert all(torch.isfinite(p).all() for p in m.parameters())
        assert spread1 < 0.25 * spread0, (spread0, spread1)
        if rank == 0:
            print('PairAveraging device-selected peer under a CUDA graph ok: spread %.4f -> %.4f' % (spread0, spread1),
                  flush=True)
    # monitors
    m = make()
    opt = kf.optimizers.MonitorGradientNoiseScaleOptimizer(FusedSGD(m.parameters(), lr=0.05), m.named_parameters(), device_batch_size=32)
    for step in range(3):
        x, y = batch(rank, step)
        opt.zero_grad(set_to_none=True)
        torch.nn.functional.cross_entropy(m(x), y).backward()
        opt.step()
    if np_ > 1:
        assert opt.noise_scale is not None and torch.isfinite(opt.noise_scale)
        plain = float(opt.noise_scale)
        # the same monitor on the overlapped (bucketed) step: norms come from per-bucket slots
        for inner_kind in ('fused', 'torch'):
            m = make()
            if inner_kind == 'fused':
                inner = FusedSGD(m.parameters(), lr=0.05, bucket_bytes=8000)
            else:
                inner = torch.optim.SGD(m.parameters(), lr=0.05)
            opt = kf.optimizers.MonitorGradientNoiseScaleOptimizer(inner, m.named_parameters(), device_batch_size=32,
                                                                   overlap=True)
            if inner_kind == 'torch':
                opt._kf_sync.bucket_bytes = 8000
            for step in range(3):
                x, y = batch(rank, step)
                opt.zero_grad(set_to_none=True)
                torch.nn.functional.cross_entropy(m(x), y).backward()
                opt.step()
            got = float(opt.noise_scale)
            assert abs(got - plain) <= 1e-4 * abs(plain), (inner_kind, got, plain)
    m = make()
    opt = kf.optimizers.MonitorGradientVarianceOptimizer(torch.optim.SGD(m.parameters(), lr=0.05), m.named_parameters())
    x, y = batch(rank, 0)
    torch.nn.functional.cross_entropy(m(x), y).backward()
    gs = kf.ops.all_gather(torch.cat([p.grad.reshape(-1) for p in m.parameters()]))
    opt.step()
    exp, off = 0.0, 0
    for p in m.parameters():
        st = gs[:, off:off + p.numel()]
        exp += float(torch.norm((st * st).mean(0) - st.mean(0) ** 2))
        off += p.numel()
    assert abs(float(opt.variance) - exp) < 1e-3 * max(exp, 1e-6), (float(opt.variance), exp)
    if rank == 0:
        print('monitors ok', flush=True)


def check_planner():
    """Strategy graphs executed on the device + online variant chooser."""
    from kungfu_b200.parallel.planner import GraphExecutor, VariantChooser
    n = 300000
    buf = comm.alloc(n * 4)
    xs = inputs(n, torch.float32, 9)
    ref = torch.stack(xs).sum(0)
    for s in ('STAR', 'RING', 'BINARY_TREE', 'CLIQUE', 'TREE'):
        kfpy.set_strategy(s)
        ex = GraphExecutor(comm)
        buf.view(torch.float32)[:n].copy_(xs[rank])
        ex.all_reduce(buf, n, torch.float32, 'sum', monitored=True)
        torch.cuda.synchronize()
        got = buf.view(torch.float32)[:n].cpu()
        assert torch.allclose(got, ref, rtol=1e-5, atol=1e-5), (s, (got - ref).abs().max())
        assert ex.throughput() > 0
    # set_tree: a chain rooted at the last rank
    father = [min(r + 1, np_ - 1) for r in range(np_)]
    kfpy.set_tree(father)
    ex = GraphExecutor(comm)
    buf.view(torch.float32)[:n].copy_(xs[rank])
    ex.all_reduce(buf, n, torch.float32, 'max')
    torch.cuda.synchronize()
    assert torch.equal(buf.view(torch.float32)[:n].cpu(), torch.stack(xs).max(0).values)
    kfpy.set_strategy('AUTO')
    # ---- the same through the PUBLIC ops on CUDA tensors (nothing is staged through the host) ----
    x = torch.arange(100003, device=dev, dtype=torch.float32) * 1e-3 + rank
    want = sum(torch.arange(100003, dtype=torch.float32) * 1e-3 + r for r in range(np_))
    for s in ('STAR', 'RING', 'BINARY_TREE_STAR'):
        kfpy.set_strategy(s)
        y = kf.ops.monitored_all_reduce(x)
        assert y.is_cuda and torch.allclose(y.cpu(), want, rtol=1e-5, atol=1e-4), (s, (y.cpu() - want).abs().max())
    star = [np_ - 1] * np_
    y = kf.ops.monitored_all_reduce(x, tree=star, op='max')
    assert torch.equal(y.cpu(), torch.arange(100003, dtype=torch.float32) * 1e-3 + (np_ - 1))
    kfpy.set_strategy('AUTO')
    assert kf.ops.consensus(torch.full((1000, ), 7.0, device=dev))
    assert kf.ops.consensus(torch.full((1000, ), float(rank), device=dev)) == (np_ == 1)
    # ---- device traffic in the monitoring API, and the interference vote on the TRAINING all-reduce ----
    from kungfu_b200.parallel.planner import planner_for
    dp = planner_for(comm)
    big = comm.alloc(16 << 20, multicast=True)
    kf.ops.calc_stats()          # opens the first window (and switches device timing on)
    kf.ops.egress_rates()

    def window(k=12):
        for _ in range(k):
            coll.all_reduce(comm, big, (16 << 20) // 2, torch.bfloat16, 'sum', scale=1.0 / np_)
        torch.cuda.synchronize()
        kf.ops.calc_stats()

    if np_ > 1:
        window()
        rates = kf.ops.egress_rates()
        assert len(rates) == np_ and float(rates.sum()) > 1e6 and float(rates[rank]) == 0.0, rates
        active = comm.stats.active
        fast = comm.stats.by_variant[active].throughput
        assert fast > 0 and not kf.ops.check_interference()
        # slow the active variant down (one CTA instead of 32; a collective kernel needs the same grid on
        # every rank, so the injection is job-wide): the window collapses, the workers vote, the planner
        # installs another variant
        coll.SLOW_VARIANT = active
        window()
        slow = comm.stats.by_variant[active].throughput
        assert slow < 0.8 * fast, (slow, fast)
        assert kf.ops.check_interference(), 'a majority saw the active variant under 80 % of its reference'
        coll.SLOW_VARIANT = None
        window()
        if comm.multicast_ok:  # (without multicast two-shot is the only large-message variant: nothing to switch to)
            assert coll.choose_algo(comm, 16 << 20, big, K.OP_SUM, K.BF16) != active
            assert comm.stats.active != active
        xs16 = inputs((16 << 20) // 2, torch.bfloat16, 5)
        big.view(torch.bfloat16)[:(16 << 20) // 2].copy_(xs16[rank])
        coll.all_reduce(comm, big, (16 << 20) // 2, torch.bfloat16, 'sum')
        torch.cuda.synchronize()
        ref16 = torch.stack([t.float() for t in xs16]).sum(0)
        got16 = big.view(torch.bfloat16)[:(16 << 20) // 2].float().cpu()
        assert torch.allclose(got16, ref16, rtol=2e-2, atol=2e-2)
        if rank == 0:
            from kungfu_b200.parallel.devstats import VARIANTS
            print('adaptive switch ok: %s %.1f GB/s -> slowed to %.1f GB/s -> %s' %
                  (VARIANTS[active], fast / 1e9, slow / 1e9, VARIANTS[comm.stats.active]), flush=True)
            kf.ops.print_strategy_stats()
        dp.chooser.disabled.clear()
        dp.probe()
    ch = VariantChooser(comm)
    ch.probe(sizes=(64 << 10, 4 << 20))
    assert ch.choose(4 << 20) in (K.ALGO_TWOSHOT, K.ALGO_NVLS, K.ALGO_ONESHOT)
    switched = ch.check_interference()
    assert switched in (False, True)
    if rank == 0:
        print('planner ok: best', ch.best, 'busbw', {a: round(v) for a, v in ch.reference.items()}, flush=True)


check_allreduce()
check_movers()
check_fused_optimizers()
check_planner()
kf.run_barrier()
print('OK rank %d/%d multicast=%s' % (rank, np_, getattr(comm, 'multicast_ok', None)), flush=True)
