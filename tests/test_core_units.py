"""Unit tests of the C++ core through its C ABI (no processes, no GPU).
Mirrors the reference's Go/C++ unit tests: /root/reference/srcs/go/plan/*_test.go,
/root/reference/tests/cpp/unit/test_{dtype,mst,op}.cpp."""
import ctypes as C
import json
import os

import numpy as np
import pytest

from kungfu_b200 import _host as H


def _s(fn, *args, cap=1 << 16):
    buf = C.create_string_buffer(cap)
    n = fn(*args, buf, cap)
    assert n >= 0, H.lib().kf_last_error().decode()
    return buf.value.decode()


def _lib():
    l = H.lib()
    l.kf_gen_strategy_graphs.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int]
    l.kf_minimum_spanning_tree.argtypes = [C.POINTER(C.c_float), C.c_int, C.POINTER(C.c_int32)]
    l.kf_transform2.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int]
    l.kf_transform2.restype = None
    l.kf_cluster_resize.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_int]
    l.kf_parse_hostfile.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
    l.kf_gen_peer_list.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_char_p, C.c_int]
    return l


PEERS_1HOST = b'127.0.0.1:10000,127.0.0.1:10001,127.0.0.1:10002,127.0.0.1:10003'
PEERS_2HOST = b'10.0.0.1:10000,10.0.0.1:10001,10.0.0.2:10000,10.0.0.2:10001,10.0.0.3:10000'


def _edges(s):
    import re
    return set((int(a), int(b)) for a, b in re.findall(r'\((\d+)->(\d+)\)', s))


def _loops(s):
    import re
    return set(int(a) for a in re.findall(r'\((\d+)\)', s))


def test_star_and_auto_single_host():
    l = _lib()
    for name in (b'STAR', b'AUTO'):
        out = _s(l.kf_gen_strategy_graphs, PEERS_1HOST, name).strip().split('\n')
        assert len(out) == 1
        red, bc = out[0].split(' bcast=')
        assert _edges(bc) == {(0, 1), (0, 2), (0, 3)}
        assert _edges(red) == {(1, 0), (2, 0), (3, 0)}
        assert _loops(red) == {0, 1, 2, 3}


def test_ring_has_np_rotations_and_each_is_a_chain():
    l = _lib()
    out = _s(l.kf_gen_strategy_graphs, PEERS_1HOST, b'RING').strip().split('\n')
    assert len(out) == 4
    for line in out:
        red, bc = line.split(' bcast=')
        assert len(_edges(red)) == 3 and len(_edges(bc)) == 3
        # reduce chain ends where the broadcast chain starts
        sinks = {b for _, b in _edges(red)} - {a for a, _ in _edges(red)}
        roots = {a for a, _ in _edges(bc)} - {b for _, b in _edges(bc)}
        assert sinks == roots and len(sinks) == 1


def test_clique_tree_binary_tree_star_multi():
    l = _lib()
    assert len(_s(l.kf_gen_strategy_graphs, PEERS_1HOST, b'CLIQUE').strip().split('\n')) == 4
    bt = _s(l.kf_gen_strategy_graphs, PEERS_1HOST, b'BINARY_TREE')
    assert _edges(bt.split(' bcast=')[1]) == {(0, 1), (0, 2), (1, 3)}
    # two-level shapes on 3 hosts: masters are ranks 0, 2, 4
    tree = _s(l.kf_gen_strategy_graphs, PEERS_2HOST, b'TREE').split(' bcast=')[1]
    assert _edges(tree) == {(0, 1), (2, 3), (0, 2), (0, 4)}
    bts = _s(l.kf_gen_strategy_graphs, PEERS_2HOST, b'BINARY_TREE_STAR').split(' bcast=')[1]
    assert _edges(bts) == {(0, 1), (2, 3), (0, 2), (0, 4)}
    assert len(_s(l.kf_gen_strategy_graphs, PEERS_2HOST, b'MULTI_BINARY_TREE_STAR').strip().split('\n')) == 3
    assert len(_s(l.kf_gen_strategy_graphs, PEERS_2HOST, b'MULTI_STAR').strip().split('\n')) == 3
    # AUTO on several hosts = BINARY_TREE_STAR
    assert _s(l.kf_gen_strategy_graphs, PEERS_2HOST, b'AUTO') == _s(l.kf_gen_strategy_graphs, PEERS_2HOST, b'BINARY_TREE_STAR')
    assert l.kf_gen_strategy_graphs(PEERS_1HOST, b'NOPE', C.create_string_buffer(16), 16) < 0


def test_every_bcast_graph_is_a_spanning_tree():
    l = _lib()
    for peers, n in ((PEERS_1HOST, 4), (PEERS_2HOST, 5)):
        for name in (b'STAR', b'TREE', b'BINARY_TREE', b'BINARY_TREE_STAR', b'MULTI_STAR', b'CLIQUE', b'RING',
                     b'MULTI_BINARY_TREE_STAR'):
            for line in _s(l.kf_gen_strategy_graphs, peers, name).strip().split('\n'):
                e = _edges(line.split(' bcast=')[1])
                assert len(e) == n - 1
                children = [b for _, b in e]
                assert len(set(children)) == n - 1  # every non-root has exactly one parent


def test_minimum_spanning_tree_total_weight():
    l = _lib()
    rng = np.random.RandomState(0)
    for n in (2, 5, 9):
        w = rng.rand(n, n).astype(np.float32)
        edges = (C.c_int32 * (2 * (n - 1)))()
        k = l.kf_minimum_spanning_tree(w.ctypes.data_as(C.POINTER(C.c_float)), n, edges)
        assert k == n - 1
        sym = (w + w.T) / 2
        got = sum(sym[edges[2 * i], edges[2 * i + 1]] for i in range(k))
        # Kruskal oracle
        parent = list(range(n))

        def find(a):
            while parent[a] != a:
                a = parent[a]
            return a
        tot = 0.0
        for wt, i, j in sorted((sym[i, j], i, j) for i in range(n) for j in range(i + 1, n)):
            a, b = find(i), find(j)
            if a != b:
                parent[a] = b
                tot += wt
        assert abs(got - tot) < 1e-5


@pytest.mark.parametrize('np_dt,code', [(np.float32, H.F32), (np.float64, H.F64), (np.int32, H.I32), (np.int64, H.I64),
                                        (np.uint8, H.U8), (np.int8, H.I8), (np.int16, H.I16), (np.float16, H.F16)])
def test_transform2_all_ops(np_dt, code):
    l = _lib()
    rng = np.random.RandomState(1)
    x = (rng.rand(1003) * 4 + 1).astype(np_dt)
    y = (rng.rand(1003) * 4 + 1).astype(np_dt)
    for op, f in ((0, np.add), (1, np.minimum), (2, np.maximum), (3, np.multiply)):
        z = np.zeros_like(x)
        l.kf_transform2(x.ctypes.data, y.ctypes.data, z.ctypes.data, x.size, code, op)
        ref = f(x, y)
        if np_dt == np.float16:
            assert np.allclose(z.astype(np.float32), ref.astype(np.float32), rtol=2e-3)
        else:
            assert np.array_equal(z, ref)


def test_transform2_bf16_sum_rounds_to_nearest_even():
    import torch
    l = _lib()
    a = torch.randn(1000).bfloat16()
    b = torch.randn(1000).bfloat16()
    z = torch.zeros(1000, dtype=torch.bfloat16)
    l.kf_transform2(a.data_ptr(), b.data_ptr(), z.data_ptr(), 1000, H.BF16, 0)
    assert torch.equal(z, (a.float() + b.float()).bfloat16())


@pytest.mark.parametrize('n', [0, 1, 7, 8, 9, 1003, 4099])
def test_transform2_16bit_ops_are_exact_and_in_place_safe(n):
    """fp16 / bf16 through all four ops: the 8-lane path and the scalar tail give the fp32 result rounded to
    nearest even, bit for bit, also when the output is one of the inputs (the graph walk reduces in place)."""
    import torch
    l = _lib()
    g = torch.Generator().manual_seed(n)
    for dt, code in ((torch.float16, H.F16), (torch.bfloat16, H.BF16)):
        a = (torch.randn(n, generator=g) * 3).to(dt)
        b = (torch.randn(n, generator=g) * 3).to(dt)
        for op, f in ((0, torch.add), (1, torch.minimum), (2, torch.maximum), (3, torch.mul)):
            ref = f(a.float(), b.float()).to(dt)
            z = torch.full((n, ), 7, dtype=dt)
            l.kf_transform2(a.data_ptr(), b.data_ptr(), z.data_ptr(), n, code, op)
            assert torch.equal(z, ref), (dt, op)
            x = a.clone()
            l.kf_transform2(x.data_ptr(), b.data_ptr(), x.data_ptr(), n, code, op)  # z is x
            assert torch.equal(x, ref), (dt, op, 'in place on x')
            y = b.clone()
            l.kf_transform2(a.data_ptr(), y.data_ptr(), y.data_ptr(), n, code, op)  # z is y
            assert torch.equal(y, ref), (dt, op, 'in place on y')
    for dt, code in ((torch.float32, H.F32), (torch.int32, H.I32), (torch.float64, H.F64), (torch.int64, H.I64)):
        a = (torch.randn(n, generator=g) * 100).to(dt)
        b = (torch.randn(n, generator=g) * 100).to(dt)
        for op, f in ((0, torch.add), (1, torch.minimum), (2, torch.maximum), (3, torch.mul)):
            x = a.clone()
            l.kf_transform2(x.data_ptr(), b.data_ptr(), x.data_ptr(), n, code, op)
            assert torch.equal(x, f(a, b)), (dt, op)


def test_cluster_resize_grow_and_shrink():
    l = _lib()
    c = json.dumps({'Runners': ['10.0.0.1:38080', '10.0.0.2:38080'], 'Workers': ['10.0.0.1:10000']}).encode()
    grown = json.loads(_s(l.kf_cluster_resize, c, 4))
    # growth goes to the least loaded runner, on the next free port
    assert grown['Workers'] == ['10.0.0.1:10000', '10.0.0.2:10000', '10.0.0.1:10001', '10.0.0.2:10001']
    shrunk = json.loads(_s(l.kf_cluster_resize, json.dumps(grown).encode(), 2))
    assert shrunk['Workers'] == grown['Workers'][:2]
    assert shrunk['Runners'] == grown['Runners']
    bad = json.dumps({'Runners': [], 'Workers': ['10.0.0.9:1']}).encode()
    assert l.kf_cluster_resize(bad, 2, C.create_string_buffer(64), 64) < 0


def test_hostfile_and_peer_list_generation():
    l = _lib()
    text = b"# comment\n10.0.0.1 slots=4\n10.0.0.2 slots=2 public_addr=gw.example\n\n"
    assert _s(l.kf_parse_hostfile, text) == '10.0.0.1:4:10.0.0.1,10.0.0.2:2:gw.example'
    assert l.kf_parse_hostfile(b"not-an-ip slots=1\n", C.create_string_buffer(64), 64) < 0
    peers = _s(l.kf_gen_peer_list, b'10.0.0.1:2,10.0.0.2:2', 3, b'10000-10010')
    assert peers == '10.0.0.1:10000,10.0.0.1:10001,10.0.0.2:10000'
    assert l.kf_gen_peer_list(b'10.0.0.1:2', 3, b'10000-10010', C.create_string_buffer(64), 64) < 0  # over capacity


def test_single_process_mode_and_ema():
    import kungfu_b200.python as kfpy
    assert kfpy.current_rank() == 0 and kfpy.current_cluster_size() == 1
    assert kfpy.current_local_rank() == 0 and kfpy.current_local_size() == 1
    assert not kfpy.detached()
    kfpy.run_barrier()
    assert kfpy.resize_cluster(1) == (False, False)
    from kungfu_b200.utils import ExponentialMovingAverage
    ema = ExponentialMovingAverage(0.9)
    vals = [round(ema.update(x), 4) for x in (2, 3, 4, 5, 6)]
    # the reference's unit test values (tests/python/unit, alpha 0.9)
    assert vals == [2.0, 2.1, 2.29, 2.561, 2.9049]


def test_python_utils_helpers():
    import io
    import time
    from kungfu_b200 import utils
    assert utils.show_rate(3 * 1024 ** 3, 1.0) == '3.00 GiB/s'
    assert utils.show_size(1536) == '1.5 KiB' and utils.show_size(12) == '12 B'
    r, dt = utils.measure(lambda x: x + 1, 41)
    assert r == 42 and dt >= 0
    t0 = time.monotonic()
    assert utils.poll(lambda: time.monotonic() - t0 > 0.1, timeout=2) is True
    assert utils.poll(lambda: False, timeout=0.1) is None
    assert isinstance(utils.list_nvidia_gpu_names(), list)
    buf = io.StringIO()
    with utils.StallDetector('slow op', period=0.05, out=buf) as sd:
        time.sleep(0.2)
    assert sd.reports >= 2 and 'slow op stalled for' in buf.getvalue()


def test_experiment_matrix_dry_run_and_noise_scale_scraper(tmp_path):
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, 'benchmarks', 'experiments.py'), '--models', 'resnet50,vgg16',
                          '--optimizers', 'ssgd,pair', '--np', '1,4', '--batch-sizes', '64', '--graph', '--dry-run'],
                         stdout=subprocess.PIPE, text=True, check=True).stdout.strip().splitlines()
    assert len(out) == 8
    assert sum('--graph' in line for line in out) == 4  # never for pair averaging
    assert all('kungfu-run' in line and 'train_bench.py' in line for line in out)
    log = tmp_path / 'run.log'
    log.write_text('step 10 np 2 loss 2.1 noise scale 120.5\nnoise\nstep 20 np 2 noise_scale=340.1\nstep 30 noise scale: 1e3\n')
    csv = tmp_path / 'gns.csv'
    res = subprocess.run([sys.executable, os.path.join(root, 'benchmarks', 'monitoring', 'noise_scale_from_log.py'), str(log),
                          '--csv', str(csv)], stdout=subprocess.PIPE, text=True, check=True).stdout
    assert '3 readings' in res
    assert csv.read_text().splitlines() == ['step,noise_scale', '10,120.5', '20,340.1', '30,1000']


def test_cuda_entry_points_reject_bad_arguments_before_touching_the_device():
    """Argument checks of the device C ABI run on the host: clear errors, no GPU needed."""
    from kungfu_b200.ops import _cuda as K
    lib = K.lib()
    buf = (C.c_char * 4096)()
    p = C.addressof(buf)
    p16 = (p + 15) // 16 * 16
    # linear: K and N must be multiples of 8; pointers 16-byte aligned; bias dtype f32/bf16
    assert lib.kfb_linear_act(p16, p16, None, 0, p16, None, 8, 12, 16, 0, None) != 0
    assert b'multiples of 8' in lib.kfb_last_error()
    assert lib.kfb_linear_act(p16 + 2, p16, None, 0, p16, None, 8, 16, 16, 0, None) != 0
    assert b'aligned' in lib.kfb_last_error()
    assert lib.kfb_linear_act(p16, p16, p16, K.I32, p16, None, 8, 16, 16, 0, None) != 0
    assert b'bias' in lib.kfb_last_error()
    assert lib.kfb_linear_act(p16, p16, None, 0, p16, None, 0, 16, 16, 0, None) == 0  # empty problem: nothing to do
    # activation backward: element count multiple of 8, known activation
    assert lib.kfb_act_bwd(p16, p16, p16, 12, 1, None) != 0
    assert lib.kfb_act_bwd(p16, p16, p16, 16, 7, None) != 0
    # add & norm: hidden size and dropout range
    assert lib.kfb_add_dropout_ln_fwd(p16, None, p16, p16, p16, None, None, None, 4, 100, 1e-5, 0.0, 1, None, 0, None) != 0
    assert b'hidden size' in lib.kfb_last_error()
    assert lib.kfb_add_dropout_ln_fwd(p16, None, p16, p16, p16, None, None, None, 4, 256, 1e-5, 1.5, 1, None, 0, None) != 0
    assert b'dropout' in lib.kfb_last_error()
    # column sum and qkv gather
    assert lib.kfb_colsum(p16, 4, 100, p16, p16, None) != 0
    assert b'multiple of 64' in lib.kfb_last_error()
    strides = (C.c_longlong * 9)(*([64, 8, 8] * 3))
    assert lib.kfb_pack_qkv_grad(p16, p16, p16, strides, 1, 1, 1, 12, p16, None) != 0
    assert b'head dim' in lib.kfb_last_error()
    assert lib.kfb_colsum_scratch_floats(768) >= 64 * 768
    assert lib.kfb_ln_partial_floats(4096, 768) % (2 * 768) == 0


def test_built_library_contains_the_blackwell_instructions():
    """The device library is sm_100a code that really uses the 5th-generation tensor cores, TMA, tensor memory,
    cluster barriers and NVLS multimem: check the SASS of the built .so (no GPU needed, cuobjdump ships with CUDA)."""
    import re
    import shutil
    import subprocess
    from kungfu_b200.ops import _cuda as K
    tool = shutil.which('cuobjdump') or '/usr/local/cuda/bin/cuobjdump'
    if not os.path.exists(tool):
        pytest.skip('cuobjdump not available')
    elf = subprocess.run([tool, '-lelf', K.LIB_PATH], stdout=subprocess.PIPE, text=True).stdout
    assert 'sm_100a' in elf, elf[:500]
    sass = subprocess.run([tool, '-sass', K.LIB_PATH], stdout=subprocess.PIPE, text=True, timeout=300).stdout
    for mnemonic, meaning in (('UTCHMMA', 'tcgen05.mma'), ('UTCHMMA.2CTA', 'tcgen05.mma.cta_group::2'),
                              ('UTMALDG.2D', 'cp.async.bulk.tensor (TMA)'), ('UTCBAR', 'tcgen05.commit'),
                              ('LDTM', 'tcgen05.ld'), ('UCGABAR_ARV', 'barrier.cluster'), ('LDGMC', 'multimem.ld_reduce'),
                              ('SYNCS.PHASECHK', 'mbarrier.try_wait')):
        assert mnemonic in sass, '%s (%s) not found in the SASS' % (mnemonic, meaning)


def test_json_and_hostfile_parsers_survive_mutated_and_deeply_nested_input():
    """Thousands of randomly mutated cluster documents, truncated ones, NUL / high bytes, and 100 000-level nesting
    (which used to overflow the recursive parser's stack): every call returns, valid documents still resize."""
    import random
    l = _lib()
    good = json.dumps({'Runners': ['127.0.0.1:38080'], 'Workers': ['127.0.0.1:10000', '127.0.0.1:10001']})
    buf = C.create_string_buffer(1 << 16)
    assert l.kf_cluster_resize(good.encode(), 3, buf, len(buf)) >= 0 and len(json.loads(buf.value.decode())['Workers']) == 3
    rng = random.Random(7)
    accepted = 0
    for _ in range(3000):
        s = list(good)
        for _ in range(rng.randint(1, 6)):
            op, pos = rng.random(), rng.randrange(len(s) + 1)
            if op < 0.4 and s:
                s.pop(min(pos, len(s) - 1))
            elif op < 0.8:
                s.insert(pos, rng.choice('{}[]":,\\0123456789.eE-+ntf\x00\xff abc'))
            else:
                s = s[:pos]
        data = ''.join(s).encode('latin1')
        accepted += l.kf_cluster_resize(data, rng.randint(-2, 9), buf, len(buf)) >= 0
        l.kf_parse_hostfile(data, buf, len(buf))
    assert accepted < 3000
    for depth in (65, 1000, 100000):
        assert l.kf_cluster_resize(b'[' * depth, 2, buf, len(buf)) < 0
        assert l.kf_cluster_resize(b'{"a":' * depth, 2, buf, len(buf)) < 0


def test_cluster_resize_properties_under_random_histories():
    """Property (hypothesis): along ANY sequence of sizes, a resized cluster has exactly that many distinct workers,
    all on hosts that have a runner; shrinking keeps a prefix; growing keeps every existing worker in place and puts
    the new ones on the least loaded host at a port nobody on that host uses."""
    from hypothesis import given, settings
    from hypothesis import strategies as st
    l = _lib()

    @settings(max_examples=80, deadline=None)
    @given(st.integers(1, 4), st.integers(2000, 60000), st.lists(st.integers(0, 24), min_size=1, max_size=10))
    def prop(n_hosts, base_port, sizes):
        runners = ['10.1.0.%d:38080' % (i + 1) for i in range(n_hosts)]
        cur = {'Runners': runners, 'Workers': ['10.1.0.1:%d' % base_port]}
        for n in sizes:
            nxt = json.loads(_s(l.kf_cluster_resize, json.dumps(cur).encode(), n))
            w_old, w_new = cur['Workers'], nxt['Workers']
            assert nxt['Runners'] == runners and len(w_new) == n and len(set(w_new)) == n
            assert all(w.split(':')[0] in {r.split(':')[0] for r in runners} for w in w_new)
            if n <= len(w_old):
                assert w_new == w_old[:n]
            else:
                assert w_new[:len(w_old)] == w_old
                for k in range(len(w_old), n):  # each addition goes to a least loaded host
                    load = {r.split(':')[0]: 0 for r in runners}
                    for w in w_new[:k]:
                        load[w.split(':')[0]] += 1
                    assert load[w_new[k].split(':')[0]] == min(load.values())
            if n > 0:
                cur = nxt

    prop()
