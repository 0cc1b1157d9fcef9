"""Real multi-GPU runs of the peer-memory path (skipped with < 2 GPUs)."""
import os

import pytest
import torch

from conftest import ROOT, run_workers

pytestmark = pytest.mark.gpu


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.parametrize('np_', [2, 4, 8])
def test_peer_memory_path(np_):
    if _ngpu() < np_:
        pytest.skip('needs %d GPUs' % np_)
    # KUNGFU_REQUIRE_NVLS: a box whose multicast set-up fails must fail here, not pass on the fall-back kernels
    res = run_workers(np_, os.path.join(ROOT, 'tests', 'workers', 'gpu_peer_worker.py'), timeout=600,
                      extra_flags=['-allow-nvlink'], env={'KUNGFU_REQUIRE_NVLS': '1', 'KUNGFU_PROBE_VARIANTS': '0'})
    assert res.returncode == 0, res.stdout[-6000:]
    assert res.stdout.count('OK rank') == np_, res.stdout[-6000:]
    assert 'multicast=True' in res.stdout, res.stdout[-2000:]


def test_peer_memory_path_ipc_fallback():
    if _ngpu() < 2:
        pytest.skip('needs 2 GPUs')
    res = run_workers(2, os.path.join(ROOT, 'tests', 'workers', 'gpu_peer_worker.py'), timeout=300,
                      extra_flags=['-allow-nvlink'], env={'KUNGFU_SYMM_MODE': 'ipc'})
    assert res.returncode == 0, res.stdout[-6000:]


def test_gpu_elastic_resize(tmp_path):
    """resize_cluster on GPUs: peer memory is re-mapped for every new membership,
    fused optimizer state is consolidated / re-broadcast, survivors keep running."""
    import collections
    import socket
    n = _ngpu()
    if n < 2:
        pytest.skip('needs 2 GPUs')
    schedule, steps = ('1:2,2:3,4:3,2:2,1:2', 12) if n >= 4 else ('1:2,2:3,1:2', 7)
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    log = str(tmp_path / 'steps.log')
    res = run_workers(1, os.path.join(ROOT, 'tests', 'workers', 'gpu_elastic_worker.py'), args=[schedule, str(steps), log],
                      timeout=400, extra_flags=['-w', '-builtin-config-port', str(port), '-allow-nvlink'])
    assert res.returncode == 0, res.stdout[-6000:]
    sizes, pids = collections.defaultdict(set), collections.defaultdict(set)
    for line in open(log):
        step, rank, np_, pid, loss = line.split()
        sizes[int(step)].add(int(np_))
        pids[int(step)].add(int(pid))
    off = 0
    for part in schedule.split(','):
        size, k = map(int, part.split(':'))
        for st in range(off, off + k):
            assert sizes[st] == {size}, (st, sizes[st])
        off += k
    first = next(iter(pids[0]))
    assert all(first in pids[st] for st in range(steps))


def test_gpu_elastic_resize_gives_memory_back(tmp_path):
    """Six grow / shrink cycles (1 -> 2 -> 1 ...): the survivor's free device memory at cluster size 1 must
    come back to where it was after the first cycles, i.e. retired communicators really unmap the peers'
    blocks and release their own (the reference destroyed its NCCL communicators and dropped the removed
    peers' connections on every resize, srcs/cpp/src/tensorflow/ops/gpu/scheduler.cpp:49-69)."""
    import socket
    if _ngpu() < 2:
        pytest.skip('needs 2 GPUs')
    cycles = 6
    schedule = ','.join(['1:1', '2:1'] * cycles + ['1:1'])
    steps = 2 * cycles + 1
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    log, memlog = str(tmp_path / 'steps.log'), str(tmp_path / 'mem.log')
    res = run_workers(1, os.path.join(ROOT, 'tests', 'workers', 'gpu_elastic_worker.py'), args=[schedule, str(steps), log],
                      timeout=300, extra_flags=['-w', '-builtin-config-port', str(port), '-allow-nvlink'],
                      env={'KUNGFU_TEST_MEMLOG': memlog})
    assert res.returncode == 0, res.stdout[-6000:]
    free_at_1 = [int(l.split()[2]) for l in open(memlog) if l.split()[1] == '1']
    assert len(free_at_1) == cycles + 1, free_at_1
    # the first two visits still grow the caching allocator; after that every visit must look the same
    base = free_at_1[2]
    worst = max(abs(f - base) for f in free_at_1[2:])
    assert worst <= 64 << 20, 'free memory at np=1 drifts over the cycles: %s' % [f >> 20 for f in free_at_1]


def test_hierarchical_two_hosts_on_loopback():
    """2 'hosts' (loopback aliases) x 2 GPUs: local NVLink kernels + cross-host control plane."""
    import subprocess
    import sys
    if _ngpu() < 4:
        pytest.skip('needs 4 GPUs')
    exe = os.path.join(ROOT, 'kungfu_b200', 'bin', 'kungfu-run')
    base = 23000 + (os.getpid() * 17) % 9000
    hosts = '127.0.0.1:2,127.0.0.2:2'
    procs = []
    for i, ip in enumerate(('127.0.0.1', '127.0.0.2')):
        env = dict(os.environ, PYTHONPATH=ROOT, KUNGFU_TEST_GPU_OFFSET=str(2 * i))
        cmd = [exe, '-q', '-np', '4', '-H', hosts, '-self', ip, '-allow-nvlink', '-timeout', '200s', '-port-range',
               '%d-%d' % (base, base + 16), '-port', str(base + 20), sys.executable,
               os.path.join(ROOT, 'tests', 'workers', 'gpu_hier_worker.py')]
        procs.append(subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env))
    outs = [p.communicate(timeout=260)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), '\n'.join(o[-3000:] for o in outs)
    assert sum(o.count('OK hier') for o in outs) == 4
