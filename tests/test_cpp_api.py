"""kungfu::Peer C++ API (include/kungfu/peer.hpp) under kungfu-run, np 1..4 x strategies."""
import os
import subprocess

import pytest

from conftest import ROOT


@pytest.fixture(scope='module')
def fake_agent(tmp_path_factory):
    out = str(tmp_path_factory.mktemp('cpp') / 'fake-agent')
    lib = os.path.join(ROOT, 'kungfu_b200', 'lib')
    subprocess.check_call(['g++', '-O1', '-std=c++17', '-pthread', '-I', os.path.join(ROOT, 'include'),
                           os.path.join(ROOT, 'tests', 'cpp', 'fake_agent.cpp'), '-o', out,
                           '-L', lib, '-lkungfu_host', '-Wl,-rpath,' + lib])
    return out


@pytest.mark.parametrize('np_,strategy', [(1, 'AUTO'), (2, 'STAR'), (3, 'RING'), (4, 'BINARY_TREE_STAR'), (4, 'CLIQUE'),
                                          (4, 'MULTI_BINARY_TREE_STAR'), (3, 'TREE')])
def test_fake_agent(fake_agent, np_, strategy):
    exe = os.path.join(ROOT, 'kungfu_b200', 'bin', 'kungfu-run')
    base = 21000 + (os.getpid() * 53) % 11000
    res = subprocess.run([exe, '-q', '-np', str(np_), '-strategy', strategy, '-timeout', '60s', '-port-range',
                          '%d-%d' % (base, base + 32), '-port', str(base + 40), fake_agent],
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=90)
    assert res.returncode == 0, res.stdout[-3000:]
    assert res.stdout.count('OK fake-agent') == np_


def test_cpp_utils_header(tmp_path):
    """include/kungfu/utils.hpp: Channel, Latch, ExponentialMovingAverage, HandleTable, trace / check macros."""
    out = str(tmp_path / 'utils-test')
    subprocess.check_call(['g++', '-O1', '-std=c++17', '-pthread', '-I', os.path.join(ROOT, 'include'),
                           os.path.join(ROOT, 'tests', 'cpp', 'utils_test.cpp'), '-o', out])
    res = subprocess.run([out], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=60)
    assert res.returncode == 0 and 'OK utils' in res.stdout, res.stdout


@pytest.mark.parametrize('tool,args,np_', [
    ('kungfu-bench-allreduce', ['-model', '6x5000', '-mode', 'par', '-epochs', '1', '-warmup', '1'], 3),
    ('kungfu-bench-allreduce', ['-model', 'resnet50', '-mode', 'seq', '-fuse', '-epochs', '1', '-warmup', '0'], 2),
    ('kungfu-bench-p2p', ['-model', '4x10000', '-epochs', '1', '-warmup', '1'], 3),
])
def test_native_bench_tools(tool, args, np_):
    """Host data-plane micro-benchmarks (reference: tests/go/cmd/kungfu-bench-allreduce, kungfu-bench-p2p)."""
    bindir = os.path.join(ROOT, 'kungfu_b200', 'bin')
    base = 21000 + (os.getpid() * 59 + len(args) * 101) % 11000
    res = subprocess.run([os.path.join(bindir, 'kungfu-run'), '-q', '-np', str(np_), '-timeout', '60s', '-port-range',
                          '%d-%d' % (base, base + 32), '-port', str(base + 40), os.path.join(bindir, tool)] + args,
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert res.returncode == 0, res.stdout[-3000:]
    assert 'OK ' + tool in res.stdout and 'rate' in res.stdout


def test_fake_adaptive_trainer(tmp_path):
    """C++ elastic trainer: kungfu::Peer::ResizeCluster along a schedule under kungfu-run -w."""
    out = str(tmp_path / 'fake-adaptive-trainer')
    lib = os.path.join(ROOT, 'kungfu_b200', 'lib')
    subprocess.check_call(['g++', '-O1', '-std=c++17', '-pthread', '-I', os.path.join(ROOT, 'include'),
                           os.path.join(ROOT, 'tests', 'cpp', 'fake_adaptive_trainer.cpp'), '-o', out,
                           '-L', lib, '-lkungfu_host', '-Wl,-rpath,' + lib])
    exe = os.path.join(ROOT, 'kungfu_b200', 'bin', 'kungfu-run')
    base = 21000 + (os.getpid() * 71) % 11000
    res = subprocess.run([exe, '-q', '-w', '-np', '1', '-H', '127.0.0.1:4', '-timeout', '90s', '-port-range',
                          '%d-%d' % (base, base + 32), '-port', str(base + 40), out, '3:1,3:3,3:4,3:2', '50000'],
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert res.returncode == 0, res.stdout[-3000:]
    assert 'resized 1 -> 3' in res.stdout and 'resized 3 -> 4' in res.stdout and 'resized 4 -> 2' in res.stdout
    assert res.stdout.count('OK fake-adaptive-trainer') == 2 and res.stdout.count('detached at step') == 2
