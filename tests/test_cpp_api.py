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
    base = 21000 + (os.getpid() * 53) % 15000
    res = subprocess.run([exe, '-q', '-np', str(np_), '-strategy', strategy, '-timeout', '60s', '-port-range',
                          '%d-%d' % (base, base + 32), '-port', str(base + 40), fake_agent],
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=90)
    assert res.returncode == 0, res.stdout[-3000:]
    assert res.stdout.count('OK fake-agent') == np_
