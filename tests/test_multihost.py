"""Multi-host behaviour on one machine: three loopback aliases act as hosts,
each with its own kungfu-run (the way kungfu-rrun would start them)."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT


@pytest.mark.parametrize('strategy', ['BINARY_TREE_STAR', 'TREE', 'MULTI_STAR', 'MULTI_BINARY_TREE_STAR', 'RING', 'AUTO'])
def test_three_hosts_on_loopback(strategy):
    exe = os.path.join(ROOT, 'kungfu_b200', 'bin', 'kungfu-run')
    base = 22000 + (os.getpid() * 31) % 10000
    hosts = '127.0.0.1:2,127.0.0.2:2,127.0.0.3:1'
    env = dict(os.environ, PYTHONPATH=ROOT)
    procs = []
    for ip in ('127.0.0.1', '127.0.0.2', '127.0.0.3'):
        cmd = [exe, '-q', '-np', '5', '-H', hosts, '-self', ip, '-strategy', strategy, '-timeout', '90s', '-port-range',
               '%d-%d' % (base, base + 16), '-port', str(base + 20), sys.executable,
               os.path.join(ROOT, 'tests', 'workers', 'multihost_worker.py')]
        procs.append(subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env))
    outs = [p.communicate(timeout=120)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), '\n'.join(o[-1500:] for o in outs)
    assert sum(o.count('OK multihost') for o in outs) == 5
