import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


@pytest.fixture(scope='session', autouse=True)
def _native_built():
    """Native libraries are built in-tree; tests never run against a stale or missing build."""
    from kungfu_b200 import _build
    _build.build_host(verbose=False)
    import torch
    if torch.cuda.is_available():
        lib = os.path.join(ROOT, 'kungfu_b200', 'lib', 'libkungfu_cuda.so')
        if not os.path.exists(lib):
            _build.build_cuda(verbose=False)
    yield


def run_workers(np_, script, args=(), strategy='AUTO', timeout=120, extra_flags=(), env=None):
    """Launch `script` on np_ loopback workers with our own kungfu-run; returns CompletedProcess."""
    import subprocess
    exe = os.path.join(ROOT, 'kungfu_b200', 'bin', 'kungfu-run')
    port_base = 20000 + (os.getpid() * 37) % 12000  # stay below the ephemeral range (32768+)
    cmd = [exe, '-q', '-np', str(np_), '-H', '127.0.0.1:%d' % max(np_, 8), '-strategy', strategy,
           '-port-range', '%d-%d' % (port_base, port_base + 64), '-port', str(port_base + 100),
           '-timeout', '%ds' % timeout] + list(extra_flags) + [sys.executable, script] + list(args)
    e = dict(os.environ)
    e['PYTHONPATH'] = ROOT + os.pathsep + e.get('PYTHONPATH', '')
    if env:
        e.update(env)
    return subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout + 30, env=e)
