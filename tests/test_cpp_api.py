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


def test_runtime_internals(tmp_path):
    """tests/cpp/runtime_unit_test.cpp: the subjects of the reference's Go unit tests (host / peer lists, cluster
    resize, hostfile, forest arrays, runner flags, CUDA_VISIBLE_DEVICES remap, versioned store window, message
    framing incl. hostile headers, mailbox rendezvous, rate accumulators) against the runtime's own classes."""
    out = str(tmp_path / 'runtime-unit')
    lib = os.path.join(ROOT, 'kungfu_b200', 'lib')
    subprocess.check_call(['g++', '-O1', '-std=c++17', '-pthread', '-I', os.path.join(ROOT, 'include'), '-I',
                           os.path.join(ROOT, 'csrc'), os.path.join(ROOT, 'tests', 'cpp', 'runtime_unit_test.cpp'), '-o', out,
                           '-L', lib, '-lkungfu_host', '-Wl,-rpath,' + lib])
    res = subprocess.run([out], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert res.returncode == 0 and 'OK runtime-unit' in res.stdout, res.stdout[-3000:]


def test_worker_environment_through_the_runner():
    """What a worker really sees: the runner's CUDA_VISIBLE_DEVICES picks and orders the devices
    (reference: kungfu/job/cuda_visible_device.go), values containing '=' survive the environment merge
    (reference: proc/proc_test.go), and the worker-side lookup turns the device into an index."""
    exe = os.path.join(ROOT, 'kungfu_b200', 'bin', 'kungfu-run')
    base = 21000 + (os.getpid() * 67) % 11000
    script = 'echo "W $KUNGFU_SELF_SPEC kf=$KUNGFU_CUDA_VISIBLE_DEVICES cuda=$CUDA_VISIBLE_DEVICES y=$Y"'

    def run(extra, env_cuda):
        env = dict(os.environ, Y='Z=2')
        env.pop('CUDA_VISIBLE_DEVICES', None)
        if env_cuda is not None:
            env['CUDA_VISIBLE_DEVICES'] = env_cuda
        r = subprocess.run([exe, '-q', '-np', '2', '-port-range', '%d-%d' % (base, base + 8), '-port', str(base + 40)] + extra +
                           ['sh', '-c', script], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env, timeout=60)
        assert r.returncode == 0, r.stdout
        return sorted(l[l.index('W '):] for l in r.stdout.splitlines() if 'W 127.0.0.1:' in l)

    a, b = run([], '4,6')
    assert a.endswith(':%d kf=4 cuda=4 y=Z=2' % base) and b.endswith(':%d kf=6 cuda=6 y=Z=2' % (base + 1))
    a, b = run(['-allow-nvlink'], '4,6')  # the node stays visible; the worker looks its device up in the list
    assert a.endswith('kf=4 cuda=4,6 y=Z=2') and b.endswith('kf=6 cuda=4,6 y=Z=2')
    a, b = run([], None)                  # nothing set: slot == device
    assert a.endswith('kf=0 cuda=0 y=Z=2') and b.endswith('kf=1 cuda=1 y=Z=2')
    a, b = run([], 'GPU-aa,GPU-bb')       # uuid form is handed on by position
    assert a.endswith('kf=0 cuda=GPU-aa y=Z=2') and b.endswith('kf=1 cuda=GPU-bb y=Z=2')

    from kungfu_b200.python import _cuda_index_from_env as idx
    never = lambda: 1 / 0  # noqa: E731
    assert idx('6', '4,6', None, never) == 1        # -allow-nvlink: device 6 is the second visible one
    assert idx('6', '6', None, never) == 0          # narrowed to one device
    assert idx('1', None, None, never) == 1         # nothing hidden
    assert idx('-1', '4,6', None, never) == -1      # the launcher had no GPU for this worker
    assert idx('1', 'GPU-aa,GPU-bb', None, never) == 1 and idx('1', 'GPU-bb', None, never) == 0
    assert idx('1', '0,1,2,3', '3', never) == 1     # a KungFu launcher wins over LOCAL_RANK
    assert idx(None, '0,1', '3', never) == 3        # torchrun: LOCAL_RANK
    assert idx(None, None, None, lambda: 2) == 2    # bare process: the local rank
