"""Loopback multi-process tests launched with our own kungfu-run, np = 1..4 x
every strategy, the way the reference tests itself
(/root/reference/scripts/tests/run-integration-tests.sh:30-38,
/root/reference/scripts/tests/run-op-tests.sh)."""
import os

import pytest
import torch

from conftest import ROOT, run_workers

W = os.path.join(ROOT, 'tests', 'workers')
STRATEGIES = ['STAR', 'MULTI_STAR', 'RING', 'CLIQUE', 'TREE', 'BINARY_TREE', 'BINARY_TREE_STAR',
              'MULTI_BINARY_TREE_STAR', 'AUTO']


def _ok(res, np_):
    assert res.returncode == 0, res.stdout[-4000:]
    assert res.stdout.count('OK') >= np_, res.stdout[-4000:]


@pytest.mark.parametrize('np_', [1, 2, 3])  # np=4 is covered strategy by strategy below
def test_ops_auto(np_):
    _ok(run_workers(np_, os.path.join(W, 'ops_worker.py')), np_)


@pytest.mark.parametrize('strategy', STRATEGIES)
def test_ops_every_strategy(strategy):
    _ok(run_workers(4, os.path.join(W, 'ops_worker.py'), strategy=strategy), 4)


@pytest.mark.parametrize('np_,kinds', [(1, 'ssgd,pair'), (2, 'ssgd,sma,monitors'), (3, 'pair'), (4, 'ssgd,sma,monitors')])
def test_optimizers(kinds, np_):
    """Several checks per launch (a worker spends most of its life importing torch)."""
    res = run_workers(np_, os.path.join(W, 'optim_worker.py'), args=[kinds])
    assert res.returncode == 0, res.stdout[-3000:]
    for kind in kinds.split(','):
        assert res.stdout.count('OK %s ' % kind) == np_, res.stdout[-3000:]


def test_mnist_slp_np_invariance(tmp_path):
    """S-SGD with np workers == one worker on the global batch, np in {1, 2, 3, 4}
    (/root/reference/tests/python/integration/test_mnist_slp.py:153-176 demands the same accuracy to 1e-6).
    In float64 the only difference between the splits is the order of a few additions: the weights agree
    to 1e-12 and the accuracy exactly. In float32 (the reference's precision) that reordering is visible
    in the last bits of every gradient sum; over 40 steps it reaches a few 1e-5 (bound used: 1e-4) and the accuracy, a count
    over 2000 samples, is still identical."""
    for np_ in (1, 2, 3, 4):  # both precisions in one launch per cluster size
        res = run_workers(np_, os.path.join(W, 'mnist_slp_worker.py'), args=['48', '40', str(tmp_path / ('slp%d.pt' % np_)), 'f64,f32'])
        assert res.returncode == 0 and res.stdout.count('OK rank') == 2 * np_, res.stdout[-3000:]
    for dt, tol in (('f64', 1e-12), ('f32', 1e-4)):
        outs = [torch.load(str(tmp_path / ('slp%d.pt.%s' % (np_, dt)))) for np_ in (1, 2, 3, 4)]
        for o in outs[1:]:
            assert torch.allclose(o['w'], outs[0]['w'], atol=tol, rtol=0), (dt, (o['w'] - outs[0]['w']).abs().max())
            assert torch.allclose(o['b'], outs[0]['b'], atol=tol, rtol=0), (dt, (o['b'] - outs[0]['b']).abs().max())
            assert abs(o['acc'] - outs[0]['acc']) <= 1e-6, (dt, o['acc'], outs[0]['acc'])
        assert outs[0]['acc'] > 0.3


def test_bad_worker_tears_down_job(tmp_path):
    """One failing worker fails the whole job quickly (reference: kungfu-bad-worker)."""
    script = tmp_path / 'bad.py'
    script.write_text("import sys, time\nimport kungfu_b200.torch as kf\n"
                      "r = kf.current_rank()\n"
                      "if r == 0:\n    sys.exit(3)\n"
                      "time.sleep(60)\n")
    res = run_workers(3, str(script), timeout=40)
    assert res.returncode != 0


@pytest.mark.parametrize('np_', [1, 3])
def test_ssgd_on_a_quadratic_follows_the_closed_form(np_):
    """x <- x - lr * mean_r(2x) = (1 - 2 lr) x on every worker, for any np, to 1e-6."""
    res = run_workers(np_, os.path.join(W, 'quadratic_worker.py'))
    assert res.returncode == 0, res.stdout[-3000:]
    assert res.stdout.count('OK quadratic') == np_


@pytest.mark.parametrize('np_', [2, 4])
def test_descriptor_passing_between_colocated_workers(np_):
    """SCM_RIGHTS over the control plane: the path every multi-GPU job uses to share its VMM handles."""
    res = run_workers(np_, os.path.join(W, 'fd_worker.py'))
    assert res.returncode == 0, res.stdout[-3000:]
    assert res.stdout.count('OK fd passing') == np_


def test_torchrun_environment_is_adopted(tmp_path):
    """bench.py is launched by torchrun (RANK / WORLD_SIZE / MASTER_PORT, no kungfu-run): the runtime derives an
    equivalent single-host peer list from that environment."""
    import subprocess
    import sys
    script = tmp_path / 'w.py'
    script.write_text("import torch\nimport kungfu_b200.torch as kf\n"
                      "r, n = kf.current_rank(), kf.current_cluster_size()\n"
                      "y = kf.ops.all_reduce_fn(torch.full((5,), float(r + 1)), name='t')\n"
                      "assert n == 3 and float(y[0]) == 6.0, (n, y)\n"
                      "kf.run_barrier()\nprint('OK torchrun-env rank', r, flush=True)\n")
    port = 29000 + os.getpid() % 900
    procs = []
    for r in range(3):
        env = dict(os.environ, RANK=str(r), LOCAL_RANK=str(r), WORLD_SIZE='3', MASTER_ADDR='127.0.0.1',
                   MASTER_PORT=str(port), PYTHONPATH=ROOT + os.pathsep + os.environ.get('PYTHONPATH', ''))
        env.pop('KUNGFU_SELF_SPEC', None)
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=120)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert sum('OK torchrun-env' in o for o in outs) == 3


def test_runner_restarts_a_worker_that_hit_the_loader_assertion(tmp_path):
    """Fault injection (reference: kungfu-bench-allreduce -rand-nccl-failure + the runner's retry hack): one worker dies
    once at start-up with glibc's dynamic-linker assertion on stderr; the runner starts it again and the job succeeds.
    A worker that fails for any other reason still fails the job."""
    script = tmp_path / 'flaky.py'
    script.write_text("import os, sys\n"
                      "flag = %r + '/crashed-once'\n"
                      "try:\n"
                      "    os.close(os.open(flag, os.O_CREAT | os.O_EXCL | os.O_WRONLY))  # exactly one worker wins\n"
                      "    first = True\n"
                      "except FileExistsError:\n"
                      "    first = False\n"
                      "if first:\n"
                      "    sys.stderr.write('Inconsistency detected by ld.so: dl-open.c: 272: dl_open_worker: Assertion failed!\\n')\n"
                      "    sys.exit(127)\n"
                      "import kungfu_b200.torch as kf\n"
                      "kf.run_barrier()\n"
                      "print('OK after restart rank', kf.current_rank(), flush=True)\n" % str(tmp_path))
    res = run_workers(2, str(script), timeout=90)
    assert res.returncode == 0, res.stdout[-3000:]
    assert res.stdout.count('OK after restart') == 2
    assert os.path.exists(str(tmp_path / 'crashed-once'))


def test_workers_survive_garbage_and_hostile_headers_on_their_ports(tmp_path):
    """A port scanner, a stray client or a corrupt peer must cost a connection, never a worker: random bytes, a valid
    hello followed by absurd name / body lengths, on every connection type, while a 2-worker job is running."""
    import socket
    import struct
    import subprocess
    import sys
    import time
    script = tmp_path / 'slow.py'
    script.write_text("import time, torch\nimport kungfu_b200.torch as kf\nkf.run_barrier()\n"
                      "open(%r + '/up-%%d' %% kf.current_rank(), 'w').close()\ntime.sleep(3)\n"
                      "y = kf.ops.all_reduce_fn(torch.ones(100000), name='after-garbage')\n"
                      "assert float(y[0]) == 2.0\nprint('OK survived', kf.current_rank(), flush=True)\n" % str(tmp_path))
    base = 20000 + (os.getpid() * 41) % 11000
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get('PYTHONPATH', ''))
    run = subprocess.Popen([os.path.join(ROOT, 'kungfu_b200', 'bin', 'kungfu-run'), '-q', '-np', '2', '-port-range',
                            '%d-%d' % (base, base + 8), '-port', str(base + 20), '-timeout', '80s', sys.executable, str(script)],
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)
    try:
        t0 = time.time()
        while time.time() - t0 < 60 and not all(os.path.exists(str(tmp_path / ('up-%d' % r))) for r in (0, 1)):
            time.sleep(0.1)
        hello = lambda t: struct.pack('<IHHII', 0x4b46b200, t, 1, 0x7f000001, 0)  # noqa: E731
        bad = [os.urandom(512), b'GET / HTTP/1.1\r\n\r\n',
               struct.pack('<I', 5) + b'hello' + struct.pack('<IQ', 0, 2 ** 63 + 7),   # body length beyond any tensor
               struct.pack('<I', 2 ** 31),                                              # name length beyond the limit
               struct.pack('<I', 3) + b'abc' + struct.pack('<IQ', 0, 2 ** 30) + b'x' * 64]  # body that never arrives
        for port in (base, base + 1):
            for t in range(5):
                for body in bad:
                    try:
                        s = socket.create_connection(('127.0.0.1', port), timeout=2)
                        s.sendall((hello(t) if body is not bad[0] else b'') + body)
                        s.close()
                    except OSError:
                        pass
        log, _ = run.communicate(timeout=90)
    finally:
        if run.poll() is None:
            run.kill()
    assert run.returncode == 0, log[-3000:]
    assert log.count('OK survived') == 2


def test_modelarts_environment_describes_the_cluster():
    import sys
    """No kungfu-run: two tasks find each other through the ModelArts variables
    (DLS_TASK_INDEX / DLS_TASK_NUMBER / BATCH_CUSTOM<i>_HOSTS), one of them by host name
    (reference: srcs/go/platforms/modelarts/modelarts.go:25-115)."""
    import subprocess
    base = 21000 + (os.getpid() * 13) % 9000
    code = ('import torch, kungfu_b200.torch as kf\n'
            'x = kf.ops.all_reduce_fn(torch.tensor([float(kf.current_rank() + 1)]))\n'
            'assert kf.current_cluster_size() == 2 and float(x) == 3.0, (kf.current_cluster_size(), x)\n'
            'kf.run_barrier()\n'
            'print("OK modelarts rank", kf.current_rank())\n')
    procs = []
    for t in range(2):
        env = dict(os.environ, PYTHONPATH=ROOT, DLS_TASK_INDEX=str(t), DLS_TASK_NUMBER='2',
                   BATCH_CUSTOM0_HOSTS='127.0.0.1:%d' % base, BATCH_CUSTOM1_HOSTS='localhost:%d' % (base + 1))
        for k in [k for k in env if k.startswith('KUNGFU_')]:
            del env[k]
        procs.append(subprocess.Popen([sys.executable, '-c', code], stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                                      text=True, env=env))
    outs = [p.communicate(timeout=120)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), '\n'.join(outs)
    assert sorted(o.strip().splitlines()[-1] for o in outs) == ['OK modelarts rank 0', 'OK modelarts rank 1']
    # unusable variables are an error, not a silent single-process run
    env = dict(os.environ, PYTHONPATH=ROOT, DLS_TASK_INDEX='0', DLS_TASK_NUMBER='2', BATCH_CUSTOM0_HOSTS='127.0.0.1:%d' % base)
    for k in [k for k in env if k.startswith('KUNGFU_')]:
        del env[k]
    r = subprocess.run([sys.executable, '-c', code], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env,
                       timeout=60)
    assert r.returncode != 0 and 'BATCH_CUSTOM1_HOSTS' in r.stdout, r.stdout


@pytest.mark.parametrize('np_', [1, 2, 4])
def test_p2p_model_ops(np_):
    """Selectors, ModelAveraging / RequestModel and their prefetching variants (native double buffer)."""
    res = run_workers(np_, os.path.join(ROOT, 'tests', 'workers', 'p2p_model_worker.py'), timeout=90)
    assert res.returncode == 0 and res.stdout.count('OK p2p-model') == np_, res.stdout[-3000:]


def _state(pid):
    try:
        with open('/proc/%d/stat' % pid) as f:
            return f.read().rsplit(')', 1)[1].split()[0]
    except OSError:
        return None  # gone (reaped)


def _children(pid):
    out = []
    for d in os.listdir('/proc'):
        if d.isdigit():
            try:
                with open('/proc/%s/stat' % d) as f:
                    if int(f.read().rsplit(')', 1)[1].split()[1]) == pid:
                        out.append(int(d))
            except (OSError, ValueError, IndexError):
                pass
    return out


def test_workers_never_outlive_their_runner_and_a_killed_worker_stops_the_job(tmp_path):
    """(1) SIGKILL of the runner: its workers are killed by the kernel (PR_SET_PDEATHSIG), nobody keeps a port or a
    GPU. (2) SIGKILL of one worker in the middle of a job: the runner stops the others and reports 128 + 9 within
    seconds (reference: local.RunAll cancels every sibling on the first failure, utils/runner/local/local.go:62-92)."""
    import signal
    import subprocess
    import time
    exe = os.path.join(ROOT, 'kungfu_b200', 'bin', 'kungfu-run')
    base = 20000 + (os.getpid() * 29) % 11000
    run = subprocess.Popen([exe, '-q', '-np', '2', '-port-range', '%d-%d' % (base, base + 8), '-port', str(base + 20), 'sleep', '60'],
                           stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    kids = []
    for _ in range(50):
        kids = _children(run.pid)
        if len(kids) == 2:
            break
        time.sleep(0.1)
    assert len(kids) == 2
    run.send_signal(signal.SIGKILL)
    run.wait(timeout=10)
    for _ in range(50):
        if all(_state(k) in (None, 'Z', 'X') for k in kids):
            break
        time.sleep(0.1)
    assert all(_state(k) in (None, 'Z', 'X') for k in kids), [(_state(k)) for k in kids]  # dead (a zombie until init reaps it)

    script = tmp_path / 'victim.py'
    script.write_text('import os, signal, time, torch\nimport kungfu_b200.torch as kf\n'
                      'for i in range(400):\n'
                      '    kf.ops.all_reduce_fn(torch.ones(1000))\n'
                      '    if kf.current_rank() == 1 and i == 5:\n'
                      '        os.kill(os.getpid(), signal.SIGKILL)\n'
                      '    time.sleep(0.05)\n')
    t0 = time.time()
    res = run_workers(3, str(script), timeout=60)
    assert res.returncode == 137 and 'exited with 137; stopping the others' in res.stdout, res.stdout[-2000:]
    assert time.time() - t0 < 40
