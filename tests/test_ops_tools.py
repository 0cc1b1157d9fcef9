"""Operator tools: kungfu-test-util (-kill / -propose), kungfu-notify-start, kungfu-config-client
(reference: tests/go/cmd/kungfu-test-util, kungfu-notify-start, kungfu-config-client-example)."""
import json
import os
import subprocess
import sys
import time
import urllib.error
import urllib.request

from conftest import ROOT

BIN = os.path.join(ROOT, 'kungfu_b200', 'bin')


def _base(salt):
    return 21000 + (os.getpid() * 61 + salt * 977) % 11000


def _get(url):
    try:
        with urllib.request.urlopen(url, timeout=5) as r:
            return r.status, r.read().decode()
    except urllib.error.HTTPError as e:
        return e.code, ''


def test_config_client_and_propose(tmp_path):
    port = _base(1)
    url = 'http://127.0.0.1:%d/config' % port
    srv = subprocess.Popen([os.path.join(BIN, 'kungfu-config-server'), '-port', str(port)],
                           stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    try:
        for _ in range(50):
            try:
                _get(url)
                break
            except Exception:
                time.sleep(0.1)
        assert _get(url)[0] == 404
        example = subprocess.run([os.path.join(BIN, 'kungfu-config-client'), '-example', '-H', '127.0.0.1:3'],
                                 stdout=subprocess.PIPE, text=True, check=True).stdout
        cluster = json.loads(example)
        assert len(cluster['Workers']) == 3 and len(cluster['Runners']) == 1
        f = tmp_path / 'c.json'
        f.write_text(example)
        out = subprocess.run([os.path.join(BIN, 'kungfu-test-util'), '-propose', str(f), '-server', url],
                             stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        assert out.returncode == 0, out.stdout
        status, body = _get(url)
        assert status == 200 and json.loads(body)['Workers'] == cluster['Workers']
        # the periodic client PUTs clusters of random size within the host list's capacity
        out = subprocess.run([os.path.join(BIN, 'kungfu-config-client'), '-server', url, '-H', '127.0.0.1:3',
                              '-period', '50ms', '-ttl', '200ms'], stdout=subprocess.PIPE, text=True, timeout=30)
        assert out.returncode == 0 and out.stdout.count('-> 200') >= 2
        assert 1 <= len(json.loads(_get(url)[1])['Workers']) <= 3
        out = subprocess.run([os.path.join(BIN, 'kungfu-config-client'), '-server', url, '-clear'],
                             stdout=subprocess.PIPE, text=True, timeout=30)
        assert out.returncode == 0
        assert _get(url)[0] == 404
    finally:
        srv.kill()
        srv.wait()


WORKER = """
import sys, time
sys.path.insert(0, %r)
import kungfu_b200.torch as kf
print('up', kf.current_rank(), kf.current_cluster_size(), flush=True)
kf.run_barrier()
time.sleep(%f)
print('done', kf.current_rank(), flush=True)
"""


def test_notify_start_initialises_an_empty_runner(tmp_path):
    base = _base(2)
    script = tmp_path / 'w.py'
    script.write_text(WORKER % (ROOT, 0.2))
    cfg = tmp_path / 'cluster.json'
    cfg.write_text(json.dumps({'Runners': ['127.0.0.1:%d' % (base + 40)],
                               'Workers': ['127.0.0.1:%d' % base, '127.0.0.1:%d' % (base + 1)]}))
    run = subprocess.Popen([os.path.join(BIN, 'kungfu-run'), '-q', '-w', '-init-version', '-1', '-H', '127.0.0.1:2',
                            '-port', str(base + 40), '-port-range', '%d-%d' % (base, base + 16), '-timeout', '60s',
                            sys.executable, str(script)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    try:
        out = subprocess.run([os.path.join(BIN, 'kungfu-notify-start'), '-config-server', str(cfg)],
                             stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=60)
        assert out.returncode == 0, out.stdout
        log, _ = run.communicate(timeout=90)
    finally:
        if run.poll() is None:
            run.kill()
    assert run.returncode == 0, log[-2000:]
    assert 'waiting to be initialized' in log
    assert log.count('up ') == 2 and log.count('done ') == 2


def test_kill_asks_one_worker_to_exit(tmp_path):
    base = _base(3)
    script = tmp_path / 'w.py'
    # after the barrier both workers drop a marker; rank 1 would then sleep for a minute, the exit
    # request ends it early (exit code 0, so the job succeeds)
    script.write_text("""
import sys, time
sys.path.insert(0, %r)
import kungfu_b200.torch as kf
kf.run_barrier()
open(%r + '/ready-%%d' %% kf.current_rank(), 'w').close()
time.sleep(4 if kf.current_rank() == 0 else 60)
print('done', kf.current_rank(), flush=True)
""" % (ROOT, str(tmp_path)))
    run = subprocess.Popen([os.path.join(BIN, 'kungfu-run'), '-q', '-np', '2', '-port', str(base + 40), '-port-range',
                            '%d-%d' % (base, base + 16), '-timeout', '100s', sys.executable, str(script)],
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    try:
        t0 = time.time()
        while time.time() - t0 < 60 and not all(os.path.exists(str(tmp_path / ('ready-%d' % r))) for r in (0, 1)):
            time.sleep(0.2)
        assert os.path.exists(str(tmp_path / 'ready-1')), 'workers never came up'
        out = subprocess.run([os.path.join(BIN, 'kungfu-test-util'), '-kill', '127.0.0.1:%d' % (base + 1)],
                             stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=30)
        assert out.returncode == 0, out.stdout
        log, _ = run.communicate(timeout=60)
    finally:
        if run.poll() is None:
            run.kill()
    assert run.returncode == 0, log[-2000:]
    assert 'done 0' in log and 'done 1' not in log


def test_cluster_manager_example():
    """examples/cluster_manager.py: config server + two empty runners on loopback aliases + notify-start + PUT resizes."""
    base = _base(4)
    res = subprocess.run([sys.executable, os.path.join(ROOT, 'examples', 'cluster_manager.py'), '--max-step', '50',
                          '--period', '3', '--sizes', '2,4,3', '--port-base', str(base)],
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=240)
    assert res.returncode == 0, res.stdout[-3000:]
    assert 'runners exited with [0, 0]' in res.stdout
    assert res.stdout.count('finished 50 steps with np=3') == 3


def test_random_resize_soak():
    """A config client PUTs a cluster of random size (1..4) every 300 ms while the workers poll the config
    server after every step (reference: the docker-compose cluster tests with kungfu-config-client-example)."""
    base = _base(5)
    cfg_port = base + 60
    url = 'http://127.0.0.1:%d/config' % cfg_port
    worker = [sys.executable, os.path.join(ROOT, 'examples', 'cluster_manager.py'), '--worker', '--max-step', '70']
    run = subprocess.Popen([os.path.join(BIN, 'kungfu-run'), '-q', '-w', '-np', '2', '-H', '127.0.0.1:4', '-port',
                            str(base + 40), '-port-range', '%d-%d' % (base, base + 16), '-builtin-config-port',
                            str(cfg_port), '-config-server', url, '-timeout', '150s'] + worker,
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    try:
        time.sleep(4)  # let the first workers import torch and start stepping
        client = subprocess.run([os.path.join(BIN, 'kungfu-config-client'), '-server', url, '-H', '127.0.0.1:4', '-port',
                                 str(base + 40), '-port-range', '%d-%d' % (base, base + 16), '-period', '300ms',
                                 '-ttl', '5s'], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=60)
        assert client.stdout.count('-> 200') >= 5, client.stdout
        log, _ = run.communicate(timeout=170)
    finally:
        if run.poll() is None:
            run.kill()
    assert run.returncode == 0, log[-3000:]
    assert 'finished 70 steps' in log
    assert 'now rank' in log or 'leaves at step' in log  # at least one membership change was observed


def test_virtual_cluster_under_a_shuffling_config_client():
    """Three loopback nodes x 2 slots, one kept-alive watch-mode runner per node, a stand-alone config server and a
    client that PUTs random SUBSETS of the slots in random order (ranks move, nodes empty out and are refilled,
    sometimes nobody survives): every worker checks every all-reduce, and after a final full cluster all six
    slots finish. Reference: the docker-compose cluster tests (benchmarks/adaptation/gen-compose.py:60-123,
    .github/workflows/cluster-test-1.sh, kungfu-config-client-example.go:96-116)."""
    res = subprocess.run([sys.executable, os.path.join(ROOT, 'benchmarks', 'adaptation', 'virtual_cluster.py'), '--nodes', '3',
                          '--cap', '2', '--np', '2', '--ttl', '6', '--period', '0.5', '--seed', str(os.getpid() % 1000),
                          '--port-base', str(_base(9))], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=280)
    assert res.returncode == 0, res.stdout[-4000:]
    assert '6 of 6 workers finished' in res.stdout and 'runners exited with [0, 0, 0]' in res.stdout, res.stdout[-2000:]


def test_config_client_shuffle_is_a_random_subset_in_random_order():
    seen = set()
    for seed in range(8):
        r = subprocess.run([os.path.join(BIN, 'kungfu-config-client'), '-H', '127.0.0.1:2,127.0.0.2:2', '-port-range', '100-110',
                            '-shuffle', '-seed', str(seed), '-example'], stdout=subprocess.PIPE, text=True, timeout=10)
        c = json.loads(r.stdout)
        assert sorted(c['Workers']) == ['127.0.0.1:100', '127.0.0.1:101', '127.0.0.2:100', '127.0.0.2:101']
        assert c['Runners'] == ['127.0.0.1:38080', '127.0.0.2:38080']
        seen.add(tuple(c['Workers']))
    assert len(seen) > 2  # orders differ between seeds


def test_elastic_hook_follows_config_server(tmp_path):
    """ElasticHook (reference: tensorflow/experimental/hook/elastic.py) under kungfu-run -w with a built-in config
    server: an operator PUTs 3 workers, then 1; the hook grows, shrinks (detached workers stop) and finishes."""
    base = _base(6)
    cfg_port = base + 60
    url = 'http://127.0.0.1:%d/config' % cfg_port
    worker = [sys.executable, os.path.join(ROOT, 'tests', 'workers', 'elastic_hook_worker.py'), '1200']
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get('PYTHONPATH', ''))
    run = subprocess.Popen([os.path.join(BIN, 'kungfu-run'), '-q', '-w', '-np', '1', '-H', '127.0.0.1:4', '-port',
                            str(base + 40), '-port-range', '%d-%d' % (base, base + 16), '-builtin-config-port',
                            str(cfg_port), '-config-server', url, '-timeout', '150s'] + worker,
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)

    def put(n):
        c = {'Runners': ['127.0.0.1:%d' % (base + 40)], 'Workers': ['127.0.0.1:%d' % (base + i) for i in range(n)]}
        f = tmp_path / ('c%d.json' % n)
        f.write_text(json.dumps(c))
        return subprocess.run([os.path.join(BIN, 'kungfu-test-util'), '-propose', str(f), '-server', url],
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=30)
    try:
        time.sleep(5)
        assert put(3).returncode == 0
        time.sleep(8)
        assert put(1).returncode == 0
        log, _ = run.communicate(timeout=170)
    finally:
        if run.poll() is None:
            run.kill()
    assert run.returncode == 0, log[-3000:]
    assert log.count('OK elastic hook: change cluster') == 2, log[-3000:]
    assert 'OK elastic hook: finished after' in log
    assert 'resize 1 -> 3' in log and 'resize 3 -> 1' in log


def test_config_server_survives_malformed_requests():
    """Garbage, truncated heads, lying or negative Content-Length, broken JSON: every one is refused or ignored and
    the server keeps serving."""
    import socket
    port = _base(7)
    url = 'http://127.0.0.1:%d/config' % port
    srv = subprocess.Popen([os.path.join(BIN, 'kungfu-config-server'), '-port', str(port)],
                           stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)

    def raw(data):
        s = socket.create_connection(('127.0.0.1', port), timeout=5)
        s.sendall(data)
        try:
            s.shutdown(socket.SHUT_WR)
            out = s.recv(4096)
        except OSError:
            out = b''
        s.close()
        return out
    try:
        for _ in range(50):
            try:
                _get(url)
                break
            except Exception:
                time.sleep(0.1)
        assert raw(b'\x00\x01\x02garbage\r\n\r\n').startswith(b'HTTP/1.1 404')
        assert raw(b'GET /config HTTP/1.1\r\n') == b''
        assert raw(b'PUT /config HTTP/1.1\r\nContent-Length: 999999999999\r\n\r\nabc') == b''
        assert raw(b'PUT /config HTTP/1.1\r\nContent-Length: -5\r\n\r\n{}') == b''
        assert raw(b'PUT /config HTTP/1.1\r\nContent-Length: 12\r\n\r\n{"Runners":1').startswith(b'HTTP/1.1 400')
        assert raw(b'A' * 200000) == b''
        assert srv.poll() is None
        good = json.dumps({'Runners': ['127.0.0.1:38080'], 'Workers': ['127.0.0.1:10000', '127.0.0.1:10001']}).encode()
        req = urllib.request.Request(url, data=good, method='PUT')
        assert urllib.request.urlopen(req, timeout=5).status == 200
        status, body = _get(url)
        assert status == 200 and len(json.loads(body)['Workers']) == 2
    finally:
        srv.kill()
        srv.wait()
