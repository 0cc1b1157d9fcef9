"""Elastic resize on loopback: kungfu-run -w with the builtin config server,
sizes 1 -> 2 -> 4 -> 2 -> 1 (the reference's bench-adaptation / step-based
schedule tests, /root/reference/tests/python/integration/test_step_based_schedule.py:7-13,
/root/reference/benchmarks/adaptation/bench-adaptation.sh)."""
import collections
import os
import subprocess
import sys
import time
import urllib.request

import pytest

from conftest import ROOT, run_workers

W = os.path.join(ROOT, 'tests', 'workers')


def _free_port():
    import socket
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize('schedule,steps', [('1:3,2:3,4:3,2:3,1:3', 15), ('2:2,1:2,3:2', 6)])
def test_resize_schedule(tmp_path, schedule, steps):
    log = str(tmp_path / 'steps.log')
    np0 = int(schedule.split(':')[0])
    cfg_port = _free_port()
    res = run_workers(np0, os.path.join(W, 'elastic_worker.py'), args=[schedule, str(steps), log], timeout=150,
                      extra_flags=['-w', '-builtin-config-port', str(cfg_port)])
    assert res.returncode == 0, res.stdout[-6000:]
    sizes = collections.defaultdict(set)
    pids = collections.defaultdict(set)
    for line in open(log):
        if line.startswith('detached'):
            continue
        step, rank, np_, pid = map(int, line.split())
        sizes[step].add(np_)
        pids[step].add(pid)
    off = 0
    for part in schedule.split(','):
        size, n = map(int, part.split(':'))
        for s in range(off, off + n):
            assert sizes[s] == {size}, (s, sizes[s], schedule)
            assert len(pids[s]) == size
        off += n
    # the worker that started as rank 0 survives every resize: it was never restarted
    first = next(iter(pids[0])) if np0 == 1 else None
    if first is not None:
        assert all(first in pids[s] for s in range(steps))


def test_config_server_rest_api(tmp_path):
    port = _free_port()
    exe = os.path.join(ROOT, 'kungfu_b200', 'bin', 'kungfu-config-server')
    p = subprocess.Popen([exe, '-port', str(port)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    url = 'http://127.0.0.1:%d/config' % port
    try:
        for _ in range(50):
            try:
                urllib.request.urlopen(url)
            except urllib.error.HTTPError as e:
                assert e.code == 404  # nothing stored yet
                break
            except Exception:
                time.sleep(0.1)
        good = b'{"Runners":["127.0.0.1:38080"],"Workers":["127.0.0.1:10000","127.0.0.1:10001"]}'
        req = urllib.request.Request(url, data=good, method='PUT')
        assert urllib.request.urlopen(req).status == 200
        got = urllib.request.urlopen(url).read().decode()
        assert '"127.0.0.1:10001"' in got
        bad = b'{"Runners":[],"Workers":["127.0.0.1:10000"]}'  # worker without a runner on its host
        with pytest.raises(urllib.error.HTTPError):
            urllib.request.urlopen(urllib.request.Request(url, data=bad, method='PUT'))
        # clearing to zero workers freezes the config (PUT -> 403)
        empty = b'{"Runners":["127.0.0.1:38080"],"Workers":[]}'
        assert urllib.request.urlopen(urllib.request.Request(url, data=empty, method='PUT')).status == 200
        with pytest.raises(urllib.error.HTTPError) as ei:
            urllib.request.urlopen(urllib.request.Request(url, data=good, method='PUT'))
        assert ei.value.code == 403
        # POST resets
        assert urllib.request.urlopen(urllib.request.Request(url, data=b'', method='POST')).status == 200
        with pytest.raises(urllib.error.HTTPError) as ei:
            urllib.request.urlopen(url)
        assert ei.value.code == 404
        urllib.request.urlopen('http://127.0.0.1:%d/stop' % port)
        assert p.wait(timeout=10) == 0
    finally:
        if p.poll() is None:
            p.kill()


def test_config_server_init_file_endpoint_and_ttl(tmp_path):
    """-init publishes a cluster from a file, -endpoint moves the REST path (the reference's examples use /get and
    /put), -ttl makes the server leave by itself (reference: cmd/kungfu-config-server/kungfu-config-server.go:19-65)."""
    port = _free_port()
    init = tmp_path / 'cluster.json'
    init.write_text('{"Runners":["127.0.0.1:38080"],"Workers":["127.0.0.1:10000","127.0.0.1:10001","127.0.0.1:10002"]}')
    exe = os.path.join(ROOT, 'kungfu_b200', 'bin', 'kungfu-config-server')
    t0 = time.time()
    p = subprocess.Popen([exe, '-port', str(port), '-init', str(init), '-endpoint', '/get', '-ttl', '2s'],
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    try:
        got = None
        for _ in range(50):
            try:
                got = urllib.request.urlopen('http://127.0.0.1:%d/get' % port).read().decode()
                break
            except urllib.error.HTTPError:
                raise
            except Exception:
                time.sleep(0.05)
        assert got is not None and got.count('127.0.0.1:1000') == 3, got
        with pytest.raises(urllib.error.HTTPError) as ei:  # the default path is not served when -endpoint moved it
            urllib.request.urlopen('http://127.0.0.1:%d/config' % port)
        assert ei.value.code == 404
        assert p.wait(timeout=10) == 0 and 1.5 < time.time() - t0 < 9
    finally:
        if p.poll() is None:
            p.kill()
    bad = tmp_path / 'bad.json'
    bad.write_text('{"Runners":[],"Workers":["127.0.0.1:10000"]}')
    r = subprocess.run([exe, '-port', str(_free_port()), '-init', str(bad)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=10)
    assert r.returncode != 0 and 'bad -init' in r.stdout


def test_regrow_onto_a_used_address_restarts_the_control_exchange():
    """1 -> 2 -> 1 -> 2 -> 1 -> 2: same size, same rank, new peer each time (found on GPUs, where the second
    grow hung in PeerComm's first all-gather because the survivor kept counting message names)."""
    res = run_workers(1, os.path.join(W, 'regrow_worker.py'), args=['1:1,2:1,1:1,2:1,1:1,2:1,1:1', '7'], timeout=120,
                      extra_flags=['-w', '-builtin-config-port', str(_free_port())])
    assert res.returncode == 0, res.stdout[-4000:]
    assert 'OK regrow worker rank 0 done at step 7' in res.stdout


def test_resizes_leave_no_descriptors_threads_or_memory_behind():
    """Five grow / shrink cycles 1 -> 3 -> 1: the survivor's open descriptors and threads at size 1 do not grow
    from the early visits to the late ones, and its resident memory does not creep (25 cycles by hand: 9
    descriptors, 24 threads, 524 MiB throughout)."""
    import re
    cycles = 5
    schedule = ','.join(['1:1', '3:1'] * cycles + ['1:1'])
    res = run_workers(1, os.path.join(W, 'leak_worker.py'), args=[schedule, str(2 * cycles + 1)], timeout=240,
                      extra_flags=['-w', '-builtin-config-port', str(_free_port())])
    assert res.returncode == 0 and 'OK leak worker done' in res.stdout, res.stdout[-4000:]
    rows = [tuple(map(int, m.groups())) for m in re.finditer(r'LEAK step (\d+) fds (\d+) threads (\d+) rss_mib (\d+)', res.stdout)]
    assert len(rows) == cycles + 1, rows
    # (the worker reports the floor over one second: connections of the peers that just left are still closing)
    early, late = rows[:3], rows[3:]
    assert max(r[1] for r in late) <= max(r[1] for r in early), rows
    assert max(r[2] for r in late) <= max(r[2] for r in early), rows
    assert rows[-1][3] - rows[2][3] <= 16, rows
