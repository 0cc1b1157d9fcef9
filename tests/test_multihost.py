"""Multi-host behaviour on one machine: three loopback aliases act as hosts,
each with its own kungfu-run (the way kungfu-rrun would start them)."""
import os
import subprocess
import sys
import time

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


@pytest.mark.parametrize('strategy', ['AUTO'])  # AUTO on several hosts is MULTI_BINARY_TREE_STAR
def test_sixteen_workers_on_four_hosts(strategy):
    """The reference's virtual cluster shape (4 nodes x 4 slots): every op of the multi-host worker at np = 16."""
    exe = os.path.join(ROOT, 'kungfu_b200', 'bin', 'kungfu-run')
    base = 22000 + (os.getpid() * 43) % 10000
    ips = ['127.0.0.%d' % (i + 1) for i in range(4)]
    hosts = ','.join(ip + ':4' for ip in ips)
    env = dict(os.environ, PYTHONPATH=ROOT)
    procs = []
    for ip in ips:
        cmd = [exe, '-q', '-np', '16', '-H', hosts, '-self', ip, '-strategy', strategy, '-timeout', '150s', '-port-range',
               '%d-%d' % (base, base + 16), '-port', str(base + 20), sys.executable,
               os.path.join(ROOT, 'tests', 'workers', 'multihost_worker.py')]
        procs.append(subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env))
    outs = [p.communicate(timeout=200)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), '\n'.join(o[-1500:] for o in outs)
    assert sum(o.count('OK multihost') for o in outs) == 16


FAKE_SSH = """#!/bin/sh
# stands in for the system ssh: drops the options and the target, runs the command on this machine
while [ "$1" = "-o" ]; do shift 2; done
echo "fake-ssh target=$1" >&2
shift
exec sh -c "$*"
"""


def test_rrun_starts_one_runner_per_host_through_ssh(tmp_path):
    """kungfu-rrun end to end (not -dry-run): the `ssh` it executes is a local
    stand-in, so the per-host command line it builds (-H, -self, forwarded
    flags, quoting of the program's arguments) is what starts the job.
    Reference: /root/reference/srcs/go/cmd/kungfu-rrun/rrun.go."""
    fake = tmp_path / 'ssh'
    fake.write_text(FAKE_SSH)
    fake.chmod(0o755)
    bindir = os.path.join(ROOT, 'kungfu_b200', 'bin')
    base = 20000 + (os.getpid() * 17) % 10000  # below the ephemeral range
    env = dict(os.environ, PYTHONPATH=ROOT, PATH='%s:%s:%s' % (tmp_path, bindir, os.environ.get('PATH', '')))
    cmd = [os.path.join(bindir, 'kungfu-rrun'), '-H', '127.0.0.1:2,127.0.0.2:1', '-u', 'nobody', '-q', '-np', '3', '-timeout', '90s',
           '-port-range', '%d-%d' % (base, base + 16), '-port', str(base + 20), sys.executable, '-c',
           "import kungfu_b200.torch as kf, torch; s = kf.ops.all_reduce_fn(torch.ones(3), 'sum'); "
           "print('OK rrun %d/%d sum=%d it\\'s quoted' % (kf.current_rank(), kf.current_cluster_size(), int(s[0])))"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env, timeout=150)
    assert r.returncode == 0, r.stdout[-3000:]
    assert r.stdout.count('fake-ssh target=nobody@127.0.0.') == 2, r.stdout[-3000:]
    for rank in range(3):
        assert "OK rrun %d/3 sum=3 it's quoted" % rank in r.stdout, r.stdout[-3000:]
    # output is prefixed with the host it came from
    assert '[127.0.0.1] ' in r.stdout and '[127.0.0.2] ' in r.stdout

    # a failing host fails the whole call and the other hosts are cancelled, as the reference does: whichever host
    # takes the lock sleeps for a minute, the other one exits 3 at once
    lock = tmp_path / 'lock'
    t0 = time.time()
    r = subprocess.run([os.path.join(bindir, 'kungfu-distribute'), '-H', '127.0.0.1:1,127.0.0.2:1', 'sh', '-c',
                        'mkdir %s 2>/dev/null && exec sleep 60 || exit 3' % lock],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env, timeout=90)
    assert r.returncode == 3 and 'exited with 3: stopping the other hosts' in r.stdout, r.stdout
    assert time.time() - t0 < 20, 'the sleeping host was not cancelled'


def test_experiment_tools_install_and_scale_through_ssh(tmp_path):
    """benchmarks/experiments.py --remote-install and --scaling over two loopback hosts, with the stand-in ssh:
    the packed tree is pushed and pip-installed per host, then a config server + kungfu-rrun job follows a
    resize schedule that spreads over both hosts (reference: experiments/cmd/kungfu-remote-install,
    kungfu-run-scaling-experiments)."""
    fake = tmp_path / 'ssh'
    fake.write_text(FAKE_SSH)
    fake.chmod(0o755)
    bindir = os.path.join(ROOT, 'kungfu_b200', 'bin')
    env = dict(os.environ, PYTHONPATH=ROOT, PATH='%s:%s:%s' % (tmp_path, bindir, os.environ.get('PATH', '')))
    exp = [sys.executable, os.path.join(ROOT, 'benchmarks', 'experiments.py')]
    # dry run of the install: one ssh per host, public address honoured
    r = subprocess.run(exp + ['--remote-install', '--hosts', '10.0.0.1:8,10.0.0.2:8:node2.example', '--user', 'me', '--dry-run'],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env, timeout=60)
    lines = r.stdout.strip().splitlines()
    assert r.returncode == 0 and len(lines) == 2 and 'me@10.0.0.1' in lines[0] and 'me@node2.example' in lines[1], r.stdout
    # the scaling experiment for real: 1 -> 3 -> 4 -> 2 workers over two hosts with two slots each
    base = 21000 + (os.getpid() * 13) % 9000  # below the ephemeral range (a port someone's outgoing connection holds cannot be bound)
    r = subprocess.run(exp + ['--scaling', '2:1,2:3,2:4,2:2', '--model-mb', '1', '--hosts', '127.0.0.1:2,127.0.0.2:2', '--config-port',
                              str(base + 90), '--port-range', '%d-%d' % (base, base + 16), '--runner-port', str(base + 80),
                              '--timeout', '200'], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env, timeout=240)
    assert r.returncode == 0, r.stdout[-4000:]
    for a, b in ((1, 3), (3, 4), (4, 2)):
        assert 'RESIZE %d -> %d took' % (a, b) in r.stdout, r.stdout[-4000:]
    assert r.stdout.count(' done') == 2 and r.stdout.count('detached at step') == 2, r.stdout[-4000:]
    # every worker of the second host listens inside the job's port range, not the default one
    assert '[127.0.0.2.%d::' % base in r.stdout and '127.0.0.2.10000' not in r.stdout, r.stdout[-4000:]
