"""The examples run as documented (small settings)."""
import os
import subprocess
import sys

from conftest import ROOT

RUN = os.path.join(ROOT, 'kungfu_b200', 'bin', 'kungfu-run')


def _base(salt):
    return 21000 + (os.getpid() * 67 + salt * 991) % 11000


def test_rl_policy_gradient_two_workers(tmp_path):
    base = _base(1)
    ckpt = str(tmp_path / 'pg.npz')
    one = subprocess.run([sys.executable, os.path.join(ROOT, 'examples', 'rl_policy_gradient.py'), '--episodes', '3', '--batch', '4',
                          '--save', ckpt], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=200)
    assert one.returncode == 0 and os.path.exists(ckpt), one.stdout[-2000:]
    res = subprocess.run([RUN, '-q', '-np', '2', '-port-range', '%d-%d' % (base, base + 16), '-port', str(base + 40),
                          sys.executable, os.path.join(ROOT, 'examples', 'rl_policy_gradient.py'), '--episodes', '6', '--batch', '4',
                          '--init', ckpt], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=200)
    assert res.returncode == 0, res.stdout[-2000:]
    assert 'np 2' in res.stdout and 'replicas identical: True' in res.stdout


def test_mnist_elastic_eager_follows_its_schedule():
    base = _base(2)
    res = subprocess.run([RUN, '-q', '-w', '-np', '1', '-H', '127.0.0.1:4', '-port-range', '%d-%d' % (base, base + 16),
                          '-port', str(base + 40), '-timeout', '150s', sys.executable,
                          os.path.join(ROOT, 'examples', 'mnist_elastic_eager.py')],
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=200)
    assert res.returncode == 0, res.stdout[-3000:]
    for n in (2, 4, 3, 1):
        assert 'cluster is now %d workers' % n in res.stdout
    assert res.stdout.count('detached at step') == 3
    assert 'finished at step 40 with np=1' in res.stdout
