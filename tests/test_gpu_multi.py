"""Real multi-GPU runs of the peer-memory path (skipped with < 2 GPUs)."""
import os

import pytest
import torch

from conftest import ROOT, run_workers

pytestmark = pytest.mark.gpu


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.parametrize('np_', [2, 4, 8])
def test_peer_memory_path(np_):
    if _ngpu() < np_:
        pytest.skip('needs %d GPUs' % np_)
    res = run_workers(np_, os.path.join(ROOT, 'tests', 'workers', 'gpu_peer_worker.py'), timeout=300,
                      extra_flags=['-allow-nvlink'])
    assert res.returncode == 0, res.stdout[-6000:]
    assert res.stdout.count('OK rank') == np_, res.stdout[-6000:]


def test_peer_memory_path_ipc_fallback():
    if _ngpu() < 2:
        pytest.skip('needs 2 GPUs')
    res = run_workers(2, os.path.join(ROOT, 'tests', 'workers', 'gpu_peer_worker.py'), timeout=300,
                      extra_flags=['-allow-nvlink'], env={'KUNGFU_SYMM_MODE': 'ipc'})
    assert res.returncode == 0, res.stdout[-6000:]
