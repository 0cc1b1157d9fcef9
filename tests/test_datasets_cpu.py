"""Dataset helpers: CIFAR pickle loader, idx tar packer, image folder + batched augmentation."""
import os
import pickle
import tarfile

import numpy as np
import torch


def test_load_cifar10_reads_python_batches_and_falls_back(tmp_path):
    from kungfu_b200.datasets import load_cifar10
    d = tmp_path / 'cifar-10-batches-py'
    d.mkdir()
    rng = np.random.RandomState(0)
    for name, n in [('data_batch_%d' % i, 20) for i in range(1, 6)] + [('test_batch', 10)]:
        with open(d / name, 'wb') as f:
            pickle.dump({b'data': rng.randint(0, 256, (n, 3072), dtype=np.uint8), b'labels': list(rng.randint(0, 10, n))}, f)
    xt, yt, xe, ye = load_cifar10(str(tmp_path), normalize=False, one_hot=True)
    assert xt.shape == (100, 3, 32, 32) and xe.shape == (10, 3, 32, 32) and yt.shape == (100, 10)
    assert 0 <= float(xt.min()) and float(xt.max()) <= 1
    xt2, _, _, _ = load_cifar10(str(tmp_path), normalize=True)
    assert abs(float(xt2.mean())) < 0.5 and float(xt2.min()) < 0
    xs, ys, xes, yes = load_cifar10(None, n_synthetic=64)
    assert xs.shape == (64, 3, 32, 32) and ys.shape == (64, ) and xes.shape[0] == 10000


def test_npz_to_idx_tar_roundtrip(tmp_path):
    from kungfu_b200.datasets import npz_to_idx_tar, read_idx
    a = np.arange(24, dtype=np.uint8).reshape(2, 3, 4)
    b = np.linspace(0, 1, 7, dtype=np.float64)
    np.savez(tmp_path / 'd.npz', images=a, weights=b)
    out = npz_to_idx_tar(str(tmp_path / 'd.npz'))
    with tarfile.open(out) as tar:
        assert sorted(tar.getnames()) == ['images.idx', 'weights.idx']
        tar.extractall(tmp_path / 'x')
    assert np.array_equal(read_idx(str(tmp_path / 'x' / 'images.idx')), a)
    assert np.allclose(read_idx(str(tmp_path / 'x' / 'weights.idx')), b.astype(np.float32))


def test_image_folder_and_batched_augmentation(tmp_path):
    from PIL import Image
    from kungfu_b200.datasets import ImageFolder, gpu_augment, random_input
    for c in ('cat', 'dog'):
        os.makedirs(tmp_path / c)
        for i in range(3):
            arr = np.full((40 + i, 50, 3), 60 if c == 'cat' else 200, dtype=np.uint8)
            Image.fromarray(arr).save(tmp_path / c / ('%d.png' % i))
    ds = ImageFolder(str(tmp_path), load_size=64)
    assert ds.classes == ['cat', 'dog'] and len(ds) == 6
    x, y = ds.batch(range(6), pin=False)
    assert x.shape == (6, 3, 64, 64) and x.dtype == torch.uint8 and y.tolist() == [0, 0, 0, 1, 1, 1]
    g = torch.Generator().manual_seed(0)
    out = gpu_augment(x, train=True, size=32, dtype=torch.float32, generator=g)
    assert out.shape == (6, 3, 32, 32) and out.is_contiguous(memory_format=torch.channels_last)
    assert torch.isfinite(out).all()
    # constant images stay (nearly) constant per image, and dogs stay brighter than cats
    assert float(out[3:].mean()) > float(out[:3].mean())
    ev = gpu_augment(x, train=False, size=32, dtype=torch.float32)
    assert torch.allclose(ev[0], ev[0].flatten()[0].expand_as(ev[0]) * 0 + ev[0], atol=0)
    assert float((ev[0, 0] - ev[0, 0, 0, 0]).abs().max()) < 1e-4
    xs, ys = random_input((3, 8, 8), 5, batch=4)
    assert xs.shape == (4, 3, 8, 8) and ys.shape == (4, 5) and float(ys.sum()) == 4
