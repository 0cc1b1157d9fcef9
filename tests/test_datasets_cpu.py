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


def test_elastic_batch_sampler_covers_every_sample_once_across_resizes():
    """Cluster sizes 2 -> 3 -> 1 -> 4 in the middle of epochs: the union over the workers of every step is one
    contiguous run of the epoch's permutation, an epoch (whole batches) touches each of its samples exactly once,
    and the permutation changes from epoch to epoch but is the same on every worker."""
    import torch
    from torch.utils.data import DataLoader, TensorDataset

    from kungfu_b200.datasets import ElasticBatchSampler
    n, gb = 1000, 24
    plan = [(2, 7), (3, 9), (1, 11), (4, 60)]  # (cluster size, steps at that size)
    consumed, seen = 0, []
    for size, steps in plan:
        workers = [ElasticBatchSampler(n, gb, r, size, shuffle=True, seed=5, consumed=consumed) for r in range(size)]
        for _ in range(steps):
            step = [w.next_indices() for w in workers]
            assert all(len(s) == gb // size for s in step)
            seen.append([i for s in step for i in s])
        consumed = workers[0].consumed
        assert all(w.consumed == consumed for w in workers)
    per_epoch = n - n % gb  # 984: whole global batches only
    flat = [i for s in seen for i in s]
    e0, e1 = flat[:per_epoch], flat[per_epoch:2 * per_epoch]
    assert len(set(e0)) == per_epoch and len(set(e1)) == per_epoch  # nothing twice within an epoch
    assert e0 != e1                                                  # reshuffled
    ref = ElasticBatchSampler(n, gb, 0, 1, shuffle=True, seed=5)
    assert [i for _ in range(len(seen)) for i in ref.next_indices()] == flat  # == what one worker alone would have seen

    # as a DataLoader batch_sampler: one epoch from the current position, then the next one
    ds = TensorDataset(torch.arange(100))
    s = ElasticBatchSampler(100, 10, rank=1, size=2)
    got = [b[0].tolist() for b in DataLoader(ds, batch_sampler=s)]
    assert len(got) == len(s) == 10 and got[0] == [5, 6, 7, 8, 9] and got[-1] == [95, 96, 97, 98, 99]
    s.reshard(0, 5)
    assert next(iter(DataLoader(ds, batch_sampler=s)))[0].tolist() == [0, 1]
    t = ElasticBatchSampler(100, 10)
    t.load_state_dict(s.state_dict())
    assert t.consumed == s.consumed == 110
    import pytest
    with pytest.raises(ValueError):
        s.reshard(0, 3)


def test_elastic_batch_sampler_random_resize_sequences():
    """Property: for ANY sequence of cluster sizes dividing the global batch, the concatenated per-step unions equal
    what a single worker would have drawn, epoch after epoch, shuffled or not."""
    from hypothesis import given, settings
    from hypothesis import strategies as st

    from kungfu_b200.datasets import ElasticBatchSampler

    @settings(max_examples=60, deadline=None)
    @given(st.integers(30, 400), st.sampled_from([12, 24, 60]), st.booleans(), st.booleans(),
           st.lists(st.tuples(st.sampled_from([1, 2, 3, 4, 6, 12]), st.integers(1, 9)), min_size=1, max_size=8))
    def prop(n, gb, shuffle, drop_last, plan):
        consumed, seen = 0, []
        for size, steps in plan:
            ws = [ElasticBatchSampler(n, gb, r, size, shuffle=shuffle, seed=3, consumed=consumed, drop_last=drop_last) for r in range(size)]
            for _ in range(steps):
                seen += [i for w in ws for i in w.next_indices()]
            consumed = ws[0].consumed
        ref = ElasticBatchSampler(n, gb, 0, 1, shuffle=shuffle, seed=3, drop_last=drop_last)
        want = [i for _ in range(len(seen) // gb) for i in ref.next_indices()]
        assert seen == want
        assert all(0 <= i < n for i in seen)
        per_epoch = max(n - n % gb, gb) if drop_last else n
        if n >= gb:
            first = seen[:per_epoch]
            assert len(set(first)) == len(first)  # an epoch never repeats a sample

    prop()
