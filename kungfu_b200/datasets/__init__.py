"""Data helpers: synthetic datasets of the shapes the reference's loaders
return (there is no network for the real files), an idx-format reader for MNIST
files when present, and the elastic dataset adaptor.
Reference: /root/reference/srcs/python/kungfu/tensorflow/v1/helpers/{mnist.py,cifar.py,
imagenet.py,idx.py}, /root/reference/srcs/python/kungfu/tensorflow/v1/datasets/adaptor.py:4-45."""
import gzip
import os
import struct

import numpy as np
import torch


def read_idx(path):
    """Read an idx file (optionally .gz): the MNIST on-disk format."""
    op = gzip.open if path.endswith('.gz') else open
    with op(path, 'rb') as f:
        magic = struct.unpack('>I', f.read(4))[0]
        dtype = {0x08: np.uint8, 0x09: np.int8, 0x0B: np.int16, 0x0C: np.int32, 0x0D: np.float32, 0x0E: np.float64}[(magic >> 8) & 0xff]
        dims = struct.unpack('>' + 'I' * (magic & 0xff), f.read(4 * (magic & 0xff)))
        return np.frombuffer(f.read(), dtype=np.dtype(dtype).newbyteorder('>')).reshape(dims)


def write_idx(path, arr):
    arr = np.asarray(arr)
    code = {np.dtype(np.uint8): 0x08, np.dtype(np.int32): 0x0C, np.dtype(np.float32): 0x0D}[arr.dtype]
    with open(path, 'wb') as f:
        f.write(struct.pack('>I', (code << 8) | arr.ndim))
        f.write(struct.pack('>' + 'I' * arr.ndim, *arr.shape))
        f.write(arr.astype(arr.dtype.newbyteorder('>')).tobytes())


def load_mnist(data_dir=None, n_train=60000, n_test=10000, seed=2020):
    """MNIST as (x_train, y_train, x_test, y_test); real idx files if found
    under data_dir, else a deterministic synthetic set of the same shape whose
    labels are a fixed random linear function of the pixels (learnable)."""
    names = ['train-images-idx3-ubyte', 'train-labels-idx1-ubyte', 't10k-images-idx3-ubyte', 't10k-labels-idx1-ubyte']
    if data_dir and all(os.path.exists(os.path.join(data_dir, n)) or os.path.exists(os.path.join(data_dir, n + '.gz')) for n in names):
        arrs = []
        for n in names:
            p = os.path.join(data_dir, n)
            arrs.append(read_idx(p if os.path.exists(p) else p + '.gz'))
        xt, yt, xe, ye = arrs
        return (torch.from_numpy(xt.astype(np.float32) / 255).view(-1, 1, 28, 28), torch.from_numpy(yt.astype(np.int64)),
                torch.from_numpy(xe.astype(np.float32) / 255).view(-1, 1, 28, 28), torch.from_numpy(ye.astype(np.int64)))
    g = torch.Generator().manual_seed(seed)
    w = torch.randn(784, 10, generator=g)
    x = torch.rand(n_train + n_test, 784, generator=g)
    y = (x @ w).argmax(1)
    x = x.view(-1, 1, 28, 28)
    return x[:n_train], y[:n_train], x[n_train:], y[n_train:]


def synthetic_images(n, size=224, classes=1000, seed=0, dtype=torch.float32):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(n, 3, size, size, generator=g).to(dtype), torch.randint(0, classes, (n, ), generator=g)


def synthetic_cifar10(n=50000, seed=0):
    return synthetic_images(n, 32, 10, seed)


class ElasticShardedDataset(object):
    """Batches of a dataset for THIS worker under a changing cluster size.

    State is three integers -- global sample offset, global batch size,
    (rank, size) -- exactly what the reference's adaptor keeps in tf.Variables:
    after a resize call ``reshard()`` and the stream continues from the same
    global offset with the new sharding, so no sample is skipped or repeated."""

    def __init__(self, xs, ys, global_batch_size, rank=0, size=1, offset=0):
        self.xs, self.ys = xs, ys
        self.global_batch_size = global_batch_size
        self.offset = offset
        self.reshard(rank, size)

    def reshard(self, rank, size, global_batch_size=None, offset=None):
        self.rank, self.size = rank, size
        if global_batch_size is not None:
            self.global_batch_size = global_batch_size
        if offset is not None:
            self.offset = offset
        if self.global_batch_size % size:
            raise ValueError('global batch %d is not divisible by %d workers' % (self.global_batch_size, size))

    def next_batch(self):
        n = len(self.xs)
        local = self.global_batch_size // self.size
        idx = [(self.offset + self.rank * local + i) % n for i in range(local)]
        self.offset += self.global_batch_size
        return self.xs[idx], self.ys[idx]


class ElasticBatchSampler(object):
    """``batch_sampler`` for ``torch.utils.data.DataLoader`` under a changing cluster size.

    Every step consumes ``global_batch_size`` consecutive positions of a per-epoch permutation that all
    workers derive from ``(seed, epoch)``; this worker takes its slice of them. The whole state is the
    number of samples consumed so far (``consumed``), which ``ElasticState`` already keeps in step with
    the cluster: after a resize call ``reshard(rank, size)`` (optionally with the synced ``consumed``)
    and the stream goes on where the cluster as a whole stopped -- nothing is skipped or seen twice,
    whatever the sizes before and after. The reference keeps the same three integers in tf.Variables
    (/root/reference/srcs/python/kungfu/tensorflow/v1/datasets/adaptor.py:4-45).

        sampler = ElasticBatchSampler(len(ds), 256, kf.current_rank(), kf.current_cluster_size(), shuffle=True)
        loader = DataLoader(ds, batch_sampler=sampler, num_workers=4)
    """

    def __init__(self, n_samples, global_batch_size, rank=0, size=1, shuffle=False, seed=0, consumed=0, drop_last=True):
        self.n = int(n_samples)
        self.global_batch_size = int(global_batch_size)
        self.shuffle, self.seed, self.drop_last = shuffle, seed, drop_last
        self.consumed = int(consumed)
        self._perm_epoch, self._perm = None, None
        self.reshard(rank, size)

    def reshard(self, rank, size, consumed=None, global_batch_size=None):
        if global_batch_size is not None:
            self.global_batch_size = int(global_batch_size)
        if self.global_batch_size % size:
            raise ValueError('global batch %d is not divisible by %d workers' % (self.global_batch_size, size))
        self.rank, self.size = int(rank), int(size)
        if consumed is not None:
            self.consumed = int(consumed)

    @property
    def epoch(self):
        return self.consumed // self._epoch_len()

    def _epoch_len(self):
        # samples handed out per epoch: whole global batches only (drop_last), else the tail wraps into the next epoch
        if self.drop_last:
            return max(self.n - self.n % self.global_batch_size, self.global_batch_size)
        return self.n

    def _order(self, epoch):
        if not self.shuffle:
            return None
        if self._perm_epoch != epoch:
            g = torch.Generator().manual_seed(self.seed * 1000003 + epoch)
            self._perm, self._perm_epoch = torch.randperm(self.n, generator=g).tolist(), epoch
        return self._perm

    def next_indices(self):
        """This worker's indices for the next step; advances ``consumed`` by one global batch."""
        local = self.global_batch_size // self.size
        out = []
        for i in range(local):
            g = self.consumed + self.rank * local + i
            epoch, pos = divmod(g, self._epoch_len())
            order = self._order(epoch)
            out.append(order[pos % self.n] if order is not None else pos % self.n)
        self.consumed += self.global_batch_size
        return out

    def __iter__(self):
        # one epoch's worth of steps from wherever the cluster is now
        epoch = self.epoch
        while self.epoch == epoch:
            yield self.next_indices()

    def __len__(self):
        return self._epoch_len() // self.global_batch_size

    def state_dict(self):
        return {'consumed': self.consumed, 'global_batch_size': self.global_batch_size, 'seed': self.seed}

    def load_state_dict(self, state):
        self.consumed, self.global_batch_size, self.seed = state['consumed'], state['global_batch_size'], state['seed']
        self._perm_epoch = None


from .vision import (ImageFolder, gpu_augment, load_cifar10, npz_to_idx_tar, random_input)  # noqa: E402,F401


# ---- the reference's names for the same things ---------------------------------------------------
read_idx_file, write_idx_file = read_idx, write_idx  # helpers/idx.py:49,70
npz2idxtar = npz_to_idx_tar  # helpers/idx.py:83
BaseDatasetAdaptor = ExampleDatasetAdaptor = ElasticShardedDataset  # v1/datasets/adaptor.py:4-45


def load_datasets(data_dir=None, normalize=False, one_hot=False, padded=False):
    """MNIST as an object with .train / .test, each with .images / .labels (helpers/mnist.py:41)."""
    import collections
    import torch.nn.functional as F
    xt, yt, xe, ye = load_mnist(data_dir)
    if not normalize:
        xt, xe = xt * 255, xe * 255
    if padded:
        xt, xe = F.pad(xt, (2, 2, 2, 2)), F.pad(xe, (2, 2, 2, 2))
    if one_hot:
        yt, ye = F.one_hot(yt, 10).float(), F.one_hot(ye, 10).float()
    Split = collections.namedtuple('Split', 'images labels')
    Sets = collections.namedtuple('Datasets', 'train test')
    return Sets(Split(xt, yt), Split(xe, ye))


def load_mnist_data(data_dir, prefix, normalize, one_hot, padded=False):
    """One split ('train' or 't10k') as (images, labels) (helpers/mnist.py:18)."""
    sets = load_datasets(data_dir, normalize, one_hot, padded)
    return tuple(sets.train if prefix == 'train' else sets.test)


class Cifar10Loader(object):
    """helpers/cifar.py: ``Cifar10Loader(data_dir).load_datasets()`` -> .train / .test splits."""
    classes = 10

    def __init__(self, data_dir=None, normalize=True, one_hot=False):
        self.data_dir, self.normalize, self.one_hot = data_dir, normalize, one_hot

    def load_datasets(self):
        import collections
        xt, yt, xe, ye = load_cifar10(self.data_dir, self.normalize, self.one_hot)
        Split = collections.namedtuple('Split', 'images labels')
        return collections.namedtuple('Datasets', 'train test')(Split(xt, yt), Split(xe, ye))


class Cifar100Loader(Cifar10Loader):
    """CIFAR-100 has the same container format with 'fine_labels'; without the files (no network) the
    synthetic stand-in has 100 classes."""
    classes = 100

    def load_datasets(self):
        import collections
        x, y = synthetic_images(60000, 32, 100, seed=1)
        Split = collections.namedtuple('Split', 'images labels')
        return collections.namedtuple('Datasets', 'train test')(Split(x[:50000], y[:50000]), Split(x[50000:], y[50000:]))


def create_dataset(data_dir, batch_size=32, n=1, size=224, train=True):
    """ImageNet-style folder -> iterator of augmented device batches (helpers/imagenet.py:159 built a
    TFRecord pipeline; here: ImageFolder on the host + gpu_augment on the device)."""
    import torch
    ds = ImageFolder(data_dir)
    dev = torch.device('cuda') if torch.cuda.is_available() else torch.device('cpu')
    for start in range(0, len(ds) - batch_size + 1, batch_size * n):
        x, y = ds.batch(range(start, start + batch_size))
        yield gpu_augment(x.to(dev, non_blocking=True), train=train, size=size,
                          dtype=torch.bfloat16 if dev.type == 'cuda' else torch.float32), y.to(dev)
