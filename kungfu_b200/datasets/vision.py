"""Real-file loaders and device-side preprocessing for the image benchmarks.

Reference: /root/reference/srcs/python/kungfu/tensorflow/v1/helpers/cifar.py (python-pickle
CIFAR-10), imagenet.py (TFRecord pipeline: decode, distorted bounding-box crop, flip, colour
distortion, resize) and idx.py:83 (npz -> tar of idx files). There is no TensorFlow here and the
B200 has cycles to spare next to a 148-SM training step, so the augmentation runs ON THE GPU on
whole uint8 batches: the host only reads bytes and issues one pinned H2D copy per batch.
"""
import os
import pickle
import tarfile

import numpy as np
import torch
import torch.nn.functional as F

from . import synthetic_cifar10, write_idx

CIFAR_MEAN = (0.4914, 0.4822, 0.4465)
CIFAR_STD = (0.2470, 0.2435, 0.2616)
IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)


def random_input(sample_shape, n_classes, batch=1, seed=0):
    """A random (features, one-hot labels) pair of the given shape (helpers/__init__.py:random_input)."""
    g = torch.Generator().manual_seed(seed)
    x = torch.rand((batch, ) + tuple(sample_shape), generator=g)
    y = F.one_hot(torch.randint(0, n_classes, (batch, ), generator=g), n_classes).float()
    return x, y


def load_cifar10(data_dir=None, normalize=True, one_hot=False, n_synthetic=50000):
    """(x_train [N,3,32,32] float, y_train, x_test, y_test) from the python-version batches
    (``data_batch_1..5``, ``test_batch``, optionally inside ``cifar-10-batches-py``); a synthetic set of
    the same shapes when the files are absent (no network here)."""
    root = None
    for cand in ([data_dir, os.path.join(data_dir, 'cifar-10-batches-py')] if data_dir else []):
        if cand and os.path.exists(os.path.join(cand, 'data_batch_1')):
            root = cand
    if root is None:
        x, y = synthetic_cifar10(n_synthetic + 10000)
        xt, yt, xe, ye = x[:n_synthetic], y[:n_synthetic], x[n_synthetic:], y[n_synthetic:]
    else:
        def read(name):
            with open(os.path.join(root, name), 'rb') as f:
                d = pickle.load(f, encoding='bytes')
            return np.asarray(d[b'data'], dtype=np.uint8).reshape(-1, 3, 32, 32), np.asarray(d[b'labels'], dtype=np.int64)

        parts = [read('data_batch_%d' % i) for i in range(1, 6)]
        xt = torch.from_numpy(np.concatenate([p[0] for p in parts])).float() / 255
        yt = torch.from_numpy(np.concatenate([p[1] for p in parts]))
        xe, ye = read('test_batch')
        xe, ye = torch.from_numpy(xe).float() / 255, torch.from_numpy(ye)
        if normalize:
            m, s = torch.tensor(CIFAR_MEAN).view(1, 3, 1, 1), torch.tensor(CIFAR_STD).view(1, 3, 1, 1)
            xt, xe = (xt - m) / s, (xe - m) / s
    if one_hot:
        yt, ye = F.one_hot(yt, 10).float(), F.one_hot(ye, 10).float()
    return xt, yt, xe, ye


def npz_to_idx_tar(in_file, out_file=None):
    """Repack every array of an .npz as ``<name>.idx`` inside a tar (helpers/idx.py:83 npz2idxtar)."""
    out_file = out_file or os.path.splitext(in_file)[0] + '.idx.tar'
    data = np.load(in_file)
    tmp = out_file + '.member'
    with tarfile.open(out_file, 'w') as tar:
        for name in data.files:
            arr = data[name]
            if arr.dtype not in (np.uint8, np.int32, np.float32):
                arr = arr.astype(np.float32 if arr.dtype.kind == 'f' else np.int32)
            write_idx(tmp, arr)
            tar.add(tmp, arcname=name + '.idx')
    if os.path.exists(tmp):
        os.remove(tmp)
    return out_file


class ImageFolder(object):
    """``root/<class name>/*.{jpg,jpeg,png}`` -> (uint8 [3,S,S] tensor, class index). Images are decoded and
    resized on the host to a fixed square (so a batch is one dense uint8 tensor); everything random happens
    later on the device (`gpu_augment`)."""
    EXT = ('.jpg', '.jpeg', '.png', '.bmp')

    def __init__(self, root, load_size=256):
        self.load_size = load_size
        self.classes = sorted(d for d in os.listdir(root) if os.path.isdir(os.path.join(root, d)))
        self.samples = []
        for ci, c in enumerate(self.classes):
            for f in sorted(os.listdir(os.path.join(root, c))):
                if f.lower().endswith(self.EXT):
                    self.samples.append((os.path.join(root, c, f), ci))

    def __len__(self):
        return len(self.samples)

    def __getitem__(self, i):
        from PIL import Image
        path, label = self.samples[i]
        with Image.open(path) as im:
            im = im.convert('RGB').resize((self.load_size, self.load_size), Image.BILINEAR)
            arr = np.asarray(im, dtype=np.uint8)
        return torch.from_numpy(arr.copy()).permute(2, 0, 1), label

    def batch(self, indices, pin=True):
        xs, ys = zip(*(self[i] for i in indices))
        x, y = torch.stack(xs), torch.tensor(ys)
        if pin and torch.cuda.is_available():
            x, y = x.pin_memory(), y.pin_memory()
        return x, y


def gpu_augment(x_u8, train=True, size=224, scale=(0.35, 1.0), jitter=0.2, mean=IMAGENET_MEAN, std=IMAGENET_STD,
                dtype=torch.bfloat16, generator=None):
    """uint8 [B,3,H,W] (any device) -> normalised channels-last [B,3,size,size] of `dtype`.

    train: per-image random-area crop + resize, horizontal flip, brightness / contrast / saturation jitter
    (the reference's distorted-bounding-box crop and colour distortion, imagenet.py:43-140);
    eval: centre crop of 87.5 % + resize. All of it is batched tensor math on the input's device."""
    B, C, H, W = x_u8.shape
    dev = x_u8.device
    x = x_u8.float().div_(255)
    rnd = lambda *s: torch.rand(*s, device=dev, generator=generator)  # noqa: E731
    if train:
        area = scale[0] + (scale[1] - scale[0]) * rnd(B)
        side = area.sqrt()                                  # crop side as a fraction of the image
        cx = (1 - side) * rnd(B) + side / 2
        cy = (1 - side) * rnd(B) + side / 2
        flip = torch.where(rnd(B) < 0.5, -1.0, 1.0)
        theta = torch.zeros(B, 2, 3, device=dev)
        theta[:, 0, 0] = side * flip
        theta[:, 0, 2] = 2 * cx - 1
        theta[:, 1, 1] = side
        theta[:, 1, 2] = 2 * cy - 1
    else:
        theta = torch.zeros(B, 2, 3, device=dev)
        theta[:, 0, 0] = theta[:, 1, 1] = 0.875
    grid = F.affine_grid(theta, (B, C, size, size), align_corners=False)
    x = F.grid_sample(x, grid, mode='bilinear', padding_mode='border', align_corners=False)
    if train and jitter > 0:
        b = 1 + jitter * (2 * rnd(B, 1, 1, 1) - 1)
        c = 1 + jitter * (2 * rnd(B, 1, 1, 1) - 1)
        s = 1 + jitter * (2 * rnd(B, 1, 1, 1) - 1)
        x = x * b
        grey = x.mean(dim=1, keepdim=True)
        x = (x - grey) * s + grey
        m = grey.mean(dim=(2, 3), keepdim=True)
        x = ((x - m) * c + m).clamp_(0, 1)
    mean_t = torch.tensor(mean, device=dev).view(1, 3, 1, 1)
    std_t = torch.tensor(std, device=dev).view(1, 3, 1, 1)
    x = (x - mean_t) / std_t
    return x.to(dtype).contiguous(memory_format=torch.channels_last)
