"""Datasets: MNIST / CIFAR-10 readers without TensorFlow, synthetic data of the same shapes, ``next_batch`` API.

Reference behaviour kept (pytorch_code/mnist/mnist.py:38-131,157-211; cifar10/cifar10.py:167-199,394-407;
datasets/__init__.py:1-51): ``read_data_sets`` returns ``Datasets(train, validation, test)``; ``DataSet.next_batch``
walks the (shuffled) data and reshuffles at every epoch end; MNIST pixels are scaled to [-0.5, 0.5] and reshaped to
(N,1,28,28); CIFAR-10 stays raw 0..255 float32 (the reference applies no normalisation), (N,3,32,32); the 10k test
set doubles as "validation".

Deliberate differences (SURVEY §2.11-6,12): the shuffle is seeded per worker (``seed + worker_id``) instead of the
constant 66478 that made every worker train on identical batches; CIFAR pickles are transposed (NHWC->NCHW) instead
of raw-reshaped; nothing is downloaded (no network here): files are read from ``train_dir`` if present, otherwise a
deterministic synthetic set of the right shape is produced (``synthetic=True`` forces that).
"""
from __future__ import annotations

import collections
import gzip
import os
import pickle

import numpy as np
import torch

SEED = 66478
Datasets = collections.namedtuple("Datasets", ["train", "validation", "test"])


class DataSet:
    def __init__(self, images: np.ndarray, labels: np.ndarray, seed: int = SEED):
        assert images.shape[0] == labels.shape[0], (images.shape, labels.shape)
        self._images, self._labels = images, labels
        self._num_examples = images.shape[0]
        self._epochs_completed = 0
        self._index_in_epoch = 0
        self._rng = np.random.RandomState(seed)
        self._shuffle()

    def _shuffle(self):
        perm = self._rng.permutation(self._num_examples)
        self._images = self._images[perm]
        self._labels = self._labels[perm]

    images = property(lambda self: self._images)
    labels = property(lambda self: self._labels)
    num_examples = property(lambda self: self._num_examples)
    epochs_completed = property(lambda self: self._epochs_completed)

    def next_batch(self, batch_size: int):
        """Next ``batch_size`` examples; at the end of an epoch the data is reshuffled and the batch restarts."""
        start = self._index_in_epoch
        self._index_in_epoch += batch_size
        if self._index_in_epoch > self._num_examples:
            self._epochs_completed += 1
            self._shuffle()
            start = 0
            self._index_in_epoch = batch_size
            assert batch_size <= self._num_examples
        end = self._index_in_epoch
        return self._images[start:end], self._labels[start:end]


# ---- synthetic ------------------------------------------------------------------------------------------
def synthetic_images(n, chw, kind, seed):
    rng = np.random.RandomState(seed)
    if kind == "mnist":
        return (rng.rand(n, *chw).astype(np.float32) - 0.5)                 # [-0.5, 0.5] like the reader
    return (rng.rand(n, *chw).astype(np.float32) * 255.0)                    # raw 0..255 like the reader


def synthetic_datasets(kind: str, n_train=4096, n_test=1024, seed=SEED, worker_id=0) -> Datasets:
    chw = (1, 28, 28) if kind == "mnist" else (3, 32, 32)
    rng = np.random.RandomState(seed + 7)
    # a learnable synthetic task: the label is a function of a fixed random projection of the image
    proj = rng.randn(int(np.prod(chw)), 10).astype(np.float32)

    def make(n, s):
        x = synthetic_images(n, chw, kind, s)
        centred = x.reshape(n, -1) - x.reshape(n, -1).mean(1, keepdims=True)
        y = (centred @ proj).argmax(1).astype(np.int64)
        return x, y

    xtr, ytr = make(n_train, seed + 1)
    xte, yte = make(n_test, seed + 2)
    return Datasets(train=DataSet(xtr, ytr, seed + worker_id), validation=DataSet(xte, yte, seed), test=None)


# ---- MNIST (IDX) ----------------------------------------------------------------------------------------
def _open(path):
    return gzip.open(path, "rb") if path.endswith(".gz") else open(path, "rb")


def extract_data(filename, num_images):
    """IDX3 images -> float32 (N, 28, 28, 1) scaled to [-0.5, 0.5] (reference mnist.py:133-146)."""
    with _open(filename) as f:
        f.read(16)
        buf = f.read(28 * 28 * num_images)
    data = np.frombuffer(buf, dtype=np.uint8).astype(np.float32)
    data = (data - 255.0 / 2.0) / 255.0
    return data.reshape(num_images, 28, 28, 1)


def extract_labels(filename, num_images):
    with _open(filename) as f:
        f.read(8)
        buf = f.read(num_images)
    return np.frombuffer(buf, dtype=np.uint8).astype(np.int64)


def _find(train_dir, names):
    for n in names:
        p = os.path.join(train_dir, n)
        if os.path.exists(p):
            return p
    return None


def read_mnist(train_dir="MNIST_data", reshape=True, synthetic=False, worker_id=0, n_workers=1, seed=SEED) -> Datasets:
    files = {}
    if not synthetic:
        for key, base in (("tri", "train-images-idx3-ubyte"), ("trl", "train-labels-idx1-ubyte"),
                          ("tei", "t10k-images-idx3-ubyte"), ("tel", "t10k-labels-idx1-ubyte")):
            files[key] = _find(train_dir, [base + ".gz", base])
    if synthetic or any(v is None for v in files.values()):
        return synthetic_datasets("mnist", seed=seed, worker_id=worker_id)
    tri, trl = extract_data(files["tri"], 60000), extract_labels(files["trl"], 60000)
    tei, tel = extract_data(files["tei"], 10000), extract_labels(files["tel"], 10000)
    if reshape:
        tri, tei = tri.transpose(0, 3, 1, 2).copy(), tei.transpose(0, 3, 1, 2).copy()
    return Datasets(train=DataSet(tri, trl, seed + worker_id), validation=DataSet(tei, tel, seed), test=None)


# ---- CIFAR-10 (python pickles) --------------------------------------------------------------------------
def read_cifar10(train_dir="cifar10_data", padding_size=0, reshape=True, synthetic=False, worker_id=0, n_workers=1,
                 seed=SEED) -> Datasets:
    root = None
    if not synthetic:
        for cand in (os.path.join(train_dir, "cifar-10-batches-py"), train_dir):
            if os.path.exists(os.path.join(cand, "data_batch_1")):
                root = cand
    if root is None:
        return synthetic_datasets("cifar10", seed=seed, worker_id=worker_id)

    def load(names):
        xs, ys = [], []
        for n in names:
            with open(os.path.join(root, n), "rb") as f:
                d = pickle.load(f, encoding="latin1")
            xs.append(np.asarray(d["data"], dtype=np.float32).reshape(-1, 3, 32, 32))   # pickles are CHW already
            ys.append(np.asarray(d["labels"], dtype=np.int64))
        return np.concatenate(xs), np.concatenate(ys)

    xtr, ytr = load([f"data_batch_{i}" for i in range(1, 6)])
    xte, yte = load(["test_batch"])
    if padding_size:
        pad = ((0, 0), (0, 0), (padding_size, padding_size), (padding_size, padding_size))
        xtr, xte = np.pad(xtr, pad), np.pad(xte, pad)
    if not reshape:
        xtr, xte = xtr.transpose(0, 2, 3, 1).copy(), xte.transpose(0, 2, 3, 1).copy()
    return Datasets(train=DataSet(xtr, ytr, seed + worker_id), validation=DataSet(xte, yte, seed), test=None)


# ---- torch-facing wrappers (reference datasets/__init__.py) ---------------------------------------------
class _NextBatchDataset(torch.utils.data.Dataset):
    def __init__(self, dataset: DataSet, transform=None):
        self.dataset, self.transform = dataset, transform

    def __len__(self):
        return self.dataset.num_examples

    def __getitem__(self, idx):
        x = torch.from_numpy(np.ascontiguousarray(self.dataset.images[idx]))
        if self.transform is not None:
            x = self.transform(x)
        return x, int(self.dataset.labels[idx])

    def next_batch(self, batch_size):
        x, y = self.dataset.next_batch(batch_size)
        return torch.from_numpy(np.ascontiguousarray(x)), torch.from_numpy(np.ascontiguousarray(y))

    def pin_memory(self):
        """Page-lock the image / label arrays once (device backend): ``next_batch`` then hands out views of pinned memory
        and the input staging copies them to the GPU without an intermediate host copy. After an epoch's reshuffle the
        arrays are new (pageable) ones and staging falls back to its own pinned buffers."""
        ds = self.dataset
        if not torch.cuda.is_available() or getattr(ds, "_pinned", None) is not None:
            return self
        imgs = torch.from_numpy(np.ascontiguousarray(ds._images, dtype=np.float32)).pin_memory()
        labs = torch.from_numpy(np.ascontiguousarray(ds._labels, dtype=np.int64)).pin_memory()
        ds._pinned = (imgs, labs)
        ds._images, ds._labels = imgs.numpy(), labs.numpy()
        return self


class MNISTDataset(_NextBatchDataset):
    pass


class Cifar10Dataset(_NextBatchDataset):
    pass


class DataLoader(torch.utils.data.DataLoader):
    """torch DataLoader plus the reference's one addition: a persistent iterator behind ``next_batch()``
    (pytorch_code/data_loader_ops/my_data_loader.py:310,318-319). Worker processes, pinned memory and batch
    re-ordering are torch's own (the reference vendors a copy of torch-0.2's implementation of the same)."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self._persistent_iter = None

    def next_batch(self):
        if self._persistent_iter is None:
            self._persistent_iter = iter(self)
        try:
            return next(self._persistent_iter)
        except StopIteration:
            self._persistent_iter = iter(self)
            return next(self._persistent_iter)


class PinnedBatchStager:
    """Double-buffered pinned-host -> device staging on a side stream (the B200 replacement for the reference's
    pin-memory thread): ``stage(x, y)`` starts the copy of the NEXT batch while the current step computes."""

    def __init__(self, batch_shape, device, n_buffers=2):
        self.device = device
        self.stream = torch.cuda.Stream(device=device)
        self.hx = [torch.empty(batch_shape, dtype=torch.float32).pin_memory() for _ in range(n_buffers)]
        self.hy = [torch.empty(batch_shape[0], dtype=torch.long).pin_memory() for _ in range(n_buffers)]
        self.dx = [torch.empty(batch_shape, dtype=torch.float32, device=device) for _ in range(n_buffers)]
        self.dy = [torch.empty(batch_shape[0], dtype=torch.long, device=device) for _ in range(n_buffers)]
        self.ready = [torch.cuda.Event() for _ in range(n_buffers)]
        self.free = [torch.cuda.Event() for _ in range(n_buffers)]
        self.i = 0
        self.used = [False] * n_buffers
        self._src = [None] * n_buffers

    def stage(self, x, y):
        b = self.i % len(self.hx)
        self.i += 1
        # a batch that already lives in page-locked memory (a pinned data set, see _NextBatchDataset.pin_memory) is copied to
        # the device straight from where it is; anything else goes through this stager's own pinned buffers first
        direct = (x.dtype == torch.float32 and y.dtype == torch.long and x.is_contiguous() and y.is_contiguous()
                  and x.numel() == self.hx[b].numel() and x.is_pinned() and y.is_pinned())
        if direct:
            sx, sy = x.reshape(self.hx[b].shape), y
            self._src[b] = (sx, sy)                          # keep the source alive until the copy has run
        else:
            # the host may be several steps ahead of the device: the DMA that last read this pinned buffer must have run
            # before the buffer is overwritten (an event that was never recorded returns at once)
            self.ready[b].synchronize()
            self.hx[b].view(-1).copy_(x.reshape(-1))         # same element count, any host layout ([B,1,28,28] for a [B,784] net)
            self.hy[b].copy_(y)
            sx, sy = self.hx[b], self.hy[b]
        with torch.cuda.stream(self.stream):
            if self.used[b]:
                self.stream.wait_event(self.free[b])
            self.dx[b].copy_(sx, non_blocking=True)
            self.dy[b].copy_(sy, non_blocking=True)
            self.ready[b].record(self.stream)
        return b

    def acquire(self, b):
        torch.cuda.current_stream(self.device).wait_event(self.ready[b])
        return self.dx[b], self.dy[b]

    def release(self, b):
        self.free[b].record(torch.cuda.current_stream(self.device))
        self.used[b] = True
