"""Prepare the datasets ONCE, before N ranks start, so that they neither race on the files nor each build their own
copy (reference pytorch_code/data_prepare.py:1-17: it calls both readers once, whose side effect is the download).

There is no network here, so "prepare" means, per dataset directory:

1. files already there      -> parse them once with this package's readers and report the sizes (a corrupt file fails
                               here, not inside rank 5 of a job);
2. a local archive is there -> unpack it (``cifar-10-python.tar.gz`` as the reference's reader would after its
                               download, cifar10/cifar10.py:126-144; gzipped MNIST idx files are read in place);
3. nothing is there         -> with ``--synthesize`` write random data in the exact on-disk formats (CIFAR-10 python
                               pickles, MNIST idx.gz) so that every rank of every later job reads the SAME files
                               through the real parsing path; without it, report that the in-memory generators of
                               ``utils.data.synthetic_datasets`` will be used.

    python -m pytorch_distributed_nn_b200.utils.data_prepare [--mnist-dir MNIST_data] [--cifar-dir cifar10_data] [--synthesize]
"""
from __future__ import annotations

import argparse
import gzip
import os
import pickle
import struct
import tarfile

import numpy as np

from .data import read_cifar10, read_mnist

MNIST_FILES = (("train-images-idx3-ubyte", 60000, True), ("train-labels-idx1-ubyte", 60000, False),
               ("t10k-images-idx3-ubyte", 10000, True), ("t10k-labels-idx1-ubyte", 10000, False))
CIFAR_FILES = tuple(f"data_batch_{i}" for i in range(1, 6)) + ("test_batch",)


def _has_mnist(d):
    return all(os.path.exists(os.path.join(d, f + ".gz")) for f, _, _ in MNIST_FILES) or \
        all(os.path.exists(os.path.join(d, f)) for f, _, _ in MNIST_FILES)


def _cifar_root(d):
    for cand in (os.path.join(d, "cifar-10-batches-py"), d):
        if all(os.path.exists(os.path.join(cand, f)) for f in CIFAR_FILES):
            return cand
    return None


def synthesize_mnist(d, seed=66478):
    """idx3 / idx1 big-endian files, gzipped (what mnist/mnist.py:133-156 of the reference parses)."""
    os.makedirs(d, exist_ok=True)
    rng = np.random.RandomState(seed)
    for name, n, is_img in MNIST_FILES:
        tmp = os.path.join(d, name + ".gz.tmp")
        with gzip.open(tmp, "wb", compresslevel=1) as f:
            if is_img:
                f.write(struct.pack(">IIII", 2051, n, 28, 28))
                f.write(rng.randint(0, 256, size=n * 784).astype(np.uint8).tobytes())
            else:
                f.write(struct.pack(">II", 2049, n))
                f.write(rng.randint(0, 10, size=n).astype(np.uint8).tobytes())
        os.replace(tmp, os.path.join(d, name + ".gz"))


def synthesize_cifar10(d, seed=66478):
    """``cifar-10-batches-py/{data_batch_1..5,test_batch}``: dict(data uint8 [10000, 3072] CHW, labels list)."""
    root = os.path.join(d, "cifar-10-batches-py")
    os.makedirs(root, exist_ok=True)
    rng = np.random.RandomState(seed)
    for name in CIFAR_FILES:
        blob = {"data": rng.randint(0, 256, size=(10000, 3072)).astype(np.uint8), "labels": [int(v) for v in rng.randint(0, 10, size=10000)]}
        tmp = os.path.join(root, name + ".tmp")
        with open(tmp, "wb") as f:
            pickle.dump(blob, f, protocol=2)
        os.replace(tmp, os.path.join(root, name))


def prepare_mnist(d, synthesize=False):
    if _has_mnist(d):
        how = "found"
    elif synthesize:
        synthesize_mnist(d)
        how = "synthesized"
    else:
        return {"dataset": "MNIST", "dir": d, "status": "absent: the in-memory synthetic generator will be used"}
    ds = read_mnist(d)
    return {"dataset": "MNIST", "dir": d, "status": how, "train": ds.train.num_examples, "test": ds.validation.num_examples,
            "image_shape": tuple(ds.train.images.shape[1:])}


def prepare_cifar10(d, synthesize=False):
    how = "found"
    if _cifar_root(d) is None:
        archive = os.path.join(d, "cifar-10-python.tar.gz")
        if os.path.exists(archive) and os.path.getsize(archive) > 0:
            with tarfile.open(archive, "r:gz") as t:
                t.extractall(d)
            how = "unpacked " + archive
        elif synthesize:
            synthesize_cifar10(d)
            how = "synthesized"
        else:
            return {"dataset": "CIFAR-10", "dir": d, "status": "absent: the in-memory synthetic generator will be used"}
    ds = read_cifar10(d)
    return {"dataset": "CIFAR-10", "dir": d, "status": how, "train": ds.train.num_examples, "test": ds.validation.num_examples,
            "image_shape": tuple(ds.train.images.shape[1:])}


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("--mnist-dir", default="MNIST_data")
    ap.add_argument("--cifar-dir", default="cifar10_data")
    ap.add_argument("--synthesize", action="store_true", help="write random data in the real file formats when a dataset is absent")
    a = ap.parse_args(argv)
    out = [prepare_mnist(a.mnist_dir, a.synthesize), prepare_cifar10(a.cifar_dir, a.synthesize)]
    for r in out:
        print(r)
    return out


if __name__ == "__main__":
    main()
