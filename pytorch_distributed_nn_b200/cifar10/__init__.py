"""``cifar10.read_data_sets(padding_size, reshape)`` (reference pytorch_code/cifar10/cifar10.py:394-407) without TF."""
from ..utils.data import DataSet, Datasets  # noqa: F401
from ..utils.data import read_cifar10 as _read


def read_data_sets(padding_size=0, reshape=True, train_dir="cifar10_data", **kw):
    return _read(train_dir, padding_size=padding_size, reshape=reshape, **kw)
