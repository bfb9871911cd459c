"""``mnist.read_data_sets`` (reference pytorch_code/mnist/mnist.py:157-211) without TensorFlow."""
from ..utils.data import DataSet, Datasets, extract_data, extract_labels, read_mnist as read_data_sets  # noqa: F401
from ..utils.data import SEED  # noqa: F401


def load_mnist(train_dir="MNIST-data", worker_id=0, n_workers=1):
    return read_data_sets(train_dir, worker_id=worker_id, n_workers=n_workers)
