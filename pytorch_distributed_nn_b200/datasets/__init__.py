"""Dataset façades with ``next_batch`` (reference pytorch_code/datasets/__init__.py:1-51)."""
from ..utils.data import Cifar10Dataset, MNISTDataset  # noqa: F401
