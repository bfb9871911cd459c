"""``DataLoader.next_batch()`` (reference pytorch_code/data_loader_ops/my_data_loader.py:310,318-319)."""
from ..utils.data import DataLoader, PinnedBatchStager  # noqa: F401
