"""Worker without the kill probe: backward always runs to the end
(reference pytorch_code/distributed_worker_normal.py: ``network.backward`` instead of ``backward_signal_kill``)."""
from .distributed_worker import DistributedWorker as _KillWorker, ModelBuffer  # noqa: F401


class DistributedWorker(_KillWorker):
    _kill_aware = False
