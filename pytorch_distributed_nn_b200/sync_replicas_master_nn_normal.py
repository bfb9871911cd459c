"""Master without straggler kill: every step waits for all n workers
(reference pytorch_code/sync_replicas_master_nn_normal.py; the only difference from the kill variant is the policy)."""
from .sync_replicas_master_nn import (MAX_NUM_ITERATIONS, STEP_START_, GradientAccumulator,  # noqa: F401
                                       SyncReplicasMaster_NN as _KillMaster, update_params_dist_version)


class SyncReplicasMaster_NN(_KillMaster):
    _policy_default = "all"

    def __init__(self, comm, **kwargs):
        kwargs = dict(kwargs)
        kwargs["straggler_policy"] = "all"
        kwargs.pop("kill_threshold", None)
        super().__init__(comm, **kwargs)

    def send_kill_signal(self, should_kill_list):
        return []
