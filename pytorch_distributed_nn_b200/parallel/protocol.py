"""Protocol constants shared by every backend (names follow the reference where it has them).

Reference tags (pytorch_code/sync_replicas_master_nn.py:16-19,235-314; distributed_worker.py:202-243):
10 = step, 11+l = weights of tensor l (Async mode), 88+l = gradient of tensor l, 77 = kill signal.
"""
STEP_START_ = 1
MAX_NUM_ITERATIONS = 20000

TAG_STEP = 10
TAG_WEIGHT_BASE = 11
TAG_KILL = 77
TAG_GRAD_BASE = 88

KILL_SIGNAL = -1

# how the aggregated gradient is normalised: north-star default 1/k; the python
# reference divides by n regardless of k (sync_replicas_master_nn.py:212-213); the C++ reference by the
# number actually received (MPI_code/src/distributed/sync_replicas_master_nn.h:124-128).
AVG_MODES = {"k": 0, "n": 1, "received": 2}

STRAGGLER_POLICIES = ("all", "kofn", "interval")
