"""Distributed entry point: rank 0 -> master, ranks >= 1 -> workers (reference pytorch_code/distributed_nn.py:69-102).

    torchrun --nproc-per-node 8 -m pytorch_distributed_nn_b200.distributed_nn --network=ResNet18 --dataset=Cifar10 \
        --batch-size=1024 --comm-type=Bcast --num-aggregate=7 --synthetic --max-steps 100

The reference selects kill / normal behaviour by editing import lines; here ``--straggler-policy`` picks the module
(``all`` -> the ``*_normal`` classes).
"""
from __future__ import annotations

import argparse
import sys

import torch

from .utils.config import add_fit_args, kwargs_from_args


def load_dataset(args, rank, world):
    from . import cifar10, mnist
    from .datasets import Cifar10Dataset, MNISTDataset
    if args.dataset.upper() == "MNIST":
        ds = mnist.read_data_sets(args.data_dir or "MNIST_data", reshape=True, synthetic=args.synthetic, worker_id=rank, n_workers=world)
        return MNISTDataset(ds.train, None), ds.validation
    if args.dataset.lower() == "cifar10":
        ds = cifar10.read_data_sets(padding_size=0, reshape=True, train_dir=args.data_dir or "cifar10_data", synthetic=args.synthetic,
                                    worker_id=rank, n_workers=world)
        return Cifar10Dataset(ds.train, None), ds.validation
    raise ValueError(f"unknown --dataset {args.dataset!r}")


def resolve_backend(args):
    if args.backend != "auto":
        return args.backend
    return "gloo" if (args.no_cuda or not torch.cuda.is_available()) else "b200"


def main(argv=None):
    parser = add_fit_args(argparse.ArgumentParser(description="B200-native synchronous-SGD parameter server"))
    args = parser.parse_args(argv)
    backend = resolve_backend(args)
    from .parallel.launch import init_distributed
    rank, world, local_rank = init_distributed(backend)
    torch.manual_seed(args.seed)
    if backend == "b200":
        from .parallel.comm import B200Communicator
        comm = B200Communicator(rank, world, None, torch.device("cuda", local_rank))
    else:
        from .parallel.comm import TorchDistCommunicator
        wire = {"fp64": torch.float64, "fp32": torch.float32, "bf16": torch.float32}[args.grad_dtype]
        comm = TorchDistCommunicator("gloo", wire_dtype=wire)
    train_set, val_set = load_dataset(args, rank, world)
    normal = args.straggler_policy == "all"
    if normal:
        from .distributed_worker_normal import DistributedWorker
        from .sync_replicas_master_nn_normal import SyncReplicasMaster_NN
    else:
        from .distributed_worker import DistributedWorker
        from .sync_replicas_master_nn import SyncReplicasMaster_NN
    # the last rank evaluates (C++ / TF reference role); on the device backend it is a passive receiver of the broadcast
    min_world = 2 if (backend == "b200" and args.colocate_worker) else 3
    evaluator_rank = world - 1 if (args.evaluator and world >= min_world) else None
    if evaluator_rank is None:
        args.evaluator = False
    result = None
    if rank == 0:
        master = SyncReplicasMaster_NN(comm=comm, **kwargs_from_args(args, master=True))
        master.build_model()
        print("I am the master: the world size is {}, cur step: {}".format(master.world_size, master.cur_step), flush=True)
        result = master.train(train_set)
        print("Done sending messages to workers! summary:", result, flush=True)
    elif rank == evaluator_rank:
        from .evaluator import B200EvaluatorNN, EvaluatorNN
        ev = (B200EvaluatorNN if backend == "b200" else EvaluatorNN)(comm, val_set, **kwargs_from_args(args, master=False))
        ev.build_model()
        result = ev.train()
    else:
        worker = DistributedWorker(comm=comm, **kwargs_from_args(args, master=False))
        worker.build_model()
        print("I am worker: {} in all {} workers, next step: {}".format(worker.rank, worker.world_size - 1, worker.next_step), flush=True)
        result = worker.train(train_loader=train_set)
        print("Now the next step is: {}".format(worker.next_step), flush=True)
    import torch.distributed as dist
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
    return result


if __name__ == "__main__":
    main()
    sys.exit(0)
