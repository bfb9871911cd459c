"""Process-group bring-up and a local multi-process launcher.

The reference is launched with ``mpirun -n <1+n_workers> [--hostfile H] python distributed_nn.py ...``
(pytorch_code/README.md:8-11). Here the same script runs under ``torchrun`` (one rank per GPU) or, for CPU tests,
under :func:`spawn_local` which forks N gloo ranks on 127.0.0.1.
"""
from __future__ import annotations

import os
import socket
import subprocess
import sys
from typing import List, Optional

import torch
import torch.distributed as dist


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def init_distributed(backend: str = "gloo"):
    """Initialise torch.distributed from the torchrun environment. Returns (rank, world, local_rank)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    if not dist.is_initialized():
        if backend == "b200":
            torch.cuda.set_device(local_rank)
            if world > 1:
                dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world,
                                        device_id=torch.device("cuda", local_rank))
        elif backend in ("nccl", "nccl-ref"):
            torch.cuda.set_device(local_rank)
            dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local_rank))
        else:
            dist.init_process_group("gloo", rank=rank, world_size=world)
    return rank, world, local_rank


def spawn_local(script_args: List[str], nprocs: int, env: Optional[dict] = None, timeout: float = 600.0, python: str = sys.executable):
    """Run ``python <script_args>`` as ``nprocs`` ranks on this host (gloo rendezvous on 127.0.0.1).
    Returns the list of (returncode, stdout+stderr) per rank."""
    port = free_port()
    procs = []
    for r in range(nprocs):
        e = dict(os.environ)
        e.update({"RANK": str(r), "WORLD_SIZE": str(nprocs), "LOCAL_RANK": str(r), "MASTER_ADDR": "127.0.0.1",
                  "MASTER_PORT": str(port), "OMP_NUM_THREADS": e.get("OMP_NUM_THREADS", "2")})
        if env:
            e.update(env)
        procs.append(subprocess.Popen([python] + script_args, env=e, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    out = []
    for p in procs:
        try:
            o, _ = p.communicate(timeout=timeout)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            o, _ = p.communicate()
            o = (o or "") + "\n[spawn_local] TIMEOUT"
        out.append((p.returncode, o))
    return out
