"""`torch.distributed` data-parallel trainer on MNIST-shaped data.

Reference: pytorch_code/backup/dist_nn_test.py:1-142 — the torch MNIST example half-way converted to
``torch.distributed`` (``--dist-backend`` default ``gloo``, line 88; ``init_process_group``, line 95) with a pasted tail
that references undefined names, so it never ran. This is the experiment it was heading for: every rank trains the same
LeNet on its own shard, gradients are averaged with one all-reduce per parameter, rank 0 reports the loss.

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 2 -m pytorch_distributed_nn_b200.backup.dist_nn_test --steps 20
"""
from __future__ import annotations

import argparse
import os

import torch
import torch.distributed as dist
import torch.nn.functional as F

from ..models.lenet import LeNet


def average_gradients(model, world):
    for p in model.parameters():
        if p.grad is not None:
            dist.all_reduce(p.grad, op=dist.ReduceOp.SUM)
            p.grad.div_(world)


def train(steps=20, batch_size=64, lr=0.01, momentum=0.5, seed=1, dist_backend="gloo", log=print):
    """Returns the list of (rank-0) losses. Works without a process group too (world size 1)."""
    distributed = "RANK" in os.environ and "WORLD_SIZE" in os.environ
    rank, world = 0, 1
    if distributed:
        if not dist.is_initialized():
            dist.init_process_group(backend=dist_backend)
        rank, world = dist.get_rank(), dist.get_world_size()
    torch.manual_seed(seed)                                   # identical initial weights on every rank
    model = LeNet()
    opt = torch.optim.SGD(model.parameters(), lr=lr, momentum=momentum)
    g = torch.Generator().manual_seed(seed + 1000 * (rank + 1))   # a different shard per rank
    # a fixed random teacher gives the labels so that the loss can actually fall
    teacher = torch.randn(28 * 28, 10, generator=torch.Generator().manual_seed(7))
    losses = []
    for step in range(steps):
        x = torch.randn(batch_size, 1, 28, 28, generator=g)
        y = (x.flatten(1) @ teacher).argmax(1)
        opt.zero_grad()
        loss = F.cross_entropy(model(x), y)
        loss.backward()
        if world > 1:
            average_gradients(model, world)
        opt.step()
        losses.append(float(loss))
        if rank == 0 and (step % 5 == 0 or step == steps - 1):
            log(f"Train step {step} world {world} loss {float(loss):.4f}")
    if world > 1:
        # replicas that started equal and applied the same averaged gradients must still be equal
        flat = torch.cat([p.detach().flatten() for p in model.parameters()])
        ref = flat.clone()
        dist.broadcast(ref, src=0)
        assert torch.allclose(flat, ref, atol=1e-6), "replicas diverged"
    return losses


def main(argv=None):
    ap = argparse.ArgumentParser(description="torch.distributed MNIST-shape example")
    ap.add_argument("--batch-size", type=int, default=64)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--lr", type=float, default=0.01)
    ap.add_argument("--momentum", type=float, default=0.5)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--dist-backend", default="gloo")
    a = ap.parse_args(argv)
    train(a.steps, a.batch_size, a.lr, a.momentum, a.seed, a.dist_backend)
    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
