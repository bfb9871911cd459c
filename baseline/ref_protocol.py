#!/usr/bin/env python3
"""Restated reference baseline ("ref-nccl-gpu" / "ref-gloo-cpu" of BASELINE.md).

The reference itself cannot run in this image (Python 2 idioms, torch <= 0.3 ``Variable``, mpi4py + mpirun,
TensorFlow-based readers). This file restates its protocol *structure* on ``torch.distributed`` with
stock PyTorch ops only -- none of this repo's kernels or engine are on this path:

* rank 0 = master, holds the weights, never computes gradients; ranks 1..n = workers
  (pytorch_code/distributed_nn.py:91-102);
* every step the master broadcasts EVERY parameter tensor separately (``--comm-type Bcast``,
  sync_replicas_master_nn.py:259-272);
* workers run forward, then the layer-split backward that sends one message per parameter tensor, deepest first,
  while backward is still running (model_ops/lenet.py:106-152, resnet_split.py:235-326);
* the master receives P x n messages, sums them, divides by n and applies ``param -= lr * grad`` (+ momentum here, so
  that both arms do the same optimiser maths), sync_replicas_master_nn.py:148-228.

Most favourable reading of "the reference's torch.distributed build": GPU tensors, NCCL point-to-point, fp32 wire
(the reference ships fp64 over TCP), bf16 autocast + channels_last compute (the reference computes in fp32 on CPU).
With one rank it degenerates to the reference's ``single_machine.py`` loop (NN_Trainer: forward, backward,
``optimizer.step()``).

    python -m torch.distributed.run --nproc-per-node N baseline/ref_protocol.py --steps K --warmup W [--network ResNet18]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import torch.nn.functional as F  # noqa: E402


class _NcclSender:
    """The 'communicator' handed to SplitNet.backward: one isend per parameter tensor (tags do not exist in NCCL;
    ordering is the protocol)."""

    def __init__(self):
        self.reqs = []

    def send_grad(self, idx, grad):
        r = dist.isend(grad.contiguous(), dst=0)
        self.reqs.append(r)
        return r


def run(args):
    from pytorch_distributed_nn_b200.models import build_model, input_shape
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    use_cuda = torch.cuda.is_available() and not args.cpu
    dev = torch.device("cuda", local) if use_cuda else torch.device("cpu")
    if use_cuda:
        torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl" if use_cuda else "gloo", rank=rank, world_size=world,
                                **({"device_id": dev} if use_cuda else {}))
    B = args.batch_size
    chw = input_shape(args.network)
    amp = use_cuda and args.amp and args.network != "LeNet"
    torch.manual_seed(1)
    net = build_model(args.network, split=world > 1).to(dev)
    if use_cuda and len(chw) == 3 and chw[0] == 3:
        net = net.to(memory_format=torch.channels_last)
    params = list(net.parameters())
    gen = torch.Generator().manual_seed(1234 + rank)
    xs = [torch.empty(B, *chw).uniform_(0, 255, generator=gen).to(dev) for _ in range(4)]
    ys = [torch.randint(0, 10, (B,), generator=gen).to(dev) for _ in range(4)]
    if use_cuda and chw[0] == 3:
        xs = [x.contiguous(memory_format=torch.channels_last) for x in xs]
    n_workers = max(1, world - 1)
    is_master = rank == 0
    mom = [torch.zeros_like(p) for p in params] if (is_master or world == 1) else None
    recv = [[torch.empty_like(p) for p in params] for _ in range(n_workers)] if (is_master and world > 1) else None
    opt = torch.optim.SGD(params, lr=args.lr, momentum=args.momentum) if world == 1 else None

    def single_step(i):
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
            out = net(xs[i % 4])
            loss = F.cross_entropy(out.float(), ys[i % 4])
        loss.backward()
        opt.step()
        return loss

    def master_step(i):
        for p in params:                                            # P broadcasts (Bcast mode)
            dist.broadcast(p.data, src=0)
        reqs = []
        for w in range(n_workers):                                  # P x n receives, reverse parameter order
            for l in range(len(params) - 1, -1, -1):
                reqs.append(dist.irecv(recv[w][l], src=w + 1))
        for r in reqs:
            r.wait()
        with torch.no_grad():
            for l, p in enumerate(params):
                g = recv[0][l].clone()
                for w in range(1, n_workers):
                    g += recv[w][l]
                g /= n_workers
                mom[l].mul_(args.momentum).add_(g)
                p.add_(mom[l], alpha=-args.lr)
        return None

    def worker_step(i):
        for p in params:
            dist.broadcast(p.data, src=0)
        net.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
            logits = net(xs[i % 4])                                 # detached leaf of the layer-split tape
        loss = F.cross_entropy(logits.float(), ys[i % 4])
        loss.backward()
        sender = _NcclSender()
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
            net.backward(logits.grad, sender, [])
        for r in sender.reqs:
            r.wait()
        return loss

    step = single_step if world == 1 else (master_step if is_master else worker_step)

    def sync():
        if use_cuda:
            torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        if use_cuda:
            torch.cuda.synchronize()

    for i in range(args.warmup):
        step(i)
    sync()
    if use_cuda:
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
    import time
    t0 = time.perf_counter()
    for i in range(args.steps):
        step(i)
    if use_cuda:
        e.record()
    sync()
    ms = s.elapsed_time(e) if use_cuda else (time.perf_counter() - t0) * 1e3
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t)
    if rank == 0:
        gb = B * n_workers
        print(json.dumps({
            "metric": "images/sec (sync-SGD parameter server, whole job) + p50/p99 step ms", "impl": "restated-reference",
            "value": round(gb * args.steps / (ms / 1e3), 1), "unit": "images/sec", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 4), "higher_is_better": True, "scaling": "weak",
            "dtype": "bf16-autocast" if amp else "fp32", "data": "synthetic",
            "config": {"model": args.network, "global_batch": gb, "per_worker_batch": B,
                       "parallelism": f"reference PS: 1 master + {n_workers} workers" if world > 1 else "single_machine.py loop",
                       "transport": ("nccl p2p + per-tensor broadcast" if use_cuda else "gloo") if world > 1 else "none"}}))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--network", default="ResNet18")
    ap.add_argument("--batch-size", type=int, default=1024)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--lr", type=float, default=0.01)
    ap.add_argument("--momentum", type=float, default=0.5)
    ap.add_argument("--no-amp", dest="amp", action="store_false", default=True, help="fp32 compute like the reference")
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--gpus", type=int, default=1)
    run(ap.parse_args())


if __name__ == "__main__":
    main()
