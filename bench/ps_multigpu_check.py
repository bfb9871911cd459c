#!/usr/bin/env python3
"""Multi-GPU checks of the device-resident parameter server (run under torchrun, one rank per GPU).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        bench/ps_multigpu_check.py [--network LeNet] [--steps 4]

Checks (asserted on every rank, summary printed by rank 0):
  1. all-n mode: after every step the master's fp32 weights equal the oracle update computed from the gradients
     all ranks actually produced, and every worker's weight buffer equals the cast of the master copy;
  2. k-of-n with an injected device-side delay on the last worker: the master selects the punctual workers,
     names the delayed one in the kill mask every step, and that worker's ``closed`` word carries the step;
  3. the update of a k-of-n step equals the oracle for the observed selection.
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from pytorch_distributed_nn_b200.parallel.b200_ps import B200ParameterServer, W_CLOSED  # noqa: E402
from pytorch_distributed_nn_b200.parallel.delays import parse_delay_spec  # noqa: E402
from pytorch_distributed_nn_b200.parallel.oracle import kofn_update  # noqa: E402


def gather_vec(t, world):
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    return out


def run_case(network, batch, rank, world, steps, k, delay_spec, grad_dtype, colocate, graph=False):
    dev = torch.device("cuda", torch.cuda.current_device())
    delay_fn = parse_delay_spec(delay_spec) if delay_spec else None
    lr, mu = 0.05, 0.5
    ps = B200ParameterServer(network, batch, rank, world, lr=lr, momentum=mu, num_aggregate=k, grad_dtype=grad_dtype,
                             colocate=colocate, delay_fn=delay_fn, device=dev)
    net = ps.net
    chw = net.in_chw
    g = torch.Generator().manual_seed(100 + rank)
    x = torch.randn(batch, *chw, generator=g).to(dev)
    y = torch.randint(0, 10, (batch,), generator=g).to(dev)
    n = net.total
    w_ref = m_ref = None
    if rank == 0:
        w_ref = net.w32.clone().cpu()
        m_ref = torch.zeros(n)
    n_workers = ps.n_workers
    ok = True
    log = []
    for s in range(1, steps + 1):
        if graph and s == 2:
            # from step 2 on the whole step is replayed from a CUDA graph; the capture itself runs step 2 for real
            ps.enable_cuda_graph(x, y)
        else:
            ps.train_step(x, y)
        # no host synchronisation inside the step: the master kernel polls while the workers are still computing
        torch.cuda.synchronize()
        dist.barrier()
        # what each worker produced this step (a killed worker's gradient is partial garbage, never used)
        my_g = net.g[:n].clone() if ps.is_worker else torch.zeros(n, device=dev)
        if grad_dtype == "bf16":
            my_g = my_g.bfloat16().float()
        grads = gather_vec(my_g, world)
        res = torch.zeros(4 + 32, device=dev, dtype=torch.int64)
        if rank == 0:
            r4 = list(ps.last_result())
            res = torch.tensor(r4 + ps.last_bucket_masks + [0] * (32 - len(ps.last_bucket_masks)), device=dev, dtype=torch.int64)
        dist.broadcast(res, 0)
        sel, kill, nrecv, st = [int(v) for v in res.tolist()[:4]]
        bucket_masks = [int(v) for v in res.tolist()[4:]]
        # master copy vs oracle
        master_w = net.w32[:n].clone() if rank == 0 else torch.empty(n, device=dev)
        dist.broadcast(master_w, 0)
        if rank == 0:
            by_worker = {w: grads[ps._worker_rank(w)].cpu() for w in range(n_workers)}
            w_new, m_new = w_ref.clone(), m_ref.clone()
            for b in net.buckets:       # every bucket is averaged over ITS first k arrivals
                who = [w for w in range(n_workers) if bucket_masks[b.index] >> w & 1]
                sl = slice(b.start, b.end)
                w_new[sl], m_new[sl], _, _ = kofn_update(w_ref[sl], m_ref[sl], {w: g[sl] for w, g in by_worker.items()}, who, len(who),
                                                         n_workers, lr=lr, mu=mu)
                if len(who) != (k or n_workers):
                    ok = False
                    log.append(f"step {s}: bucket {b.index} averaged {len(who)} sets, expected {k or n_workers}")
            w_ref, m_ref = w_new, m_new
            err = float((master_w.cpu() - w_ref).abs().max())
            # the in-switch bf16 reduction rounds the SUM to bf16 once more than the oracle does
            tol = 3e-4 if (ps.nvls_gather and grad_dtype == "bf16") else 2e-5
            if err > tol:
                ok = False
                log.append(f"step {s}: master vs oracle max abs err {err:.3e}")
        # every receiver's landing buffer of the next step's parity == cast(master copy)
        if ps.is_worker or ps.is_listener:
            expect = master_w.to(net.w.dtype)
            if not torch.equal(ps.land[ps.step & 1][:n], expect):
                ok = False
                log.append(f"rank {rank} step {s}: weight buffer differs from the master broadcast")
        if st != s:
            ok = False
            log.append(f"step {s}: master reported step {st}")
        if delay_spec and delay_spec.startswith("const"):
            late = n_workers - 1
            if not (kill >> late & 1) or (sel >> late & 1) or nrecv != k:
                ok = False
                log.append(f"step {s}: expected worker {late} killed, got sel={sel:b} kill={kill:b} nrecv={nrecv}")
            if ps.is_worker and ps.worker_id == late and int(ps.flags[W_CLOSED]) != s:
                ok = False
                log.append(f"rank {rank} step {s}: closed word = {int(ps.flags[W_CLOSED])}")
        elif delay_spec:
            # randomised delays: whoever is selected, exactly k sets count, the rest are named in the kill mask, and the
            # update above already matched the oracle for that selection
            if nrecv != k or bin(sel).count("1") != k or (sel | kill) != (1 << n_workers) - 1 or (sel & kill):
                ok = False
                log.append(f"step {s}: inconsistent masks sel={sel:b} kill={kill:b} nrecv={nrecv}")
        else:
            if sel != (1 << n_workers) - 1 or kill != 0:
                ok = False
                log.append(f"step {s}: expected all selected, got sel={sel:b} kill={kill:b}")
    if rank == 0:
        print(f"PSMODE k={ps.k}/{n_workers} grad={grad_dtype} colocate={colocate} nvls_broadcast={ps.nvls} nvls_gather={ps.nvls_gather}", flush=True)
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if not ok:
        print(f"[rank {rank}] FAIL {log}", flush=True)
    ps.close()
    return bool(int(flag))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--network", default="LeNet")
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--steps", type=int, default=4)
    args = ap.parse_args()
    rank = int(os.environ["RANK"])
    world = int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world, device_id=torch.device("cuda", torch.cuda.current_device()))
    results = {}
    results["all_n_colocated_fp32"] = run_case(args.network, args.batch, rank, world, args.steps, None, "", "fp32", True)
    results["all_n_colocated_bf16"] = run_case(args.network, args.batch, rank, world, args.steps, None, "", "bf16", True)
    results["all_n_colocated_cuda_graph"] = run_case(args.network, args.batch, rank, world, args.steps + 1, None, "", "fp32", True, graph=True)
    n_workers = world
    if n_workers >= 2:
        # the last worker is delayed by 30 ms on the device; k = n-1
        results["k_of_n_kill_colocated"] = run_case(args.network, args.batch, rank, world, args.steps, n_workers - 1,
                                                    f"const:ranks={n_workers}:ms=30", "fp32", True)
        # stress: random per-(worker, step) delays, k = n-1: the selection changes from step to step
        results["k_of_n_random_delays"] = run_case(args.network, args.batch, rank, world, 8, n_workers - 1,
                                                   "uniform:lo=0:hi=20:p=0.7:seed=5", "fp32", True)
    if world >= 3:
        results["all_n_master_only_rank0"] = run_case(args.network, args.batch, rank, world, args.steps, None, "", "fp32", False)
    if rank == 0:
        for k, v in results.items():
            print(f"PSCHECK {k} {'ok' if v else 'FAIL'}", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if all(results.values()) else 1)


if __name__ == "__main__":
    main()
