#!/usr/bin/env python3
"""Gradient push/aggregate + parameter broadcast sweep, 1 KB - 1 GB (BASELINE.json config 5).

    python -m torch.distributed.run --nproc-per-node N bench/comm_sweep.py [--max-mb 1024] [--iters 10]

For every size S (bytes of bf16 parameters) two arms are timed on the device (CUDA events, max over ranks):

  ours       every rank (n = N workers, master colocated on GPU 0) pushes its fp32 gradient as bf16 into the master's
             slot (`pdnn_ps_push`, NVLink peer stores + release flag); the master kernel (`pdnn_ps_master_step`)
             waits for all n flags, reduces, applies momentum SGD on the fp32 copy and stores the bf16 weights into
             every worker's buffer + ready flag; the rank's stream waits for that flag.
  reference  what the reference's torch.distributed path does for the same payload: `dist.reduce` of the fp32
             gradient to rank 0 (stand-in for n point-to-point receives + host adds), SGD update with torch ops on
             rank 0, `dist.broadcast` of the weights (one tensor; the reference sends one message per layer).

Roofline per step: bytes that must cross NVLink at the master = n-1 slots in + (n-1) weight copies out; reported as
achieved GB/s at the master's ingress+egress against 770 GB/s per direction (measured peer-copy figure of
/opt/skills/guides/B200_PROFILING.md).
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from pytorch_distributed_nn_b200.parallel.b200_ps import B200ParameterServer  # noqa: E402


def timed(fn, iters, world):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / iters], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=int, default=1024)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--out", default="")
    ap.add_argument("--no-cuda-graph", action="store_true")
    ap.add_argument("--n-buckets", type=int, default=1)
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    rows = []
    size = 1024
    while size <= args.max_mb * (1 << 20):
        n = size // 2                                   # bf16 parameters
        ps = B200ParameterServer(f"Flat{n}", 1, rank, world, lr=0.01, momentum=0.5, grad_dtype="bf16", colocate=True, n_buckets=args.n_buckets)
        ps.net.g.normal_()
        if not args.no_cuda_graph:          # the step's kernels as one graph launch: device-side latency, not host launch rate
            ps.train_step(None, None)
            ps.enable_cuda_graph(None, None)

        def ours_step():
            ps.train_step(None, None)
            ps.join()

        ours_ms = timed(ours_step, args.iters, world)
        timeline = None
        if rank == 0:
            ps.last_result()
            timeline = ps.last_timing_us
        nvls = bool(ps.nvls_gather)
        ps.close()
        del ps
        torch.cuda.empty_cache()
        g = torch.randn(n, device="cuda")
        w = torch.randn(n, device="cuda")
        m = torch.zeros(n, device="cuda")

        def ref_step():
            if world > 1:
                dist.reduce(g, dst=0)
            if rank == 0:
                m.mul_(0.5).add_(g, alpha=1.0 / world)
                w.add_(m, alpha=-0.01)
            if world > 1:
                dist.broadcast(w, src=0)

        ref_ms = timed(ref_step, args.iters, world)
        del g, w, m
        link_bytes = (world - 1) * n * 2 * 2             # slots in + weights out at the master (bf16 both ways)
        row = {"bytes_bf16": size, "n_params": n, "n_gpus": world, "ours_us": ours_ms * 1e3, "ref_us": ref_ms * 1e3,
               "speedup": ref_ms / ours_ms, "ours_master_link_GBps": link_bytes / (ours_ms * 1e-3) / 1e9 if world > 1 else None,
               # fused update: fp32 master weights + momentum read and written, bf16 weights written (per parameter 18 B of HBM)
               "ours_master_hbm_GBps": n * 18 / (ours_ms * 1e-3) / 1e9, "nvls_gather": nvls, "device_timeline_us": timeline}
        rows.append(row)
        if rank == 0:
            print("SWEEP " + json.dumps(row), flush=True)
        size *= 4
    if rank == 0 and args.out:
        with open(args.out, "w") as f:
            json.dump(rows, f, indent=1)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
