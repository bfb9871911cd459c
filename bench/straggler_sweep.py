#!/usr/bin/env python3
"""The straggler study in ONE job (counterpart of distributed_TF/tools/benchmark.py:60-111,226-265): every config of
configs/k_sweep, configs/interval_sweep and configs/time_cdf is run back to back inside one torchrun launch, each as a
fresh parameter server, for ``n_iters`` steps (reference: 300) replayed from CUDA graphs with the delays sampled on
the device.

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 bench/straggler_sweep.py configs/k_sweep/* \
        configs/interval_sweep/* configs/time_cdf/* [--iters 300] [--master-only] [--out results/straggler_study.json]

Per config (rank 0 prints ``STUDY {json}``): images/sec from the gradient sets the master really applied, p50 / p99 /
max of the device-timed step (release to release on the master's stream), how often each worker was among the first k, and the
per-(worker, step) compute-time statistics the reference's driver plots (stdev, max, p80..p99, slowest-worker-per-
iteration percentiles). ``--master-only`` makes GPU 0 a pure parameter server as in the reference (n = N-1 workers).
"""
import argparse
import ast
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from pytorch_distributed_nn_b200.parallel.b200_ps import B200ParameterServer  # noqa: E402
from pytorch_distributed_nn_b200.parallel.delays import DeviceDelay, summarize_ms  # noqa: E402
from tools.cluster import Cfg  # noqa: E402


def run_config(cfg, rank, world, dev, iters, master_only, warmup=5):
    B = int(cfg["batch_size"])
    spec = cfg["inject_delay"] if cfg["inject_delay"] not in ("", "none", None) else ""
    colocate = not master_only
    n_workers = world if colocate else world - 1
    k = min(int(cfg["num_aggregate"]), n_workers) if cfg["straggler_policy"] == "kofn" else None
    deadline = float(cfg["interval_ms"]) if cfg["straggler_policy"] == "interval" else 0.0
    ps = B200ParameterServer(cfg["network"], B, rank, world, lr=float(cfg["lr"]), momentum=0.5, num_aggregate=k, colocate=colocate,
                             device_delay=DeviceDelay(spec, seed=1) if spec else None, device=dev, deadline_ms=deadline)
    x = y = None
    if ps.is_worker:
        g = torch.Generator().manual_seed(7 + rank)
        x = torch.empty(B, *ps.net.in_chw).uniform_(0, 255, generator=g).to(dev)
        y = torch.randint(0, 10, (B,), generator=g).to(dev)
    for _ in range(2):
        ps.train_step(x, y)
    ps.enable_cuda_graph(x, y)
    for _ in range(warmup):
        ps.train_step(x, y)
    torch.cuda.synchronize()
    dist.barrier()
    sets0 = 0
    sel0 = [0] * n_workers
    if rank == 0:
        ps.last_result()
        sets0, sel0 = ps.total_sets_applied, list(ps.selected_counts)
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
    evs[0].record()
    for i in range(iters):
        ps.train_step(x, y)
        ps.join()
        evs[i + 1].record()
    torch.cuda.synchronize()
    dist.barrier()
    mine = torch.tensor([evs[i].elapsed_time(evs[i + 1]) for i in range(iters)], device=dev, dtype=torch.float64)
    allt = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(allt, mine)
    out = None
    if rank == 0:
        ps.last_result()
        mat = torch.stack(allt).cpu()                                   # [rank][iteration]
        # the step clock is the MASTER's: rank 0's interval ends when its last bucket has released the workers. (A max over
        # ranks per step index would count one delay twice: ranks are up to a step apart, so the same wait shows up in
        # neighbouring intervals of different ranks.)
        step_ms = mat[0]
        total_s = float(mat.sum(dim=1).max()) / 1e3
        sets = ps.total_sets_applied - sets0
        workers = mat[(0 if colocate else 1):]                          # ranks that train
        st = summarize_ms(step_ms.tolist())
        out = {"name": cfg["name"], "network": cfg["network"], "per_worker_batch": B, "n_gpus": world, "n_workers": n_workers,
               "policy": cfg["straggler_policy"], "k": ps.k, "interval_ms": deadline, "inject_delay": spec or "none", "iters": iters,
               "cuda_graph": True, "master_mode": ps.master_mode, "images_per_sec": B * sets / total_s, "sets_per_step": sets / iters,
               "step_ms": {"p50": st["p50"], "p99": st["p99"], "max": st["max"], "mean": st["mean"]},
               "selected_per_worker": [a - b for a, b in zip(ps.selected_counts, sel0)],
               "worker_times": {"all": summarize_ms(workers.flatten().tolist()),
                                "slowest_worker_per_iteration": summarize_ms(workers.max(dim=0).values.tolist())}}
    ps.close()
    del ps
    torch.cuda.empty_cache()
    dist.barrier()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("configs", nargs="+")
    ap.add_argument("--iters", type=int, default=0)
    ap.add_argument("--master-only", action="store_true")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world, device_id=dev)
    rows = []
    for path in a.configs:
        with open(path) as f:
            cfg = Cfg(ast.literal_eval(f.read()))
        try:
            r = run_config(cfg, rank, world, dev, a.iters or int(cfg["n_iters"]), a.master_only)
        except Exception as ex:  # noqa: BLE001 -- one broken config must not lose the rest of the study
            r = {"name": cfg["name"], "error": repr(ex)} if rank == 0 else None
        if rank == 0:
            rows.append(r)
            print("STUDY " + json.dumps(r), flush=True)
            if a.out:                                                   # rewritten after every config: a cut-off job keeps its rows
                os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
                with open(a.out, "w") as f:
                    json.dump(rows, f, indent=1)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
