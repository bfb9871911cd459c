#!/usr/bin/env python3
"""Benchmark driver for the straggler experiments (counterpart of distributed_TF/tools/benchmark.py).

The reference driver launches an EC2 cluster per config, runs 300 iterations, downloads the logs and plots
time-loss / time-step curves and compute-time CDFs with stdev, max, p80/p90/p95/p99 and per-iteration p95/p99/p100.
Here a config (configs/k_sweep/*, configs/interval_sweep/*, configs/time_cdf/*: python-literal dicts, %(key)s
interpolation) maps onto one `bench.py` run on the local box; results land in ``results/<name>.json`` plus a summary
table (images/sec, p50/p99 step ms, who got killed, compute-time percentiles).

    python bench/straggler_study.py configs/k_sweep/* [--gpus 8] [--iters 100] [--dry-run]
"""
import argparse
import ast
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tools.cluster import Cfg  # noqa: E402


def load(path):
    with open(path) as f:
        return Cfg(ast.literal_eval(f.read()))


def command(cfg, gpus, iters, port):
    args = ["bench.py", "--gpus", str(gpus), "--steps", str(iters), "--warmup", "5", "--network", cfg["network"],
            "--batch-size", str(cfg["batch_size"]), "--lr", str(cfg["lr"])]
    if cfg["straggler_policy"] == "kofn":
        args += ["--num-aggregate", str(min(cfg["num_aggregate"], gpus))]
    if cfg["straggler_policy"] == "interval":
        args += ["--interval-ms", str(cfg["interval_ms"])]
    if cfg["inject_delay"] and cfg["inject_delay"] != "none":
        args += ["--inject-delay", cfg["inject_delay"]]
    if cfg["worker_times_cdf"]:
        args += ["--worker-times-cdf"]
    if gpus == 1:
        return [sys.executable] + args
    return [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={gpus}", "--master-addr", "127.0.0.1",
            "--master-port", str(port)] + args


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("configs", nargs="+")
    ap.add_argument("--gpus", type=int, default=0, help="override n_gpus of the configs")
    ap.add_argument("--iters", type=int, default=0, help="override n_iters of the configs (reference: 300)")
    ap.add_argument("--out", default="results")
    ap.add_argument("--dry-run", action="store_true")
    a = ap.parse_args()
    os.makedirs(a.out, exist_ok=True)
    table = []
    for i, path in enumerate(a.configs):
        cfg = load(path)
        gpus = a.gpus or cfg["n_gpus"]
        iters = a.iters or cfg["n_iters"]
        cmd = command(cfg, gpus, iters, 29600 + i)
        print("#", cfg["name"], " ".join(cmd), flush=True)
        if a.dry_run:
            continue
        r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=1800)
        line = next((l for l in reversed(r.stdout.splitlines()) if l.startswith("{")), None)
        if r.returncode != 0 or line is None:
            print("FAILED", cfg["name"], r.stderr[-2000:])
            continue
        res = json.loads(line)
        with open(os.path.join(a.out, cfg["name"] + ".json"), "w") as f:
            json.dump(res, f, indent=1)
        cdf = res.get("worker_times_cdf", {}).get("all", {})
        table.append((cfg["name"], res["value"], res["p50_step_ms"], res["p99_step_ms"], res["last_master_result"]["kill_mask"],
                      cdf.get("stdev"), cdf.get("p99"), cdf.get("max")))
    if table:
        print("\n%-36s %12s %9s %9s %6s %9s %9s %9s" % ("config", "images/sec", "p50 ms", "p99 ms", "kill", "cdf stdev", "cdf p99", "cdf max"))
        for t in table:
            print("%-36s %12.1f %9.3f %9.3f %6s %9s %9s %9s" % tuple(x if x is not None else "-" for x in t))


if __name__ == "__main__":
    main()
