#!/usr/bin/env python3
"""Analysis half of the straggler study (reference distributed_TF/tools/benchmark.py:60-263): from the files a run leaves
in its ``--train-dir`` -- ``master_rank0.jsonl`` (step, step_ms, ts), ``worker_rank<r>.jsonl`` (step, loss, total_s, ts) and the
evaluator's ``time_loss_out_*`` rows (step time_ms loss err) -- produce the reference's five figures

    time_loss.png  time_step.png  time_precision.png  step_losses.png  time_cdfs.png

and print, per run, the compute-time statistics of its ``print_worker_sorted_times``: stdev, max, p80 / p90 / p95 / p99 over
all (worker, iteration) samples and the percentiles of the slowest worker per iteration.

    python tools/plot_study.py k6=/tmp/run_k6 k8=/tmp/run_k8 [--out figs] [--study results/study.json]

Without matplotlib the same series are written as CSV files (one per figure) and summarised as text. ``--study`` adds the
table of a ``bench/straggler_sweep.py`` result file (one row per config).
"""
import argparse
import glob
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pytorch_distributed_nn_b200.parallel.delays import summarize_ms  # noqa: E402


def _jsonl(path):
    rows = []
    if os.path.exists(path):
        with open(path) as f:
            for line in f:
                line = line.strip()
                if line:
                    rows.append(json.loads(line))
    return rows


def load_run(train_dir):
    """-> dict of series; times are seconds since the run's first record."""
    master = [r for r in _jsonl(os.path.join(train_dir, "master_rank0.jsonl")) if "step" in r and "event" not in r]
    workers = {}
    for p in sorted(glob.glob(os.path.join(train_dir, "worker_rank*.jsonl"))):
        rank = int(os.path.basename(p)[len("worker_rank"):-len(".jsonl")])
        workers[rank] = [r for r in _jsonl(p) if "step" in r and "event" not in r]
    t0 = min([r["ts"] for r in master] + [r["ts"] for rs in workers.values() for r in rs] or [0.0])
    run = {"time_step": [(r["ts"] - t0, r["step"]) for r in master],
           "step_ms": [r["step_ms"] for r in master if "step_ms" in r],
           "time_loss": [], "step_loss": [], "time_precision": [], "compute_s": {}}
    # the loss curve is the first worker's (every worker sees the same weights at the same step)
    for rank in sorted(workers):
        rs = [r for r in workers[rank] if "loss" in r]
        if rs and not run["time_loss"]:
            run["time_loss"] = [(r["ts"] - t0, r["loss"]) for r in rs]
            run["step_loss"] = [(r["step"], r["loss"]) for r in rs]
        run["compute_s"][rank] = {r["step"]: r["total_s"] for r in workers[rank] if "total_s" in r}
    if not run["time_loss"]:                                  # device backend: the master logs the loss it reads back
        run["time_loss"] = [(r["ts"] - t0, r["loss"]) for r in master if r.get("loss") is not None]
        run["step_loss"] = [(r["step"], r["loss"]) for r in master if r.get("loss") is not None]
    for p in sorted(glob.glob(os.path.join(train_dir, "time_loss_out_*")) + glob.glob(os.path.join(train_dir, "outfiles", "time_loss_out_*"))):
        for line in open(p):
            v = line.split()
            if len(v) == 4:
                run["time_precision"].append((float(v[1]) / 1e3, 100.0 * (1.0 - float(v[3]))))
    return run


def compute_time_stats(compute_s):
    """benchmark.py:60-111: all samples, and per iteration the slowest worker."""
    all_ms = [1e3 * t for per in compute_s.values() for t in per.values()]
    steps = sorted(set(s for per in compute_s.values() for s in per))
    slowest = [1e3 * max(per[s] for per in compute_s.values() if s in per) for s in steps]
    return {"all": summarize_ms(all_ms) if all_ms else None, "slowest_worker_per_iteration": summarize_ms(slowest) if slowest else None}


def cdf(samples):
    s = sorted(samples)
    n = len(s)
    return [(v, (i + 1) / n) for i, v in enumerate(s)]


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("runs", nargs="*", help="label=train_dir")
    ap.add_argument("--out", default=".")
    ap.add_argument("--study", default="", help="result file of bench/straggler_sweep.py")
    a = ap.parse_args(argv)
    os.makedirs(a.out, exist_ok=True)
    runs = {}
    for spec in a.runs:
        label, _, d = spec.partition("=")
        runs[label] = load_run(d or label)
    figs = {"time_loss": ("time (s)", "loss", True), "time_step": ("time (s)", "step", False),
            "time_precision": ("time (s)", "precision (%)", False), "step_losses": ("step", "losses", True),
            "time_cdfs": ("Time (s)", "P(X <= x)", False)}
    series = {k: {} for k in figs}
    report = {}
    for label, run in runs.items():
        series["time_loss"][label] = run["time_loss"]
        series["time_step"][label] = run["time_step"]
        series["time_precision"][label] = run["time_precision"]
        series["step_losses"][label] = run["step_loss"]
        samples = [t for per in run["compute_s"].values() for t in per.values()]
        series["time_cdfs"][label] = cdf(samples)
        report[label] = {"steps": len(run["time_step"]), "step_ms": summarize_ms(run["step_ms"]) if run["step_ms"] else None,
                         "compute_times_ms": compute_time_stats(run["compute_s"])}
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        for name, (xl, yl, logy) in figs.items():
            plt.figure()
            plt.xlabel(xl)
            plt.ylabel(yl)
            if logy:
                plt.yscale("log")
            for label, pts in series[name].items():
                if pts:
                    plt.plot([p[0] for p in pts], [p[1] for p in pts], linestyle="solid", label=label)
            plt.legend(fontsize=8)
            plt.savefig(os.path.join(a.out, name + ".png"))
            plt.close()
        wrote = "png"
    except ImportError:
        for name in figs:
            with open(os.path.join(a.out, name + ".csv"), "w") as f:
                f.write("run,x,y\n")
                for label, pts in series[name].items():
                    for x, y in pts:
                        f.write(f"{label},{x:.6g},{y:.6g}\n")
        wrote = "csv"
    if a.study:
        with open(a.study) as f:
            rows = [r for r in json.load(f) if r and "error" not in r]
        report["study"] = [{"name": r["name"], "policy": r["policy"], "k": r["k"], "interval_ms": r["interval_ms"], "delay": r["inject_delay"],
                            "images_per_sec": round(r["images_per_sec"], 1), "p50_ms": r["step_ms"]["p50"], "p99_ms": r["step_ms"]["p99"],
                            "sets_per_step": r["sets_per_step"]} for r in rows]
    report["figures"] = [os.path.join(a.out, n + "." + wrote) for n in figs]
    print(json.dumps(report, indent=1))
    return report


if __name__ == "__main__":
    main()
