"""Structured per-rank metrics (JSONL) next to the reference's human-readable print lines.

The reference prints per-step timings per rank (distributed_worker.py:195-197, sync_replicas_master_nn.py:208-209)
and scrapes logs with regexes afterwards (distributed_TF/tools/benchmark.py:24-34). Here every rank appends JSON
objects to ``{log_dir}/metrics_rank{r}.jsonl``; ``StepTimer`` accumulates p50/p99 step time and images/sec;
``Timeline`` records (time_ms, step, event, worker) rows like the C++ master's timeline file
(MPI_code/src/distributed/sync_replicas_master_nn.h:61-63,81-83).
"""
from __future__ import annotations

import json
import os
import time
from typing import List, Optional


class JsonlLogger:
    def __init__(self, log_dir: Optional[str], rank: int, name: str = "metrics"):
        self.f = None
        if log_dir:
            os.makedirs(log_dir, exist_ok=True)
            self.f = open(os.path.join(log_dir, f"{name}_rank{rank}.jsonl"), "a")
        self.rank = rank

    def log(self, **kw):
        if self.f is not None:
            kw.setdefault("ts", time.time())
            kw.setdefault("rank", self.rank)
            self.f.write(json.dumps(kw) + "\n")
            self.f.flush()

    def close(self):
        if self.f is not None:
            self.f.close()
            self.f = None


def percentile(xs: List[float], p: float) -> float:
    if not xs:
        return float("nan")
    s = sorted(xs)
    return s[min(len(s) - 1, max(0, int(round(p / 100.0 * (len(s) - 1)))))]


class StepTimer:
    def __init__(self, images_per_step: int):
        self.images_per_step = images_per_step
        self.ms: List[float] = []

    def add(self, ms: float):
        self.ms.append(ms)

    def summary(self):
        tot = sum(self.ms)
        return {"steps": len(self.ms), "p50_ms": percentile(self.ms, 50), "p99_ms": percentile(self.ms, 99),
                "mean_ms": tot / max(1, len(self.ms)), "images_per_sec": self.images_per_step * len(self.ms) / max(1e-9, tot / 1e3)}


class Timeline:
    """``time_ms step event who`` rows; ``visualize`` tooling reads them (tools/plot_timeline.py)."""

    def __init__(self, path: Optional[str]):
        self.rows = []
        self.path = path
        self.t0 = time.monotonic()

    def mark(self, step: int, event: str, who: int = 0):
        self.rows.append(((time.monotonic() - self.t0) * 1e3, step, event, who))

    def flush(self):
        if self.path:
            os.makedirs(os.path.dirname(self.path) or ".", exist_ok=True)
            with open(self.path, "a") as f:
                for r in self.rows:
                    f.write("%.3f %d %s %d\n" % r)
        self.rows = []
