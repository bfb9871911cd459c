"""Straggler (delay) injection: the reference's delay models plus parametric heavy-tail ones.

What the reference does:
  * deterministic subset: ranks 1, 2, 3 ``time.sleep(0.5)`` before every backward layer
    (pure_py_code/distributed_worker.py:130-132)                                   -> ``const``
  * uniform 300..1200 ms per step, commented out (distributed_TF/src/distributed_train.py:288-290) -> ``uniform``
  * real EC2 heterogeneity summarised as stdev/max/p80..p99 (distributed_TF/tools/benchmark.py:83-109); no samples
    are committed, so heavy-tail behaviour is modelled parametrically: ``exp`` (shifted exponential),
    ``lognormal``, ``pareto`` (documented additions, arXiv:1604.00981 is the modelling reference the TF code cites).

A spec is ``kind[:key=value]*``; times are milliseconds and are multiplied by ``scale`` (B200 steps are ~10^3x
shorter than the reference's CPU steps, hence ``--delay-scale``). Examples::

    const:ranks=1,2,3:ms=500:scale=0.01        # the reference's subset delay, scaled to 5 ms
    uniform:lo=300:hi=1200:scale=0.005
    exp:shift=0:mean=2:p=0.2                   # 20% of (worker, step) pairs get an exponential delay
    lognormal:mu=0:sigma=1
    pareto:xm=1:alpha=1.5

The returned function maps ``(worker_id, step) -> nanoseconds`` deterministically (same seed => same schedule on
every rank, so tests can predict which workers the master must drop). On the B200 backend the delay is executed
on the device by ``pdnn_delay`` (a %globaltimer busy-wait that ends early when the step is closed); on CPU backends
it is a ``time.sleep``.
"""
from __future__ import annotations

import math
import random
from typing import Callable, Dict


def _parse(spec: str) -> (str, Dict[str, str]):
    parts = [p for p in spec.split(":") if p]
    kind = parts[0].lower()
    kv = {}
    for p in parts[1:]:
        k, _, v = p.partition("=")
        kv[k.strip()] = v.strip()
    return kind, kv


def parse_delay_spec(spec: str, seed: int = 0) -> Callable[[int, int], int]:
    kind, kv = _parse(spec)
    scale = float(kv.get("scale", 1.0))
    prob = float(kv.get("p", 1.0))
    ranks = None
    if "ranks" in kv:
        ranks = {int(r) for r in kv["ranks"].split(",") if r != ""}
    seed = int(kv.get("seed", seed))

    def rng(worker, step):
        return random.Random((seed * 1000003 + worker) * 1000003 + step)

    def sample_ms(r: random.Random) -> float:
        if kind == "const":
            return float(kv.get("ms", 500.0))
        if kind == "uniform":
            return r.uniform(float(kv.get("lo", 300.0)), float(kv.get("hi", 1200.0)))
        if kind == "exp":
            return float(kv.get("shift", 0.0)) + r.expovariate(1.0 / float(kv.get("mean", 1.0)))
        if kind == "lognormal":
            return r.lognormvariate(float(kv.get("mu", 0.0)), float(kv.get("sigma", 1.0)))
        if kind == "pareto":
            return float(kv.get("xm", 1.0)) * r.paretovariate(float(kv.get("alpha", 1.5)))
        if kind in ("none", "off"):
            return 0.0
        raise ValueError(f"unknown delay kind {kind!r}")

    def fn(worker_id: int, step: int) -> int:
        # the reference counts ranks (rank 0 = master): ``ranks=1,2,3`` are workers 0,1,2 when the master is not
        # colocated. We address *workers by their rank-like id + 1* to keep the reference's spelling.
        if ranks is not None and (worker_id + 1) not in ranks:
            return 0
        r = rng(worker_id, step)
        if prob < 1.0 and r.random() >= prob:
            return 0
        return int(max(0.0, sample_ms(r)) * scale * 1e6)

    fn.spec = spec
    return fn


def summarize_ms(samples_ms):
    """stdev, max and the percentiles the reference's benchmark driver reports (benchmark.py:83-109)."""
    xs = sorted(samples_ms)
    n = len(xs)
    if n == 0:
        return {}
    mean = sum(xs) / n
    var = sum((x - mean) ** 2 for x in xs) / n

    def pct(p):
        return xs[min(n - 1, int(math.ceil(p / 100.0 * n)) - 1)]

    return {"n": n, "mean": mean, "stdev": var ** 0.5, "max": xs[-1], "p50": pct(50), "p80": pct(80), "p90": pct(90),
            "p95": pct(95), "p99": pct(99)}
