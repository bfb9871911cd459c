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


_KINDS = {"none": 0, "off": 0, "const": 1, "uniform": 2, "exp": 3, "lognormal": 4, "pareto": 5}


def philox4x32_10(counter, key):
    """Philox4x32-10 (Salmon et al., SC'11) on Python ints; the device kernel ``ps_delay_sampled_kernel`` runs the same rounds."""
    c0, c1, c2, c3 = [int(v) & 0xFFFFFFFF for v in counter]
    k0, k1 = [int(v) & 0xFFFFFFFF for v in key]
    for _ in range(10):
        p0, p1 = 0xD2511F53 * c0, 0xCD9E8D57 * c2
        c0, c1, c2, c3 = ((p1 >> 32) ^ c1 ^ k0) & 0xFFFFFFFF, p1 & 0xFFFFFFFF, ((p0 >> 32) ^ c3 ^ k1) & 0xFFFFFFFF, p0 & 0xFFFFFFFF
        k0, k1 = (k0 + 0x9E3779B9) & 0xFFFFFFFF, (k1 + 0xBB67AE85) & 0xFFFFFFFF
    return c0, c1, c2, c3


class DeviceDelay:
    """A delay spec whose samples are drawn ON the GPU (``pdnn_delay_sampled``): deterministic in
    (seed, worker, step, layer), identical launch every step, hence part of the step's CUDA graph.

    Same spec language as :func:`parse_delay_spec` plus ``per=layer`` (a fresh draw before every backward layer, what
    ``pure_py_code/distributed_worker.py:130-132`` does) -- default ``per=step`` (one draw per step, taken after the
    forward pass). :meth:`sample_ns` is the host replica of the device draw (exact for const / uniform, to float rounding
    for the transcendental tails) so that tests and studies can predict the schedule."""

    def __init__(self, spec: str, seed: int = 0):
        from ..ops import _lib
        kind, kv = _parse(spec)
        if kind not in _KINDS:
            raise ValueError(f"unknown delay kind {kind!r}")
        self.spec = spec
        self.kind = _KINDS[kind]
        self.per_layer = kv.get("per", "step") == "layer"
        a, b = {1: (float(kv.get("ms", 500.0)), 0.0),
                2: (float(kv.get("lo", 300.0)), float(kv.get("hi", 1200.0))),
                3: (float(kv.get("shift", 0.0)), float(kv.get("mean", 1.0))),
                4: (float(kv.get("mu", 0.0)), float(kv.get("sigma", 1.0))),
                5: (float(kv.get("xm", 1.0)), float(kv.get("alpha", 1.5)))}.get(self.kind, (0.0, 0.0))
        self.a, self.b = a, b
        self.scale = float(kv.get("scale", 1.0))
        self.prob = float(kv.get("p", 1.0))
        self.seed = int(kv.get("seed", seed))
        self.rank_mask = 0
        if "ranks" in kv:
            for r in kv["ranks"].split(","):
                if r != "":
                    self.rank_mask |= 1 << int(r)
        c = _lib.DelaySpec()
        c.kind, c.a, c.b, c.scale, c.prob, c.rank_mask, c.seed = self.kind, a, b, self.scale, self.prob, self.rank_mask, self.seed
        self._c = c
        self.log = None            # optional device tensor [n_layers] of the drawn nanoseconds (set by the caller)

    def launch(self, worker, layer_idx, step_word, closed_word, stream):
        import ctypes as C
        from ..ops import _lib
        if self.kind == 0 or (layer_idx != 0 and not self.per_layer):
            return
        _lib.call("pdnn_delay_sampled", C.byref(self._c), int(worker), int(layer_idx), step_word.data_ptr(),
                  closed_word.data_ptr() if closed_word is not None else None,
                  self.log.data_ptr() if self.log is not None else None, stream)

    def sample_ns(self, worker: int, step: int, layer: int = 0) -> int:
        import numpy as np
        f32 = np.float32
        c = philox4x32_10((step, layer, 0, 0), (self.seed & 0xFFFFFFFF, ((self.seed >> 32) ^ (worker * 0x9E3779B9)) & 0xFFFFFFFF))
        u0, u1, u2 = [f32(v >> 8) * f32(1.0 / 16777216.0) for v in c[:3]]
        hit = (self.rank_mask == 0 or (self.rank_mask >> (worker + 1)) & 1) and (self.prob >= 1.0 or u0 < f32(self.prob))
        ms = f32(0.0)
        if hit:
            a, b = f32(self.a), f32(self.b)
            if self.kind == 1:
                ms = a
            elif self.kind == 2:
                ms = a + (b - a) * u1
            elif self.kind == 3:
                ms = a - b * np.log(f32(1.0) - u1, dtype=f32)
            elif self.kind == 4:
                ms = np.exp(a + b * np.sqrt(f32(-2.0) * np.log(f32(1.0) - u1, dtype=f32), dtype=f32) * np.cos(f32(2.0 * np.pi) * u2, dtype=f32), dtype=f32)
            elif self.kind == 5:
                ms = a * np.power(f32(1.0) - u1, f32(-1.0) / b, dtype=f32)
        return int(max(f32(0.0), ms) * f32(self.scale) * f32(1e6))


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
