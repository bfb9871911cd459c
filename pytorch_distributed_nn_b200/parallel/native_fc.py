"""Python face of the native (C++/CUDA) fully-connected parameter server (csrc/fc_ps_native.cu).

Counterpart of running the reference's ``mpirun -n 8 ./distributed_nn`` (MPI_code/Makefile:17-19): master, evaluator
and workers are host threads of one process, each driving a GPU with this library's kernels; Python only fills the
config and reads the result.

    python -m pytorch_distributed_nn_b200.parallel.native_fc --workers 6 --collect 2 --iters 100
"""
from __future__ import annotations

import argparse
import ctypes as C

from ..models.mlp import DIMS_CPP
from ..ops import _lib


def run_native_fc_ps(dims=DIMS_CPP, batch=128, lr=1e-3, n_workers=6, n_to_collect=2, iters=100, shortcircuit=True,
                     timeline=False, evaluator=True, eval_batches=4, delay_workers=(), delay_us_per_layer=0, seed=1,
                     synthetic_examples=8192, max_devices=0, data_dir="", out_dir="outfiles") -> dict:
    cfg = _lib.FcPsConfig()
    cfg.n_dims = len(dims)
    for i, d in enumerate(dims):
        cfg.dims[i] = int(d)
    cfg.batch, cfg.lr, cfg.n_workers, cfg.n_to_collect, cfg.n_train_iters = batch, lr, n_workers, n_to_collect, iters
    cfg.shortcircuit, cfg.generate_timeline, cfg.with_evaluator, cfg.eval_batches = int(shortcircuit), int(timeline), int(evaluator), eval_batches
    mask = 0
    for w in delay_workers:
        mask |= 1 << int(w)
    cfg.delay_worker_mask, cfg.delay_us_per_layer = mask, int(delay_us_per_layer)
    cfg.seed, cfg.synthetic_examples, cfg.max_devices = seed, synthetic_examples, max_devices
    cfg.data_dir, cfg.out_dir = data_dir.encode(), out_dir.encode()
    res = _lib.FcPsResult()
    _lib.call("pdnn_fc_ps_run", C.byref(cfg), C.byref(res))
    return {f: (getattr(res, f).decode() if f == "scheme" else getattr(res, f)) for f, _ in res._fields_}


def selftest(iters=60):
    a, b = C.c_float(), C.c_float()
    _lib.call("pdnn_fc_ps_selftest", iters, C.byref(a), C.byref(b))
    return a.value, b.value


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--workers", type=int, default=6)
    ap.add_argument("--collect", type=int, default=2)
    ap.add_argument("--iters", type=int, default=100)
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--lr", type=float, default=1e-3)
    ap.add_argument("--delay-workers", default="", help="comma separated worker ids that sleep before each backward layer")
    ap.add_argument("--delay-us", type=int, default=0)
    ap.add_argument("--no-shortcircuit", action="store_true")
    ap.add_argument("--timeline", action="store_true")
    ap.add_argument("--data-dir", default="")
    ap.add_argument("--out-dir", default="outfiles")
    a = ap.parse_args(argv)
    delay = [int(x) for x in a.delay_workers.split(",") if x]
    r = run_native_fc_ps(batch=a.batch, lr=a.lr, n_workers=a.workers, n_to_collect=a.collect, iters=a.iters,
                         shortcircuit=not a.no_shortcircuit, timeline=a.timeline, delay_workers=delay,
                         delay_us_per_layer=a.delay_us, data_dir=a.data_dir, out_dir=a.out_dir)
    print(r)
    return r


if __name__ == "__main__":
    main()
