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
                     synthetic_examples=8192, max_devices=0, data_dir="", out_dir="outfiles", train=None, test=None,
                     init_weights=None, want_weights=False, want_log=False, device_protocol=True) -> dict:
    """``train`` / ``test``: optional ``(images float32 [n, dims[0]], labels int64 [n])`` CPU tensors; ``init_weights``:
    list of ``[(dims[l] + 1), dims[l+1]]`` float32 CPU tensors (bias = last row). With ``want_weights`` / ``want_log`` the
    result gains ``final_weights`` (same list form) and ``accept_log`` (rows ``(step, worker, layer, batch_start)`` in
    the order the master accumulated them), which is what an oracle needs to replay a timing-dependent run."""
    import torch
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
    cfg.device_protocol = int(device_protocol)
    res = _lib.FcPsResult()
    io = _lib.FcPsIO()
    keep = []
    sizes = [(dims[l] + 1) * dims[l + 1] for l in range(len(dims) - 1)]
    if train is not None:
        x, y = train[0].contiguous().float(), train[1].contiguous().long()
        keep += [x, y]
        io.train_images, io.train_labels, io.n_train = x.data_ptr(), y.data_ptr(), x.shape[0]
    if test is not None:
        x, y = test[0].contiguous().float(), test[1].contiguous().long()
        keep += [x, y]
        io.test_images, io.test_labels, io.n_test = x.data_ptr(), y.data_ptr(), x.shape[0]
    if init_weights is not None:
        flat = torch.cat([w.reshape(-1).float() for w in init_weights]).contiguous()
        assert flat.numel() == sum(sizes)
        keep.append(flat)
        io.init_weights = flat.data_ptr()
    final = log = None
    if want_weights:
        final = torch.zeros(sum(sizes), dtype=torch.float32)
        io.final_weights = final.data_ptr()
    if want_log:
        log = torch.zeros((iters * n_workers * (len(dims) - 1), 4), dtype=torch.int32)
        io.accept_log, io.accept_log_rows = log.data_ptr(), log.shape[0]
    _lib.call("pdnn_fc_ps_run_ex", C.byref(cfg), C.byref(res), C.byref(io))
    out = {f: (getattr(res, f).decode() if f == "scheme" else getattr(res, f)) for f, _ in res._fields_}
    if final is not None:
        out["final_weights"] = [t.view(dims[l] + 1, dims[l + 1]) for l, t in enumerate(final.split(sizes))]
    if log is not None:
        out["accept_log"] = log[: io.accept_log_used].tolist()
    return out


def oracle_replay(dims, batch, lr, init_weights, train, accept_log, iters):
    """fp32 PyTorch restatement of the native master's update for exactly the gradients it accumulated:
    per step and layer, W -= lr / count * sum of the accepted workers' gradients, each computed on that worker's
    batch with the weights of that step (reference sync_replicas_master_nn.h:71-128, nn_layer.h:97-178)."""
    import torch
    X, Y = train[0].double(), train[1].long()
    W = [w.clone().double() for w in init_weights]
    L = len(dims) - 1
    by_step = {}
    for step, w, l, start in accept_log:
        by_step.setdefault(step, []).append((w, l, start))

    def grads(start):
        z = X[start: start + batch]
        acts = [z]
        for l in range(L):
            s = torch.cat([acts[-1], torch.ones(batch, 1, dtype=z.dtype)], 1) @ W[l]
            acts.append(torch.sigmoid(s) if l < L - 1 else s)
        d = torch.softmax(acts[-1], 1)
        d[torch.arange(batch), Y[start: start + batch]] -= 1.0
        d /= batch
        out = [None] * L
        for l in range(L - 1, -1, -1):
            out[l] = torch.cat([acts[l], torch.ones(batch, 1, dtype=z.dtype)], 1).t() @ d
            if l > 0:
                d = (d @ W[l][:-1].t()) * acts[l] * (1 - acts[l])
        return out

    for step in range(1, iters + 1):
        cache, acc, cnt = {}, [torch.zeros_like(w) for w in W], [0] * L
        for w, l, start in by_step.get(step, []):
            if (w, start) not in cache:
                cache[(w, start)] = grads(start)
            acc[l] += cache[(w, start)][l]
            cnt[l] += 1
        for l in range(L):
            W[l] = W[l] - lr / max(1, cnt[l]) * acc[l]
    return [w.float() for w in W]


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
