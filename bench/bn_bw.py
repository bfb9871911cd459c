#!/usr/bin/env python3
"""Achieved HBM bandwidth of the five BatchNorm passes on the ResNet activation shapes (batch 1024).

    python bench/bn_bw.py            # prints one line per (shape, pass): microseconds and GB/s of mandatory traffic
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from pytorch_distributed_nn_b200.ops import _lib  # noqa: E402


def timed(fn, flush, iters=10):
    ts = []
    for _ in range(iters):
        flush.zero_()          # 512 MB > L2: the next pass starts cold
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    torch.cuda.set_device(0)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    tot = {}
    for (m, c) in [(1024 * 32 * 32, 256), (1024 * 32 * 32, 64), (1024 * 16 * 16, 128), (1024 * 8 * 8, 256), (1024 * 4 * 4, 512)]:
        x = torch.randn(m, c, device="cuda").bfloat16()
        dy = torch.randn(m, c, device="cuda").bfloat16()
        y = torch.empty_like(x); dx = torch.empty_like(x); dres = torch.empty_like(x)
        stats = torch.zeros(2 * c, device="cuda"); mi = torch.zeros(2 * c, device="cuda"); sums = torch.zeros(2 * c, device="cuda")
        g = torch.ones(c, device="cuda"); b = torch.zeros(c, device="cuda"); dg = torch.zeros(c, device="cuda"); db = torch.zeros(c, device="cuda")
        nb = x.numel() * 2
        P = lambda t: t.data_ptr()
        passes = {
            "fwd_stats": (1, lambda: _lib.call("pdnn_bn_fwd_stats", P(x), P(stats), m, c, st)),
            "fwd_apply": (2, lambda: _lib.call("pdnn_bn_fwd_apply", P(x), P(stats), P(g), P(b), 0, P(y), P(mi), 0, 0, m, c, 1e-5, 0.1, 1, st)),
            "fwd_apply_res": (3, lambda: _lib.call("pdnn_bn_fwd_apply", P(x), P(stats), P(g), P(b), P(dy), P(y), P(mi), 0, 0, m, c, 1e-5, 0.1, 1, st)),
            "bwd_reduce_y": (3, lambda: _lib.call("pdnn_bn_bwd_reduce", P(x), P(y), P(dy), P(mi), P(sums), m, c, 1, P(g), P(b), st)),
            "bwd_reduce_x": (2, lambda: _lib.call("pdnn_bn_bwd_reduce", P(x), 0, P(dy), P(mi), P(sums), m, c, 1, P(g), P(b), st)),
            "bwd_apply_x": (3, lambda: _lib.call("pdnn_bn_bwd_apply", P(x), 0, P(dy), P(mi), P(g), P(sums), P(dx), 0, P(dg), P(db), m, c, 1, P(b), st)),
            "bwd_apply_y_res": (5, lambda: _lib.call("pdnn_bn_bwd_apply", P(x), P(y), P(dy), P(mi), P(g), P(sums), P(dx), P(dres), P(dg), P(db), m, c, 1, P(b), st)),
        }
        for name, (k, fn) in passes.items():
            fn(); torch.cuda.synchronize()
            us = timed(fn, flush)
            tot[name] = tot.get(name, 0) + us
            print(f"BN m={m} c={c} {name:16s} {us:8.1f} us  {k * nb / us / 1e3:7.0f} GB/s")
    print("TOTAL", " ".join(f"{k}={v:.0f}us" for k, v in tot.items()), "lib=", os.environ.get("PDNN_B200_LIB", "in-tree"))


if __name__ == "__main__":
    main()
