#!/usr/bin/env python3
"""Epilogue-bound GEMM shapes (tiny K, N = 64/128): time per call and output bandwidth. Used to A/B epilogue variants
(PDNN_B200_LIB=<other build> python bench/epi_perf.py)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from pytorch_distributed_nn_b200.ops import functional as F  # noqa: E402


def time_us(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3


def main():
    torch.cuda.set_device(0)
    tag = os.path.basename(os.environ.get("PDNN_B200_LIB", "in-tree"))
    for (M, N, K) in [(1 << 20, 64, 32), (1 << 20, 64, 64), (1 << 18, 128, 64), (1 << 20, 64, 576), (1 << 18, 128, 1152)]:
        a = torch.randn(M, K, device="cuda").bfloat16()
        b = torch.randn(N, K, device="cuda").bfloat16()
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        us = time_us(lambda: F.gemm(a, b, out=out))
        print(f"EPI[{tag}] {M}x{N}x{K}: {us:.1f} us  out {M * N * 2 / us / 1e3:.0f} GB/s  {2 * M * N * K / us / 1e6:.0f} TFLOP/s", flush=True)


if __name__ == "__main__":
    main()
