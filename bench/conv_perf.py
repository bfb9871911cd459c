#!/usr/bin/env python3
"""Isolated timings of the implicit-GEMM conv kernels on the ResNet / VGG stage shapes (batch 1024) and of one big GEMM.

    python bench/conv_perf.py            # PDNN_CLUSTER=2 CTA pair (cta_group::2)
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from pytorch_distributed_nn_b200.ops import functional as F  # noqa: E402


def time_ms(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main():
    torch.cuda.set_device(0)
    tag = {"2": "pair"}.get(os.environ.get("PDNN_CLUSTER", ""), "single")
    for (n, h, c, k) in [(1024, 16, 128, 128), (1024, 8, 256, 256), (1024, 4, 512, 512), (1024, 16, 256, 256), (1024, 8, 512, 512)]:
        x = torch.randn(n, h, h, c, device="cuda").bfloat16()
        w = (torch.randn(k, 3, 3, c, device="cuda") / (3 * c ** 0.5)).bfloat16()
        out = torch.empty(n, h, h, k, device="cuda", dtype=torch.bfloat16)
        ms = time_ms(lambda: F.conv_fprop(x, w, pad=1, out=out))
        fl = 2 * n * h * h * k * 9 * c
        print(f"PERF[{tag}] conv3x3 n{n} {h}x{h} c{c} k{k}: {ms * 1e3:.1f} us  {fl / ms / 1e9:.0f} TFLOP/s", flush=True)
    for (M, N, K) in [(8192, 8192, 8192), (65536, 256, 2304), (16384, 512, 4608)]:
        a = torch.randn(M, K, device="cuda").bfloat16()
        b = torch.randn(N, K, device="cuda").bfloat16()
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        ms = time_ms(lambda: F.gemm(a, b, out=out))
        msc = time_ms(lambda: torch.matmul(a, b.t(), out=out))
        print(f"PERF[{tag}] gemm_nt {M}x{N}x{K}: {ms * 1e3:.1f} us  {2 * M * N * K / ms / 1e9:.0f} TFLOP/s  (cuBLAS {msc * 1e3:.1f} us)", flush=True)


if __name__ == "__main__":
    main()
