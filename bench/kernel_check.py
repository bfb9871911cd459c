#!/usr/bin/env python3
"""Numerics + timing checks of every sm_100a kernel against plain PyTorch fp32 references.

    python bench/kernel_check.py all            # each case in its own subprocess (a trap cannot cascade)
    python bench/kernel_check.py gemm_nt        # one case in-process

Prints one line per case: ``CASE <name> ok|FAIL max_rel=<..> [ms=<..> tflops=<..>]``.
The pytest suite (tests/test_gpu_*.py) asserts the same comparisons; this script is the fast diagnostic
used under gpurun.
"""
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

# fp32 references must be true fp32 (no TF32 in cuDNN / cuBLAS)
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False

CASES = {}


def case(fn):
    CASES[fn.__name__] = fn
    return fn


def rel_err(got, ref):
    got = got.float()
    ref = ref.float()
    denom = ref.abs().max().clamp_min(1e-6)
    return float(((got - ref).abs().max() / denom).detach())


def report(name, err, tol, extra=""):
    ok = err == err and err <= tol
    print(f"CASE {name} {'ok' if ok else 'FAIL'} max_rel={err:.3e} tol={tol:.1e} {extra}", flush=True)
    return ok


def time_ms(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    s = torch.cuda.Event(enable_timing=True)
    e = torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


def _F():
    from pytorch_distributed_nn_b200.ops import functional as F
    return F


@case
def gemm_nt():
    F = _F()
    ok = True
    for (M, N, K) in [(128, 64, 64), (256, 128, 128), (384, 256, 512), (1000, 504, 800), (200, 16, 504), (4096, 4096, 4096)]:
        a = torch.randn(M, K, device="cuda").bfloat16()
        b = torch.randn(N, K, device="cuda").bfloat16()
        bias = torch.randn(N, device="cuda")
        out = F.gemm(a, b, bias=bias, relu=True)
        ref = torch.relu(a.float() @ b.float().t() + bias)
        ok &= report(f"gemm_nt[{M}x{N}x{K}]", rel_err(out, ref), 1.5e-2)
    M = N = K = 8192
    a = torch.randn(M, K, device="cuda").bfloat16()
    b = torch.randn(N, K, device="cuda").bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    ms = time_ms(lambda: F.gemm(a, b, out=out))
    ref_ms = time_ms(lambda: torch.matmul(a, b.t()))
    print(f"PERF gemm_nt 8192^3 ms={ms:.3f} tflops={2 * M * N * K / ms / 1e9:.1f} cublas_ms={ref_ms:.3f} "
          f"cublas_tflops={2 * M * N * K / ref_ms / 1e9:.1f}", flush=True)
    return ok


@case
def gemm_f32out():
    F = _F()
    a = torch.randn(300, 192, device="cuda").bfloat16()
    b = torch.randn(100, 192, device="cuda").bfloat16()
    out = F.gemm(a, b, out_dtype=torch.float32)
    return report("gemm_f32out", rel_err(out, a.float() @ b.float().t()), 2e-3)


@case
def gemm_nn():
    F = _F()
    ok = True
    for (M, N, K) in [(128, 64, 64), (256, 192, 128), (1000, 800, 504), (512, 512, 16)]:
        a = torch.randn(M, K, device="cuda").bfloat16()
        kb = (K + 7) // 8 * 8
        b = torch.randn(K, N, device="cuda").bfloat16()      # stored [K][N] (MN-major)
        out = F.gemm(a, b, b_mn_major=True)
        ref = a.float() @ b.float()
        ok &= report(f"gemm_nn[{M}x{N}x{K}]", rel_err(out, ref), 1.5e-2)
    return ok


@case
def gemm_tn():
    F = _F()
    ok = True
    for (M, N, K) in [(128, 64, 64), (64, 64, 256), (504, 800, 1024), (16, 504, 256), (512, 256, 65536)]:
        a = torch.randn(K, M, device="cuda").bfloat16()      # stored [K][M]
        b = torch.randn(K, N, device="cuda").bfloat16()      # stored [K][N]
        out = torch.zeros(M, N, device="cuda")
        F.gemm(a, b, a_mn_major=True, b_mn_major=True, out=out, accumulate=True)
        ref = a.float().t() @ b.float()
        ok &= report(f"gemm_tn[{M}x{N}x{K}]", rel_err(out, ref), 3e-3)
    return ok


def _conv_ref(x_nhwc, w_krsc, pad):
    x = x_nhwc.float().permute(0, 3, 1, 2)
    w = w_krsc.float().permute(0, 3, 1, 2)
    return torch.nn.functional.conv2d(x, w, padding=pad).permute(0, 2, 3, 1).contiguous()


@case
def conv_fprop():
    F = _F()
    ok = True
    for (n, h, c, k) in [(4, 32, 64, 64), (8, 16, 64, 128), (16, 8, 128, 256), (32, 4, 256, 512), (64, 2, 512, 512)]:
        x = torch.randn(n, h, h, c, device="cuda").bfloat16()
        w = (torch.randn(k, 3, 3, c, device="cuda") / (3 * c ** 0.5)).bfloat16()
        out = F.conv_fprop(x, w, pad=1)
        ok &= report(f"conv_fprop[n{n} {h}x{h} c{c} k{k}]", rel_err(out, _conv_ref(x, w, 1)), 1.5e-2)
    # 1x1
    x = torch.randn(8, 16, 16, 128, device="cuda").bfloat16()
    w = (torch.randn(256, 1, 1, 128, device="cuda") / 11).bfloat16()
    ok &= report("conv_fprop[1x1]", rel_err(F.conv_fprop(x, w, pad=0), _conv_ref(x, w, 0)), 1.5e-2)
    n, h, c, k = 256, 32, 64, 64
    x = torch.randn(n, h, h, c, device="cuda").bfloat16()
    w = (torch.randn(k, 3, 3, c, device="cuda") / 24).bfloat16()
    out = torch.empty(n, h, h, k, device="cuda", dtype=torch.bfloat16)
    ms = time_ms(lambda: F.conv_fprop(x, w, pad=1, out=out))
    xc = x.permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last)
    wc = w.permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last)
    cms = time_ms(lambda: torch.nn.functional.conv2d(xc, wc, padding=1))
    fl = 2 * n * h * h * k * 9 * c
    print(f"PERF conv_fprop n256 32x32 c64 k64 ms={ms:.3f} tflops={fl / ms / 1e9:.1f} cudnn_ms={cms:.3f}", flush=True)
    n, h, c, k = 256, 8, 256, 256
    x = torch.randn(n, h, h, c, device="cuda").bfloat16()
    w = (torch.randn(k, 3, 3, c, device="cuda") / 48).bfloat16()
    out = torch.empty(n, h, h, k, device="cuda", dtype=torch.bfloat16)
    ms = time_ms(lambda: F.conv_fprop(x, w, pad=1, out=out))
    xc = x.permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last)
    wc = w.permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last)
    cms = time_ms(lambda: torch.nn.functional.conv2d(xc, wc, padding=1))
    fl = 2 * n * h * h * k * 9 * c
    print(f"PERF conv_fprop n256 8x8 c256 k256 ms={ms:.3f} tflops={fl / ms / 1e9:.1f} cudnn_ms={cms:.3f}", flush=True)
    return ok


@case
def conv_halo_diag():
    """Every filter tap on its own: each is a differently shifted view of the same halo buffer (descriptor start address
    moved by whole 128-byte rows, base-offset field 0)."""
    F = _F()
    n, h, c, k = 2, 32, 64, 64
    ok = True
    x = torch.randn(n, h, h, c, device="cuda").bfloat16()
    for (r, s) in [(1, 1), (1, 0), (1, 2), (0, 1), (2, 1), (0, 0), (2, 2)]:
        w = torch.zeros(k, 3, 3, c, device="cuda")
        w[:, r, s, :] = torch.randn(k, c, device="cuda") / 8
        w = w.bfloat16()
        out = F.conv_halo_fprop(x, w)
        ok = report(f"conv_halo_tap({r},{s})", rel_err(out, _conv_ref(x, w, 1)), 1.5e-2) and ok
    return ok


def _cluster_cases(mode, tag):
    from pytorch_distributed_nn_b200.ops import _lib
    F = _F()
    ok = True
    lib = _lib.lib()
    prev = lib.pdnn_set_cluster_mode(mode)
    try:
        for (M, N, K) in [(1024, 128, 192), (37 * 256, 384, 320), (512, 256, 512), (4096, 512, 1024)]:
            a = torch.randn(M, K, device="cuda").bfloat16()
            b = torch.randn(N, K, device="cuda").bfloat16()
            res = torch.randn(M, N, device="cuda").bfloat16()
            out = F.gemm(a, b, residual=res, relu=True)
            ref = torch.relu(a.float() @ b.float().t() + res.float())
            ok &= report(f"gemm_{tag}[{M}x{N}x{K}]", rel_err(out, ref), 1.5e-2)
        for (n, h, c, k) in [(8, 16, 128, 128), (64, 8, 256, 256), (304, 4, 512, 512)]:
            x = torch.randn(n, h, h, c, device="cuda").bfloat16()
            w = (torch.randn(k, 3, 3, c, device="cuda") / (3 * c ** 0.5)).bfloat16()
            out = F.conv_fprop(x, w, pad=1)
            ok &= report(f"conv_{tag}[n{n} {h}x{h} c{c} k{k}]", rel_err(out, _conv_ref(x, w, 1)), 1.5e-2)
    finally:
        lib.pdnn_set_cluster_mode(prev)
    return ok


@case
def gemm_pair():
    """The CTA-pair schedule (cta_group::2, M = 256 MMAs over both CTAs' shared memory) against the fp32 reference."""
    return _cluster_cases(2, "pair")


@case
def flip_multi():
    """Mirrored filter banks of several convs in one launch (tiled transpose and the element-wise fallback)."""
    F = _F()
    ok = True
    shapes = [(64, 3, 3, 64), (128, 3, 3, 64), (128, 1, 1, 64), (512, 3, 3, 256), (48, 3, 3, 40)]
    total = sum(k * r * s * c for (k, r, s, c) in shapes)
    flat = torch.randn(total + 64, device="cuda").bfloat16()
    rows, outs, off = [], [], 16
    for (k, r, s, c) in shapes:
        o = torch.zeros(c, r, s, k, device="cuda", dtype=torch.bfloat16)
        outs.append(o)
        rows.append([off, o.data_ptr(), k, r, s, c])
        off += k * r * s * c
    table = torch.tensor(rows, dtype=torch.long).cuda()
    F.wgt_flip_transpose_multi(flat, table, len(shapes))
    off = 16
    for (k, r, s, c), o in zip(shapes, outs):
        w = flat[off: off + k * r * s * c].view(k, r, s, c)
        ref = F.wgt_flip_transpose(w)
        tref = w.flip(1, 2).permute(3, 1, 2, 0).contiguous()
        ok &= report(f"flip_multi[{k}x{r}x{s}x{c}]", max(rel_err(o, ref), rel_err(o, tref)), 1e-6)
        off += k * r * s * c
    return ok


@case
def conv_stem():
    """Stem conv (C <= 3 -> 64): output, the im2col by-product and the BatchNorm statistics."""
    F = _F()
    ok = True
    for (n, h, c) in [(8, 32, 3), (4, 16, 3), (2, 8, 3), (16, 8, 1), (8, 16, 2), (300, 32, 3)]:
        x = torch.randn(n, h, h, c, device="cuda").bfloat16()
        w = (torch.randn(64, 3, 3, c, device="cuda") / (3 * c ** 0.5)).bfloat16()
        w_pad = torch.zeros(64, 32, device="cuda", dtype=torch.bfloat16)
        w_pad[:, : 9 * c] = w.reshape(64, -1)
        col = torch.full((n * h * h, 32), 7.0, device="cuda", dtype=torch.bfloat16)
        stats = torch.zeros(2, 64, device="cuda")
        assert F.conv_stem_supported(x.shape, 64)
        out = F.conv_stem(x, w_pad, col=col, bn_stats=stats)
        ref = _conv_ref(x, w, 1)
        ok &= report(f"conv_stem[n{n} {h}x{h} c{c}]", rel_err(out, ref), 1.5e-2)
        col_ref = F.im2col(x, 3, 3, 1, 1, k_pad=32)
        ok &= report(f"conv_stem_col[n{n} {h}x{h} c{c}]", rel_err(col, col_ref), 1e-6)
        of = out.float().reshape(-1, 64)
        ok &= report(f"conv_stem_stats[n{n} {h}x{h} c{c}]", max(rel_err(stats[0], of.sum(0)), rel_err(stats[1], (of * of).sum(0))), 1e-3)
        bias = torch.randn(64, device="cuda")
        out2 = F.conv_stem(x, w_pad, bias=bias, relu=True)
        ok &= report(f"conv_stem_bias_relu[n{n} {h}x{h} c{c}]", rel_err(out2, torch.relu(ref + bias)), 1.5e-2)
    x = torch.randn(1024, 32, 32, 3, device="cuda").bfloat16()
    w_pad = torch.zeros(64, 32, device="cuda", dtype=torch.bfloat16)
    w_pad[:, :27] = (torch.randn(64, 27, device="cuda") / 5).bfloat16()
    out = torch.empty(1024, 32, 32, 64, device="cuda", dtype=torch.bfloat16)
    col = torch.empty(1024 * 32 * 32, 32, device="cuda", dtype=torch.bfloat16)
    stats = torch.zeros(2, 64, device="cuda")
    ms0 = time_ms(lambda: F.conv_stem(x, w_pad, out=out))
    ms1 = time_ms(lambda: F.conv_stem(x, w_pad, out=out, col=col, bn_stats=stats))
    def two_pass():
        c2 = F.im2col(x, 3, 3, 1, 1, k_pad=32, out=col)
        F.gemm(c2, w_pad, out=out.view(-1, 64))
    ms2 = time_ms(two_pass)
    print(f"PERF conv_stem n1024 32x32 c3: plain {ms0 * 1e3:.1f} us, +col+stats {ms1 * 1e3:.1f} us, im2col+GEMM {ms2 * 1e3:.1f} us", flush=True)
    return ok


@case
def conv_halo():
    F = _F()
    ok = True
    # the large-batch entries give every CTA many tiles (pipeline wrap-around, all issuer warps busy)
    for (n, h, c, k) in [(2, 32, 64, 64), (3, 16, 128, 128), (4, 16, 64, 128), (2, 16, 128, 64), (5, 8, 64, 64), (600, 16, 64, 128),
                         (160, 32, 64, 64), (400, 16, 128, 128)]:
        x = torch.randn(n, h, h, c, device="cuda").bfloat16()
        w = (torch.randn(k, 3, 3, c, device="cuda") / (3 * c ** 0.5)).bfloat16()
        bias = torch.randn(k, device="cuda")
        res = torch.randn(n, h, h, k, device="cuda").bfloat16()
        assert F.conv_halo_supported(x.shape, k)
        out = F.conv_halo_fprop(x, w, bias=bias, residual=res, relu=True)
        ref = torch.relu(_conv_ref(x, w, 1) + bias + res.float())
        ok &= report(f"conv_halo[n{n} {h}x{h} c{c} k{k}]", rel_err(out, ref), 1.5e-2)
        # BatchNorm statistics from the epilogue: sum and sum of squares of exactly what was stored
        stats = torch.zeros(2, k, device="cuda")
        out = F.conv_halo_fprop(x, w, bn_stats=stats)
        of = out.float().reshape(-1, k)
        ok &= report(f"conv_halo_stats_sum[n{n} {h}x{h} c{c} k{k}]", rel_err(stats[0], of.sum(0)), 1e-3)
        ok &= report(f"conv_halo_stats_sq[n{n} {h}x{h} c{c} k{k}]", rel_err(stats[1], (of * of).sum(0)), 1e-3)
    for (n, h, c, k) in [(256, 32, 64, 64), (256, 16, 128, 128)]:
        x = torch.randn(n, h, h, c, device="cuda").bfloat16()
        w = (torch.randn(k, 3, 3, c, device="cuda") / 24).bfloat16()
        out = torch.empty(n, h, h, k, device="cuda", dtype=torch.bfloat16)
        ms = time_ms(lambda: F.conv_halo_fprop(x, w, out=out))
        ms_old = time_ms(lambda: F.conv_fprop(x, w, pad=1, out=out))
        xc = x.permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last)
        wc = w.permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last)
        cms = time_ms(lambda: torch.nn.functional.conv2d(xc, wc, padding=1))
        fl = 2 * n * h * h * k * 9 * c
        print(f"PERF conv_halo n{n} {h}x{h} c{c} k{k} halo_ms={ms:.3f} ({fl / ms / 1e9:.0f} TF) generic_ms={ms_old:.3f} cudnn_ms={cms:.3f}", flush=True)
    return ok


@case
def conv_halo_wgrad():
    F = _F()
    ok = True
    for (n, h, c, k) in [(2, 32, 64, 64), (3, 16, 128, 128), (4, 16, 64, 128), (2, 32, 128, 64), (5, 8, 64, 64), (37, 32, 64, 64)]:
        x = torch.randn(n, h, h, c, device="cuda").bfloat16()
        dy = torch.randn(n, h, h, k, device="cuda").bfloat16()
        assert F.conv_halo_wgrad_supported(x.shape, k)
        w = torch.zeros(k, c, 3, 3, device="cuda", requires_grad=True)
        y = torch.nn.functional.conv2d(x.float().permute(0, 3, 1, 2), w, padding=1)
        y.backward(dy.float().permute(0, 3, 1, 2))
        ref = w.grad.permute(0, 2, 3, 1)
        dw = F.conv_halo_wgrad(x, dy)
        ok &= report(f"conv_halo_wgrad[n{n} {h}x{h} c{c} k{k}]", rel_err(dw, ref), 5e-3)
    for (n, h, c, k) in [(256, 32, 64, 64), (256, 16, 128, 128)]:
        x = torch.randn(n, h, h, c, device="cuda").bfloat16()
        dy = torch.randn(n, h, h, k, device="cuda").bfloat16()
        out = torch.zeros(k, 3, 3, c, device="cuda")
        ms = time_ms(lambda: F.conv_halo_wgrad(x, dy, out=out))
        ms_old = time_ms(lambda: F.conv_wgrad(x, dy, R=3, S=3, pad=1, out=out))
        fl = 2 * n * h * h * k * 9 * c
        print(f"PERF conv_halo_wgrad n{n} {h}x{h} c{c} k{k} halo_ms={ms:.3f} ({fl / ms / 1e9:.0f} TF) generic_ms={ms_old:.3f}", flush=True)
    return ok


@case
def conv_strided():
    """Stride-2 convs as implicit GEMM: TMA element strides instead of an im2col buffer."""
    F = _F()
    ok = True
    for (n, h, c, k, r, pad) in [(8, 32, 64, 128, 3, 1), (8, 16, 128, 256, 3, 1), (16, 8, 256, 512, 3, 1), (8, 32, 64, 128, 1, 0),
                                 (16, 8, 256, 512, 1, 0)]:
        x = torch.randn(n, h, h, c, device="cuda").bfloat16()
        w = (torch.randn(k, r, r, c, device="cuda") / (r * c ** 0.5)).bfloat16()
        xr = x.float().permute(0, 3, 1, 2)
        wr = w.float().permute(0, 3, 1, 2).requires_grad_(True)
        yr = torch.nn.functional.conv2d(xr, wr, stride=2, padding=pad)
        y = F.conv_fprop_strided(x, w, pad=pad, stride=2)
        ok &= report(f"conv_s2_fprop[n{n} {h}x{h} c{c} k{k} r{r}]", rel_err(y, yr.permute(0, 2, 3, 1)), 1.5e-2)
        dy = torch.randn(y.shape, device="cuda").bfloat16()
        yr.backward(dy.float().permute(0, 3, 1, 2))
        dw = F.conv_wgrad_strided(x, dy, R=r, S=r, pad=pad, stride=2)
        ok &= report(f"conv_s2_wgrad[n{n} {h}x{h} c{c} k{k} r{r}]", rel_err(dw, wr.grad.permute(0, 2, 3, 1)), 5e-3)
        xg = x.float().permute(0, 3, 1, 2).requires_grad_(True)
        torch.nn.functional.conv2d(xg, w.float().permute(0, 3, 1, 2), stride=2, padding=pad).backward(dy.float().permute(0, 3, 1, 2))
        res = torch.randn(x.shape, device="cuda").bfloat16() if r == 3 else None     # the 1x1 form writes a quarter of dx
        dx = F.conv_dgrad_s2(dy, F.wgt_flip_transpose(w), R=r, S=r, pad=pad, residual=res)
        ref_dx = xg.grad.permute(0, 2, 3, 1) + (res.float() if res is not None else 0)
        ok &= report(f"conv_s2_dgrad[n{n} {h}x{h} c{c} k{k} r{r}]", rel_err(dx, ref_dx), 1.5e-2)
    return ok


@case
def conv_dgrad():
    F = _F()
    ok = True
    for (n, h, c, k) in [(4, 32, 64, 64), (8, 16, 64, 128), (16, 8, 128, 256)]:
        x = torch.randn(n, c, h, h, device="cuda", requires_grad=True)
        w = (torch.randn(k, 3, 3, c, device="cuda") / (3 * c ** 0.5)).bfloat16()
        dy = torch.randn(n, h, h, k, device="cuda").bfloat16()
        y = torch.nn.functional.conv2d(x, w.float().permute(0, 3, 1, 2), padding=1)
        y.backward(dy.float().permute(0, 3, 1, 2))
        ref = x.grad.permute(0, 2, 3, 1)
        wt = F.wgt_flip_transpose(w)
        dx = F.conv_fprop(dy, wt, pad=1)
        ok &= report(f"conv_dgrad[n{n} {h}x{h} c{c} k{k}]", rel_err(dx, ref), 1.5e-2)
    return ok


@case
def conv_wgrad():
    F = _F()
    ok = True
    for (n, h, c, k) in [(4, 32, 64, 64), (8, 16, 64, 128), (16, 8, 128, 256), (32, 4, 256, 512), (64, 32, 64, 64)]:
        x = torch.randn(n, h, h, c, device="cuda").bfloat16()
        dy = torch.randn(n, h, h, k, device="cuda").bfloat16()
        w = torch.zeros(k, c, 3, 3, device="cuda", requires_grad=True)
        y = torch.nn.functional.conv2d(x.float().permute(0, 3, 1, 2), w, padding=1)
        y.backward(dy.float().permute(0, 3, 1, 2))
        ref = w.grad.permute(0, 2, 3, 1)
        dw = F.conv_wgrad(x, dy, R=3, S=3, pad=1)
        ok &= report(f"conv_wgrad[n{n} {h}x{h} c{c} k{k}]", rel_err(dw, ref), 5e-3)
    n, h, c, k = 256, 32, 64, 64
    x = torch.randn(n, h, h, c, device="cuda").bfloat16()
    dy = torch.randn(n, h, h, k, device="cuda").bfloat16()
    out = torch.zeros(k, 3, 3, c, device="cuda")
    ms = time_ms(lambda: F.conv_wgrad(x, dy, R=3, S=3, pad=1, out=out))
    fl = 2 * n * h * h * k * 9 * c
    print(f"PERF conv_wgrad n256 32x32 c64 k64 ms={ms:.3f} tflops={fl / ms / 1e9:.1f}", flush=True)
    return ok


@case
def im2col_paths():
    F = _F()
    ok = True
    # stride-2 3x3 conv lowered: im2col + GEMM, dgrad via GEMM + col2im
    n, h, c, k = 8, 16, 64, 128
    x = torch.randn(n, h, h, c, device="cuda").bfloat16()
    w = (torch.randn(k, 3, 3, c, device="cuda") / 24).bfloat16()
    col = F.im2col(x, 3, 3, 2, 1)
    y = F.gemm(col, w.view(k, -1))
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    wr = w.float().permute(0, 3, 1, 2).requires_grad_(True)
    yr = torch.nn.functional.conv2d(xr, wr, stride=2, padding=1)
    ok &= report("lowered_conv_s2_fwd", rel_err(y.view(n, 8, 8, k), yr.permute(0, 2, 3, 1)), 1.5e-2)
    dy = torch.randn(n, 8, 8, k, device="cuda").bfloat16()
    yr.backward(dy.float().permute(0, 3, 1, 2))
    dcol = F.gemm(dy.view(-1, k), w.view(k, -1), b_mn_major=True)
    dx = F.col2im(dcol, x.shape, 3, 3, 2, 1)
    ok &= report("lowered_conv_s2_dgrad", rel_err(dx, xr.grad.permute(0, 2, 3, 1)), 2e-2)
    dw = torch.zeros(k, 9 * c, device="cuda")
    F.gemm(dy.view(-1, k), col, a_mn_major=True, b_mn_major=True, out=dw, accumulate=True)
    ok &= report("lowered_conv_s2_wgrad", rel_err(dw.view(k, 3, 3, c), wr.grad.permute(0, 2, 3, 1)), 5e-3)
    # stem: C=3 padded K
    x3 = torch.randn(8, 32, 32, 3, device="cuda").bfloat16()
    w3 = (torch.randn(64, 3, 3, 3, device="cuda") / 5).bfloat16()
    col3 = F.im2col(x3, 3, 3, 1, 1, k_pad=64)
    wp = torch.zeros(64, 64, device="cuda", dtype=torch.bfloat16)
    wp[:, :27] = w3.view(64, 27)
    y3 = F.gemm(col3, wp)
    r3 = _conv_ref(x3, w3, 1)
    ok &= report("stem_conv_c3", rel_err(y3.view(8, 32, 32, 64), r3), 1.5e-2)
    # fp32 variants
    xf = torch.randn(4, 12, 12, 20, device="cuda")
    colf = F.im2col(xf, 5, 5, 1, 0)
    wf = torch.randn(50, 5, 5, 20, device="cuda") / 22
    yf = F.gemm_simt(colf, wf.view(50, -1), trans_b=True, M=colf.shape[0], N=50, K=500)
    rf = torch.nn.functional.conv2d(xf.permute(0, 3, 1, 2), wf.permute(0, 3, 1, 2)).permute(0, 2, 3, 1)
    ok &= report("simt_conv5x5_f32", rel_err(yf.view(4, 8, 8, 50), rf), 1e-4)
    return ok


@case
def batchnorm():
    F = _F()
    ok = True
    for (n, h, c) in [(8, 16, 64), (4, 8, 256), (2, 4, 512), (64, 32, 64)]:
        x = (torch.randn(n, h, h, c, device="cuda") * 2 + 0.5).bfloat16()
        res = torch.randn(n, h, h, c, device="cuda").bfloat16()
        gamma = torch.rand(c, device="cuda") + 0.5
        beta = torch.randn(c, device="cuda")
        rm = torch.zeros(c, device="cuda")
        rv = torch.ones(c, device="cuda")
        y, mi = F.bn_forward(x, gamma, beta, relu=True, residual=res, running_mean=rm, running_var=rv)
        xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
        rr = res.float().permute(0, 3, 1, 2).requires_grad_(True)
        g = gamma.clone().requires_grad_(True)
        b = beta.clone().requires_grad_(True)
        rm2 = torch.zeros(c, device="cuda")
        rv2 = torch.ones(c, device="cuda")
        yr = torch.relu(torch.nn.functional.batch_norm(xr, rm2, rv2, g, b, training=True) + rr)
        ok &= report(f"bn_fwd[{n}x{h}x{h}x{c}]", rel_err(y, yr.permute(0, 2, 3, 1)), 1.5e-2)
        ok &= report(f"bn_running[{c}]", max(rel_err(rm, rm2), rel_err(rv, rv2)), 2e-3)
        dy = torch.randn(n, h, h, c, device="cuda").bfloat16()
        yr.backward(dy.float().permute(0, 3, 1, 2))
        dx, dres, dg, db = F.bn_backward(x, y, dy, gamma, mi, relu=True, want_dresidual=True)
        # the kernel's relu mask comes from the bf16 y; compare loosely
        ok &= report(f"bn_bwd_dx[{c}]", rel_err(dx, xr.grad.permute(0, 2, 3, 1)), 3e-2)
        ok &= report(f"bn_bwd_dres[{c}]", rel_err(dres, rr.grad.permute(0, 2, 3, 1)), 3e-2)
        ok &= report(f"bn_bwd_dgamma[{c}]", rel_err(dg, g.grad), 2e-2)
        ok &= report(f"bn_bwd_dbeta[{c}]", rel_err(db, b.grad), 2e-2)
        # no-residual form: the ReLU mask is rebuilt from x (y = None)
        y2, mi2 = F.bn_forward(x, gamma, beta, relu=True)
        xr2 = x.float().permute(0, 3, 1, 2).requires_grad_(True)
        yr2 = torch.relu(torch.nn.functional.batch_norm(xr2, None, None, gamma, beta, training=True))
        yr2.backward(dy.float().permute(0, 3, 1, 2))
        dx2, _, dg2, db2 = F.bn_backward(x, None, dy, gamma, mi2, relu=True, beta=beta)
        ok &= report(f"bn_bwd_maskx_dx[{c}]", rel_err(dx2, xr2.grad.permute(0, 2, 3, 1)), 3e-2)
    return ok


@case
def elementwise():
    F = _F()
    ok = True
    x = torch.randn(4, 8, 8, 64, device="cuda").bfloat16()
    ok &= report("relu", rel_err(F.relu(x), torch.relu(x.float())), 1e-6)
    dy = torch.randn_like(x)
    y = F.relu(x)
    ok &= report("relu_bwd", rel_err(F.relu_backward(y, dy), dy.float() * (x.float() > 0)), 1e-6)
    p = F.maxpool2(x)
    pr = torch.nn.functional.max_pool2d(x.float().permute(0, 3, 1, 2), 2).permute(0, 2, 3, 1)
    ok &= report("maxpool", rel_err(p, pr), 1e-6)
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    pr2 = torch.nn.functional.max_pool2d(xr, 2)
    dp = torch.randn(4, 4, 4, 64, device="cuda").bfloat16()
    pr2.backward(dp.float().permute(0, 3, 1, 2))
    ok &= report("maxpool_bwd", rel_err(F.maxpool2_backward(x, p, dp), xr.grad.permute(0, 2, 3, 1)), 1e-6)
    a = F.avgpool(x)
    ok &= report("avgpool", rel_err(a, x.float().mean(dim=(1, 2))), 1e-2)
    da = torch.randn(4, 64, device="cuda").bfloat16()
    ok &= report("avgpool_bwd", rel_err(F.avgpool_backward(da, 64), (da.float() / 64)[:, None, :].expand(4, 64, 64)), 1e-2)
    d = torch.randn(1000, 24, device="cuda").bfloat16()
    ok &= report("col_sum", rel_err(F.col_sum(d[:, :10], Ccols=10, ld=24), d[:, :10].float().sum(0)), 1e-3)
    f = torch.randn(3, 7, 28, 28, device="cuda")
    ok &= report("nchw_to_nhwc", rel_err(F.nchw_to_nhwc(f, c_pad=8)[..., :7], f.permute(0, 2, 3, 1)), 1e-2)
    f3 = torch.randn(5, 3, 32, 32, device="cuda")          # image staging path: one thread per pixel
    ok &= report("nchw_to_nhwc_rgb", rel_err(F.nchw_to_nhwc(f3), f3.permute(0, 2, 3, 1)), 1e-2)
    x12 = torch.randn(3, 4, 4, 12, device="cuda").bfloat16()   # C % 8 != 0: scalar pooling kernels
    ok &= report("avgpool_c12", rel_err(F.avgpool(x12), x12.float().mean(dim=(1, 2))), 1e-2)
    d12 = torch.randn(3, 12, device="cuda").bfloat16()
    ok &= report("avgpool_bwd_c12", rel_err(F.avgpool_backward(d12, 16), (d12.float() / 16)[:, None, :].expand(3, 16, 12)), 1e-2)
    w = torch.randn(16, 5, 3, 3, device="cuda")
    k = F.oihw_to_krsc(w)
    ok &= report("oihw_to_krsc", rel_err(k, w.permute(0, 2, 3, 1)), 0)
    ok &= report("krsc_to_oihw", rel_err(F.oihw_to_krsc(k, inverse=True), w), 0)
    s = torch.randn(1000, device="cuda").bfloat16()
    ok &= report("sigmoid", rel_err(F.sigmoid(s), torch.sigmoid(s.float())), 1e-2)
    c32 = torch.randn(1001, device="cuda")
    ok &= report("cast", rel_err(F.cast_to_f32(F.cast_to_bf16(c32)), c32.bfloat16().float()), 0)
    return ok


@case
def loss_and_sgd():
    F = _F()
    ok = True
    B, classes, ld = 512, 10, 16
    logits = torch.zeros(B, ld, device="cuda")
    logits[:, :classes] = torch.randn(B, classes, device="cuda") * 3
    lb = logits.bfloat16()
    labels = torch.randint(0, classes, (B,), device="cuda")
    loss, dl, correct = F.cross_entropy(lb, labels, classes)
    lr_ = lb.float()[:, :classes].clone().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(lr_, labels)
    ref.backward()
    ok &= report("ce_loss", abs(float(loss) - float(ref)) / abs(float(ref)), 1e-3)
    ok &= report("ce_grad", rel_err(dl[:, :classes], lr_.grad), 1e-2)
    ok &= report("ce_pad_zero", float(dl[:, classes:].float().abs().max()), 0)
    hits = int((lr_.argmax(1) == labels).sum())
    ok &= report("ce_correct", abs(int(correct) - hits), 0)
    l32 = torch.randn(B, 10, device="cuda")
    tk = F.topk_correct(l32, labels, (1, 5))
    top5 = l32.topk(5, dim=1).indices
    r1 = int((top5[:, 0] == labels).sum())
    r5 = int((top5 == labels[:, None]).any(1).sum())
    ok &= report("topk", abs(int(tk[0]) - r1) + abs(int(tk[1]) - r5), 0)
    n = 100003
    w = torch.randn(n, device="cuda")
    g = torch.randn(n, device="cuda")
    m = torch.zeros(n, device="cuda")
    wr = w.clone().requires_grad_(True)
    opt = torch.optim.SGD([wr], lr=0.1, momentum=0.9)
    wb = torch.empty(n, device="cuda", dtype=torch.bfloat16)
    for it in range(3):
        wr.grad = g.clone() * 0.5
        opt.step()
        F.sgd_momentum(w, g, m, lr=0.1, mu=0.9, scale=0.5, w_bf16=wb)
    ok &= report("sgd_momentum", rel_err(w, wr.detach()), 1e-5)
    ok &= report("sgd_bf16_copy", rel_err(wb, w), 1e-2)
    a = torch.randn(70, 33, device="cuda")
    b = torch.randn(33, 45, device="cuda")
    ok &= report("simt_nn", rel_err(F.gemm_simt(a, b, M=70, N=45, K=33), a @ b), 1e-5)
    b2 = torch.randn(70, 20, device="cuda")
    ok &= report("simt_tn", rel_err(F.gemm_simt(a, b2, trans_a=True, M=33, N=20, K=70), a.t() @ b2), 1e-5)
    return ok


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which == "all":
        rc = 0
        for name in CASES:
            t0 = time.time()
            try:
                r = subprocess.run([sys.executable, os.path.abspath(__file__), name], timeout=240, capture_output=True, text=True)
                out = r.stdout + ("\nSTDERR: " + r.stderr[-3000:] if r.returncode != 0 else "")
                print(out.strip(), flush=True)
                print(f"== {name}: rc={r.returncode} in {time.time() - t0:.1f}s", flush=True)
                rc |= r.returncode != 0
            except subprocess.TimeoutExpired:
                print(f"== {name}: TIMEOUT", flush=True)
                rc = 1
        sys.exit(rc)
    torch.manual_seed(0)
    ok = CASES[which]()
    torch.cuda.synchronize()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
