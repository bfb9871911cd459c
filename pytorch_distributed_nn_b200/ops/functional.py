"""Tensor-level wrappers over the sm_100a kernels (one Python function per C entry point family).

Conventions: activations NHWC bf16 (``[N,H,W,C]`` contiguous), conv weights KRSC (``[Cout,R,S,Cin]``),
parameters / gradients fp32. Every function launches on the current CUDA stream and returns its output
tensor(s); ``out=`` lets the engine pass preallocated buffers so a training step allocates nothing.

These are the ops the reference gets from torch.nn.functional / numpy / CBLAS (pytorch_code/model_ops/lenet.py:27-33,
resnet.py:19-36,94; pure_py_code/nn/nn_layer.py:58-145; MPI_code/src/util/util.h:21-144); the per-op mapping is in the
header of csrc/nn_kernels.cu and in include/pdnn_b200.h.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import Epilogue, call, ptr, stream_ptr

BF16 = torch.bfloat16
F32 = torch.float32

# abort guard registered by the worker role: (closed_word, step_word) device tensors or None
_GUARD = [None, None]


def set_abort_guard(closed_word, step_word) -> None:
    """Every guarded kernel launched afterwards returns immediately once *closed >= *step (GEMM / conv kernels through
    their epilogue arguments, the layer kernels of nn_kernels.cu through the library's per-thread guard)."""
    _GUARD[0], _GUARD[1] = closed_word, step_word
    call("pdnn_set_abort_guard", ptr(closed_word), ptr(step_word))


def _epi(out, ldo, out_mode, relu=False, alpha=1.0, bias=None, residual=None, max_ctas=0):
    e = Epilogue()
    e.out = ptr(out)
    e.ldo = int(ldo)
    e.out_mode = int(out_mode)
    e.relu = int(bool(relu))
    e.alpha = float(alpha)
    e.bias = ptr(bias)
    e.residual = ptr(residual)
    e.guard_closed = ptr(_GUARD[0])
    e.guard_step = ptr(_GUARD[1])
    e.max_ctas = int(max_ctas)
    return e


def _out_mode(out):
    return 0 if out.dtype == BF16 else 1


def gemm(a, b, *, a_mn_major=False, b_mn_major=False, M=None, N=None, K=None, out=None, out_dtype=BF16,
         bias=None, relu=False, residual=None, alpha=1.0, accumulate=False, k_splits=0, lda=None, ldb=None,
         ldo=None, max_ctas=0):
    """D[M,N] = alpha * A @ B^T-style product on tcgen05 tensor cores.

    ``a``: [M,K] (K contiguous) or, with ``a_mn_major``, stored [K,M]. ``b``: [N,K] or, with
    ``b_mn_major``, stored [K,N]. ``accumulate=True`` adds into an fp32 ``out`` with red.add (split-K).
    """
    if M is None:
        M = a.shape[1] if a_mn_major else a.shape[0]
    if K is None:
        K = a.shape[0] if a_mn_major else a.shape[1]
    if N is None:
        N = b.shape[1] if b_mn_major else b.shape[0]
    lda = a.stride(0) if lda is None else lda
    ldb = b.stride(0) if ldb is None else ldb
    if out is None:
        out = torch.zeros(M, N, device=a.device, dtype=F32) if accumulate else torch.empty(M, N, device=a.device, dtype=out_dtype)
    ldo = out.stride(0) if ldo is None else ldo
    mode = 2 if accumulate else _out_mode(out)
    e = _epi(out, ldo, mode, relu, alpha, bias, residual, max_ctas)
    call("pdnn_gemm_bf16", a, int(a_mn_major), lda, b, int(b_mn_major), ldb, M, N, K, C.byref(e), k_splits, stream_ptr())
    return out


def conv_fprop(x, w, *, pad, out=None, bias=None, relu=False, residual=None, out_dtype=BF16):
    """Stride-1 'same' conv. x [N,H,W,C] bf16 (C%64==0), w [Kout,R,S,C] bf16 -> [N,H,W,Kout]."""
    n, h, wd, c = x.shape
    kout, r, s, _ = w.shape
    if out is None:
        out = torch.empty(n, h, wd, kout, device=x.device, dtype=out_dtype)
    e = _epi(out, kout, _out_mode(out), relu, 1.0, bias, residual)
    call("pdnn_conv_fprop_bf16", x, w, n, h, wd, c, kout, r, s, pad, C.byref(e), stream_ptr())
    return out


def conv_fprop_strided(x, w, *, pad, stride, out=None, bias=None, relu=False):
    """Strided conv forward as implicit GEMM (TMA element strides, no im2col). x [N,H,W,C], w [Kout,R,S,C]."""
    n, h, wd, c = x.shape
    kout, r, s, _ = w.shape
    oh, ow = (h + 2 * pad - r) // stride + 1, (wd + 2 * pad - s) // stride + 1
    if out is None:
        out = torch.empty(n, oh, ow, kout, device=x.device, dtype=BF16)
    e = _epi(out, kout, 0, relu, 1.0, bias, None)
    call("pdnn_conv_fprop_strided_bf16", x, w, n, h, wd, c, kout, r, s, pad, stride, C.byref(e), stream_ptr())
    return out


def conv_wgrad_strided(x, dy, *, R, S, pad, stride, out=None, k_splits=0):
    """dw[Kout,R,S,C] (fp32) += weight gradient of the strided conv; out must be zero-filled."""
    n, h, wd, c = x.shape
    kout = dy.shape[-1]
    if out is None:
        out = torch.zeros(kout, R, S, c, device=x.device, dtype=F32)
    e = _epi(out, R * S * c, 2)
    call("pdnn_conv_wgrad_strided_bf16", x, dy, n, h, wd, c, kout, R, S, pad, stride, C.byref(e), k_splits, stream_ptr())
    return out


def conv_dgrad_s2(dy, wt, *, R, S, pad, out=None, residual=None):
    """dx of a stride-2 conv by output parity classes (no col2im). dy [N,OH,OW,Kout], wt = wgt_flip_transpose(w)
    [C,R,S,Kout] -> dx [N,2*OH,2*OW,C]."""
    n, oh, ow, kout = dy.shape
    c = wt.shape[0]
    if out is None:
        out = torch.empty(n, 2 * oh, 2 * ow, c, device=dy.device, dtype=BF16)
    e = _epi(out, c, 0, False, 1.0, None, residual)
    call("pdnn_conv_dgrad_s2_bf16", dy, wt, n, oh, ow, kout, c, R, S, pad, C.byref(e), stream_ptr())
    return out


def conv_stem_supported(x_shape, kout, R=3, S=3, stride=1, pad=1) -> bool:
    n, h, w, c = x_shape
    return bool(_lib.lib().pdnn_conv_stem_supported(n, h, w, c, kout, R, S, stride, pad))


def conv_stem(x, w_pad, *, out=None, bias=None, relu=False, col=None, bn_stats=None):
    """Stem conv (3x3/s1/p1, C <= 3 -> 64 channels): x [N,H,W,C] bf16, w_pad [64,32] bf16 (KRSC rows padded with zeros).
    ``col`` ([N*H*W, 32] bf16) additionally receives the im2col matrix (for the weight-gradient GEMM); ``bn_stats``
    ([2,64] fp32, zeroed by the caller) the per-channel sum / sum of squares of the output."""
    n, h, wd, c = x.shape
    kout = w_pad.shape[0]
    if out is None:
        out = torch.empty(n, h, wd, kout, device=x.device, dtype=BF16)
    e = _epi(out, kout, 0, relu, 1.0, bias, None)
    call("pdnn_conv_stem_bf16", x, w_pad, n, h, wd, c, kout, C.byref(e), col, bn_stats, stream_ptr())
    return out


def conv_halo_supported(x_shape, kout, R=3, S=3, pad=1) -> bool:
    n, h, w, c = x_shape
    return bool(_lib.lib().pdnn_conv_halo_supported(n, h, w, c, kout, R, S, pad))


def conv_halo_fprop(x, w, *, out=None, bias=None, relu=False, residual=None, bn_stats=None):
    """3x3/s1/p1 conv through the halo-reuse kernel (C in {64,128}): x [N,H,W,C] bf16, w [Kout,3,3,C] bf16.
    ``bn_stats`` ([2, Kout] fp32, zeroed by the caller) additionally receives the per-channel sum / sum of squares of
    the output, i.e. the statistics pass of the BatchNorm that follows."""
    n, h, wd, c = x.shape
    kout = w.shape[0]
    if out is None:
        out = torch.empty(n, h, wd, kout, device=x.device, dtype=BF16)
    e = _epi(out, kout, 0, relu, 1.0, bias, residual)
    if bn_stats is None:
        call("pdnn_conv_halo_fprop_bf16", x, w, n, h, wd, c, kout, C.byref(e), stream_ptr())
    else:
        call("pdnn_conv_halo_fprop_stats_bf16", x, w, n, h, wd, c, kout, C.byref(e), bn_stats, stream_ptr())
    return out


def conv_halo_wgrad_supported(x_shape, kout, R=3, S=3, pad=1) -> bool:
    n, h, w, c = x_shape
    return bool(_lib.lib().pdnn_conv_halo_wgrad_supported(n, h, w, c, kout, R, S, pad))


def conv_halo_wgrad(x, dy, out=None):
    """dw[Kout,3,3,C] (fp32) += wgrad of a 3x3/s1/p1 conv through the halo-reuse kernel; out must be zero-filled."""
    n, h, wd, c = x.shape
    kout = dy.shape[-1]
    if out is None:
        out = torch.zeros(kout, 3, 3, c, device=x.device, dtype=F32)
    call("pdnn_conv_halo_wgrad_bf16", x, dy, n, h, wd, c, kout, out, ptr(_GUARD[0]), ptr(_GUARD[1]), 0, stream_ptr())
    return out


def conv_wgrad(x, dy, *, R, S, pad, out=None, k_splits=0):
    """dw[Kout,R,S,C] (fp32) += dy^T * shifted(x); out must be zero-filled (or hold a running sum)."""
    n, h, wd, c = x.shape
    kout = dy.shape[-1]
    if out is None:
        out = torch.zeros(kout, R, S, c, device=x.device, dtype=F32)
    e = _epi(out, R * S * c, 2)
    call("pdnn_conv_wgrad_bf16", x, dy, n, h, wd, c, kout, R, S, pad, C.byref(e), k_splits, stream_ptr())
    return out


def wgt_flip_transpose_multi(w_flat, table, n_layers, blocks_per_layer=64):
    """One launch for all mirrored filter banks of a network (see pdnn_wgt_flip_transpose_multi for the table)."""
    call("pdnn_wgt_flip_transpose_multi", w_flat, table, n_layers, blocks_per_layer, stream_ptr())


def wgt_flip_transpose(w, out=None):
    kout, r, s, c = w.shape
    if out is None:
        out = torch.empty(c, r, s, kout, device=w.device, dtype=w.dtype)
    call("pdnn_wgt_flip_transpose", w, out, kout, r, s, c, stream_ptr())
    return out


def cast_to_bf16(src, out=None):
    if out is None:
        out = torch.empty(src.shape, device=src.device, dtype=BF16)
    call("pdnn_cast_f32_to_bf16", src, out, src.numel(), stream_ptr())
    return out


def cast_to_f32(src, out=None):
    if out is None:
        out = torch.empty(src.shape, device=src.device, dtype=F32)
    call("pdnn_cast_bf16_to_f32", src, out, src.numel(), stream_ptr())
    return out


def nchw_to_nhwc(x, c_pad=None, out=None, dtype=BF16):
    n, c, h, w = x.shape
    c_pad = c if c_pad is None else c_pad
    if out is None:
        out = torch.empty(n, h, w, c_pad, device=x.device, dtype=dtype)
    name = "pdnn_nchw_f32_to_nhwc_bf16" if out.dtype == BF16 else "pdnn_nchw_f32_to_nhwc_f32"
    call(name, x, out, n, c, h, w, c_pad, stream_ptr())
    return out


def oihw_to_krsc(src, out=None, inverse=False):
    """fp32 OIHW -> KRSC (or back with inverse=True; src is then KRSC [K,R,S,C])."""
    if not inverse:
        k, c, r, s = src.shape
        if out is None:
            out = torch.empty(k, r, s, c, device=src.device, dtype=F32)
        call("pdnn_oihw_to_krsc_f32", src, out, k, c, r, s, 1, stream_ptr())
    else:
        k, r, s, c = src.shape
        if out is None:
            out = torch.empty(k, c, r, s, device=src.device, dtype=F32)
        call("pdnn_oihw_to_krsc_f32", src, out, k, c, r, s, 0, stream_ptr())
    return out


def _sfx(t):
    return "bf16" if t.dtype == BF16 else "f32"


def im2col(x, R, S, stride, pad, k_pad=None, out=None):
    n, h, w, c = x.shape
    oh = (h + 2 * pad - R) // stride + 1
    ow = (w + 2 * pad - S) // stride + 1
    k = R * S * c
    k_pad = k if k_pad is None else k_pad
    if out is None:
        out = torch.empty(n * oh * ow, k_pad, device=x.device, dtype=x.dtype)
    call("pdnn_im2col_" + _sfx(x), x, out, n, h, w, c, R, S, stride, pad, oh, ow, k_pad, stream_ptr())
    return out


def col2im(dcol, x_shape, R, S, stride, pad, out=None):
    n, h, w, c = x_shape
    oh = (h + 2 * pad - R) // stride + 1
    ow = (w + 2 * pad - S) // stride + 1
    if out is None:
        out = torch.empty(n, h, w, c, device=dcol.device, dtype=dcol.dtype)
    call("pdnn_col2im_" + _sfx(dcol), dcol, out, n, h, w, c, R, S, stride, pad, oh, ow, dcol.shape[1], stream_ptr())
    return out


def bn_forward(x, gamma, beta, *, relu=False, residual=None, eps=1e-5, momentum=0.1, running_mean=None,
               running_var=None, stats=None, mean_invstd=None, out=None, stats_ready=False, prezeroed=False):
    """Training-mode BatchNorm over all but the last dim of NHWC bf16 x. Returns (y, mean_invstd).
    ``stats_ready``: ``stats`` already holds sum / sum of squares (produced by the conv epilogue): skip that pass.
    ``prezeroed``: the caller cleared ``stats`` already (the engine clears one pool per step instead of 2 fills per layer)."""
    c = x.shape[-1]
    m = x.numel() // c
    if stats is None:
        stats = torch.zeros(2, c, device=x.device, dtype=F32)
    elif not stats_ready and not prezeroed:
        stats.zero_()
    if mean_invstd is None:
        mean_invstd = torch.empty(2, c, device=x.device, dtype=F32)
    if out is None:
        out = torch.empty_like(x)
    st = stream_ptr()
    if not stats_ready:
        call("pdnn_bn_fwd_stats", x, stats, m, c, st)
    call("pdnn_bn_fwd_apply", x, stats, gamma, beta, residual, out, mean_invstd, running_mean, running_var, m, c,
         eps, momentum, int(relu), st)
    return out, mean_invstd


def bn_backward(x, y, dy, gamma, mean_invstd, *, relu=False, want_dresidual=False, sums=None, dx=None,
                dresidual=None, dgamma=None, dbeta=None, beta=None, prezeroed=False):
    """``y=None`` (allowed when the forward had no residual and ``beta`` is given) rebuilds the ReLU mask from x."""
    c = x.shape[-1]
    m = x.numel() // c
    if sums is None:
        sums = torch.zeros(2, c, device=x.device, dtype=F32)
    elif not prezeroed:
        sums.zero_()
    if dx is None:
        dx = torch.empty_like(x)
    if want_dresidual and dresidual is None:
        dresidual = torch.empty_like(x)
    if dgamma is None:
        dgamma = torch.empty(c, device=x.device, dtype=F32)
    if dbeta is None:
        dbeta = torch.empty(c, device=x.device, dtype=F32)
    st = stream_ptr()
    call("pdnn_bn_bwd_reduce", x, y, dy, mean_invstd, sums, m, c, int(relu), gamma, beta, st)
    call("pdnn_bn_bwd_apply", x, y, dy, mean_invstd, gamma, sums, dx, dresidual if want_dresidual else None,
         dgamma, dbeta, m, c, int(relu), beta, st)
    return dx, dresidual, dgamma, dbeta


def relu(x, out=None):
    if out is None:
        out = torch.empty_like(x)
    call("pdnn_relu_fwd" + ("" if x.dtype == BF16 else "_f32"), x, out, x.numel(), stream_ptr())
    return out


def relu_backward(y, dy, out=None):
    if out is None:
        out = torch.empty_like(dy)
    call("pdnn_relu_bwd" + ("" if y.dtype == BF16 else "_f32"), y, dy, out, y.numel(), stream_ptr())
    return out


def add(a, b, out=None):
    if out is None:
        out = torch.empty_like(a)
    call("pdnn_add_bf16", a, b, out, a.numel(), stream_ptr())
    return out


def maxpool2(x, out=None):
    n, h, w, c = x.shape
    if out is None:
        out = torch.empty(n, h // 2, w // 2, c, device=x.device, dtype=x.dtype)
    call("pdnn_maxpool2_fwd" + ("" if x.dtype == BF16 else "_f32"), x, out, n, h, w, c, stream_ptr())
    return out


def maxpool2_backward(x, y, dy, out=None):
    n, h, w, c = x.shape
    if out is None:
        out = torch.zeros_like(x)
    call("pdnn_maxpool2_bwd" + ("" if x.dtype == BF16 else "_f32"), x, y, dy, out, n, h, w, c, stream_ptr())
    return out


def avgpool(x, out=None):
    n, h, w, c = x.shape
    if out is None:
        out = torch.empty(n, c, device=x.device, dtype=x.dtype)
    call("pdnn_avgpool_fwd", x, out, n, h * w, c, stream_ptr())
    return out


def avgpool_backward(dy, hw, out=None):
    n, c = dy.shape
    if out is None:
        out = torch.empty(n, hw, c, device=dy.device, dtype=dy.dtype)
    call("pdnn_avgpool_bwd", dy, out, n, hw, c, stream_ptr())
    return out


def col_sum(dy, M=None, Ccols=None, ld=None, out=None, accumulate=False):
    M = dy.shape[0] if M is None else M
    Ccols = dy.shape[1] if Ccols is None else Ccols
    ld = dy.stride(0) if ld is None else ld
    if out is None:
        out = torch.zeros(Ccols, device=dy.device, dtype=F32)
    call("pdnn_col_sum" + ("" if dy.dtype == BF16 else "_f32"), dy, ld, out, M, Ccols, int(accumulate), stream_ptr())
    return out


def cross_entropy(logits, labels, classes=None, *, loss=None, dlogits=None, correct=None, want_grad=True):
    """Mean softmax-CE and its gradient in one pass. logits [B, ld] (bf16 or fp32), labels int64."""
    b, ld = logits.shape
    classes = ld if classes is None else classes
    if loss is None:
        loss = torch.zeros(1, device=logits.device, dtype=F32)
    if correct is None:
        correct = torch.zeros(1, device=logits.device, dtype=torch.int32)
    if want_grad and dlogits is None:
        dlogits = torch.empty_like(logits)
    name = "pdnn_cross_entropy_fwd_bwd" + ("" if logits.dtype == BF16 else "_f32")
    call(name, logits, labels, loss, dlogits if want_grad else None, correct, b, classes, ld, stream_ptr())
    return loss, dlogits, correct


def topk_correct(logits, labels, ks=(1, 5), classes=None):
    b, ld = logits.shape
    classes = ld if classes is None else classes
    ks_arr = (C.c_int * len(ks))(*ks)
    correct = torch.zeros(len(ks), device=logits.device, dtype=torch.int32)
    call("pdnn_topk_correct", logits, labels, b, classes, ld, ks_arr, len(ks), correct, stream_ptr())
    return correct


def sgd_momentum(w, g, m, *, lr, mu, scale=1.0, first_step=False, w_bf16=None):
    call("pdnn_sgd_momentum", w, g, m, w_bf16, w.numel(), lr, mu, scale, int(first_step), stream_ptr())
    return w


def sigmoid(x, out=None):
    if out is None:
        out = torch.empty_like(x)
    call("pdnn_sigmoid_fwd" + ("" if x.dtype == BF16 else "_f32"), x, out, x.numel(), stream_ptr())
    return out


def sigmoid_backward(y, dy, out=None):
    if out is None:
        out = torch.empty_like(dy)
    call("pdnn_sigmoid_bwd" + ("" if y.dtype == BF16 else "_f32"), y, dy, out, y.numel(), stream_ptr())
    return out


def gemm_simt(a, b, *, trans_a=False, trans_b=False, M, N, K, out=None, bias=None, relu=False, alpha=1.0,
              accumulate=False, lda=None, ldb=None, ldc=None):
    """Generic fp32 GEMM for shapes the tensor-core path cannot take. C = alpha*op(A)@op(B)."""
    lda = a.stride(0) if lda is None else lda
    ldb = b.stride(0) if ldb is None else ldb
    if out is None:
        out = torch.zeros(M, N, device=a.device, dtype=F32)
    ldc = out.stride(0) if ldc is None else ldc
    call("pdnn_gemm_simt_f32", a, int(trans_a), lda, b, int(trans_b), ldb, out, ldc, M, N, K, alpha, bias, int(relu),
         int(accumulate), stream_ptr())
    return out
