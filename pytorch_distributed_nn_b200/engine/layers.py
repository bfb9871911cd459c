"""Layer objects of the B200 engine: explicit forward/backward on preallocated NHWC buffers.

The reference obtains per-layer control over backward by detaching before every op and replaying the
chain rule by hand (``LeNetSplit.backward`` / ``ResNetSplit.backward``, pytorch_code/model_ops/lenet.py:106-152,
resnet_split.py:235-326) so that each parameter gradient can leave for the master as soon as it exists.
Here that is the native execution model: there is no autograd tape at all. Every layer owns its
output / input-gradient buffers (static shapes, allocated once) and launches hand-written sm_100a kernels;
``backward`` reports each finished parameter gradient to the network's bucket manager, which pushes
completed buckets to the master while the rest of backward is still running.

dtype: ``torch.bfloat16`` layers run on tcgen05 tensor cores (ResNet / VGG / MLP); ``torch.float32`` layers
run the small-shape SIMT path (LeNet: N = 10/20/50, K = 25/500 do not form tensor-core tiles).
"""
from __future__ import annotations

import math
import os
from typing import List, Optional

import torch

from ..ops import functional as F

BF16 = torch.bfloat16
F32 = torch.float32


def _round_up(v: int, m: int) -> int:
    return (v + m - 1) // m * m


class Param:
    """A parameter tensor living inside the network's flat buffers.

    ``native_shape`` is the layout the kernels use (conv: KRSC); ``torch_shape`` is what ``state_dict``
    exposes (conv: OIHW). ``w``/``g`` are views into the flat weight (compute dtype) / gradient (fp32)
    buffers, bound by ``Network.finalize``.
    """

    def __init__(self, name: str, torch_shape, kind: str):
        self.name = name
        self.torch_shape = tuple(torch_shape)
        self.kind = kind  # 'conv' | 'linear' | 'vec' | 'linear_hwc'
        self.numel = int(math.prod(torch_shape))
        self.offset = -1
        self.w: Optional[torch.Tensor] = None
        self.g: Optional[torch.Tensor] = None
        self.index = -1
        self.extra = {}

    @property
    def native_shape(self):
        if self.kind == "conv":
            o, i, r, s = self.torch_shape
            return (o, r, s, i)
        return self.torch_shape


class Layer:
    name = "layer"

    def params(self) -> List[Param]:
        return []

    def prepare(self, net) -> None:
        """Called after new weights landed (once per step): derived weight layouts."""

    def forward(self, x):
        raise NotImplementedError

    def backward(self, dy):
        raise NotImplementedError


class Conv2d(Layer):
    """Convolution without bias (ResNet/VGG style) or with bias (LeNet).

    Three execution modes, chosen from the geometry:
      * ``implicit``: stride 1, 'same' padding, Cin % 64 == 0 -> TMA implicit GEMM on tcgen05
        (fprop, dgrad through mirrored filters, split-K wgrad); no im2col buffer exists.
      * ``lowered``: anything else in bf16 -> im2col + tcgen05 GEMM (stride-2 3x3, 1x1 shortcuts, stems).
      * ``simt``: fp32 networks (LeNet) -> im2col + SIMT GEMM.
    """

    def __init__(self, net, name, cin, cout, k, stride=1, pad=0, bias=False, in_shape=None, first=False):
        self.net = net
        self.name = name
        self.cin, self.cout, self.k, self.stride, self.pad = cin, cout, k, stride, pad
        self.first = first  # no input gradient needed
        n, h, w, c = in_shape
        assert c == cin, (name, c, cin)
        self.in_shape = in_shape
        self.oh = (h + 2 * pad - k) // stride + 1
        self.ow = (w + 2 * pad - k) // stride + 1
        self.out_shape = (n, self.oh, self.ow, cout)
        self.weight = Param(name + ".weight", (cout, cin, k, k), "conv")
        self.bias = Param(name + ".bias", (cout,), "vec") if bias else None
        dt = net.dtype
        dev = net.device
        self.kdim = k * k * cin
        if dt == F32:
            self.mode = "simt"
        elif stride == 1 and 2 * pad == k - 1 and cin % 64 == 0 and self._tileable(n, h, w):
            self.mode = "implicit"
            # fprop / dgrad through the halo-reuse kernel where it applies (C, Cout in {64,128}: the traffic-bound stages)
            # PDNN_HALO_MAX_C (default 64): with 128 input channels only one halo stage per issuer fits beside the resident
            # filters and the kernel measures 140 us against 102 us for the generic implicit GEMM (ResNet-18 layer2, b1024)
            cmax = int(os.environ.get("PDNN_HALO_MAX_C", "64"))
            self.halo_f = net.use_halo and net.device.type == "cuda" and cin <= cmax and F.conv_halo_supported(in_shape, cout, k, k, pad)
            self.halo_b = (not first) and net.use_halo and net.device.type == "cuda" and cout <= cmax and \
                F.conv_halo_supported((n, h, w, cout), cin, k, k, pad)
            self.halo_w = net.use_halo and net.device.type == "cuda" and os.environ.get("PDNN_NO_HALO_WGRAD") != "1" and \
                F.conv_halo_wgrad_supported(in_shape, cout, k, k, pad)
        elif stride == 2 and cin % 64 == 0 and net.use_strided and self._tileable(n, self.oh, self.ow):
            # implicit GEMM with TMA element strides (fprop, wgrad): no im2col matrix; dgrad stays GEMM + col2im
            self.mode = "strided"
        else:
            self.mode = "lowered"
        self.y = torch.empty(self.out_shape, device=dev, dtype=dt)
        self.dx = None if first else torch.empty(in_shape, device=dev, dtype=dt)
        if self.mode == "strided":
            self.kpad = self.kdim
            self.lowered_copy = True      # the dgrad GEMM output (dcol) lives in the shared scratch
        if self.mode == "lowered":
            self.kpad = _round_up(self.kdim, 8)
            self.lowered_copy = not (k == 1 and stride == 1)   # 1x1/s1 reads x in place
            if self.kpad != self.kdim:
                self.w_pad = torch.zeros(cout, self.kpad, device=dev, dtype=dt)
                self.dw_pad = torch.zeros(cout, self.kpad, device=dev, dtype=F32)
        elif self.mode == "simt":
            self.kpad = self.kdim
            self.lowered_copy = True
        # stem (C <= 3 -> 64 channels): im2col rows built in registers inside the conv kernel (conv_stem.cu)
        self.stem = (self.mode == "lowered" and first and net.device.type == "cuda" and self.kpad == 32
                     and os.environ.get("PDNN_NO_STEM") != "1" and F.conv_stem_supported(in_shape, cout, k, k, stride, pad))
        if self.mode in ("implicit", "strided") and not first:
            self.w_t = torch.empty(cin, k, k, cout, device=dev, dtype=dt)
        # stride-2 dgrad by output parity classes (no dcol GEMM + col2im) for the 3x3/pad1 and 1x1/pad0 forms
        self.parity_dgrad = (self.mode == "strided" and not first and cout % 64 == 0 and h == 2 * self.oh and w == 2 * self.ow
                             and ((k == 3 and pad == 1) or (k == 1 and pad == 0)) and self._tileable(n, self.oh, self.ow)
                             and os.environ.get("PDNN_NO_PARITY_DGRAD") != "1")
        self.x = None
        self._colbuf = None

    @staticmethod
    def _tileable(n, h, w):
        hw = h * w
        for pix in (128, 64):
            if w > pix or pix % w:
                return False
            if hw >= pix:
                if hw % pix:
                    return False
            else:
                if pix % hw or n % (pix // hw):
                    return False
        return (n * hw) % 128 == 0

    def params(self):
        return [self.weight] + ([self.bias] if self.bias is not None else [])

    def scratch_elems(self):
        if self.mode == "implicit" or not self.lowered_copy or (self.mode == "strided" and (self.first or self.parity_dgrad)):
            return 0
        n = self.out_shape[0] * self.oh * self.ow
        return n * self.kpad

    def prepare(self, net):
        if self.mode in ("implicit", "strided") and not self.first:
            if net.flip_table is not None:
                return                      # the network mirrors every filter bank in one launch
            F.wgt_flip_transpose(self.weight.w.view(self.cout, self.k, self.k, self.cin), out=self.w_t)
        elif self.mode == "lowered" and self.kpad != self.kdim:
            self.w_pad[:, : self.kdim].copy_(self.weight.w.view(self.cout, self.kdim))

    # -- forward ---------------------------------------------------------------------------------
    def forward(self, x, bn=None):
        """``bn``: the BatchNorm that consumes this conv's output; when the kernel can, it fills ``bn.stats`` from its
        epilogue (``bn.stats_ready``) so that the BatchNorm skips its statistics pass."""
        self.x = x
        w = self.weight.w
        bias = self.bias.wf if self.bias is not None else None
        if bn is not None:
            bn.stats_ready = False
        if self.mode == "implicit":
            if self.halo_f:
                stats = None
                if bn is not None and self.net.fuse_bn_stats:
                    if not self.net.pool_fresh:
                        bn.stats.zero_()
                    bn.stats_ready = True
                    stats = bn.stats
                F.conv_halo_fprop(x, w.view(self.cout, self.k, self.k, self.cin), out=self.y, bias=bias, bn_stats=stats)
            else:
                F.conv_fprop(x, w.view(self.cout, self.k, self.k, self.cin), pad=self.pad, out=self.y, bias=bias)
        elif self.mode == "strided":
            F.conv_fprop_strided(x, w.view(self.cout, self.k, self.k, self.cin), pad=self.pad, stride=self.stride, out=self.y, bias=bias)
        elif self.mode == "lowered" and self.stem:
            col = None
            if self.net.keep_col:        # the kernel writes the im2col matrix the weight-gradient GEMM reads as a by-product
                if self._colbuf is None:
                    self._colbuf = torch.empty(self.out_shape[0] * self.oh * self.ow, self.kpad, device=self.net.device, dtype=self.net.dtype)
                col = self._colbuf
            stats = None
            if bn is not None and self.net.fuse_bn_stats:
                if not self.net.pool_fresh:
                    bn.stats.zero_()
                bn.stats_ready = True
                stats = bn.stats
            F.conv_stem(x, self.w_pad, out=self.y, bias=bias, col=col, bn_stats=stats)
        elif self.mode == "lowered":
            col = self._col(x)
            wmat = self.w_pad if self.kpad != self.kdim else w.view(self.cout, self.kdim)
            F.gemm(col, wmat, out=self.y.view(-1, self.cout), M=col.shape[0], N=self.cout, K=self.kpad, bias=bias)
        else:
            col = self._col(x)
            F.gemm_simt(col, w.view(self.cout, self.kdim), trans_b=True, M=col.shape[0], N=self.cout, K=self.kdim,
                        out=self.y.view(-1, self.cout), bias=bias)
        return self.y

    def _col(self, x, reuse=False):
        """im2col of x. The matrix is kept per layer (net.keep_col) so that backward does not rebuild it;
        with keep_col off it lives in the shared scratch and is recomputed for wgrad."""
        if not self.lowered_copy:
            return x.view(-1, self.cin)
        m = self.out_shape[0] * self.oh * self.ow
        if self.net.keep_col:
            if self._colbuf is None:
                self._colbuf = torch.empty(m, self.kpad, device=self.net.device, dtype=self.net.dtype)
            col = self._colbuf
            if reuse:
                return col
        else:
            col = self.net.scratch(m * self.kpad).view(m, self.kpad)
        F.im2col(x, self.k, self.k, self.stride, self.pad, k_pad=self.kpad, out=col)
        return col

    def _wgrad(self, x, dy, dy2, m):
        net = self.net
        if self.mode == "implicit":
            if self.halo_w:
                F.conv_halo_wgrad(x, dy, out=self.weight.g.view(self.cout, self.k, self.k, self.cin))
            else:
                F.conv_wgrad(x, dy, R=self.k, S=self.k, pad=self.pad, out=self.weight.g.view(self.cout, self.k, self.k, self.cin))
        elif self.mode == "strided":
            F.conv_wgrad_strided(x, dy, R=self.k, S=self.k, pad=self.pad, stride=self.stride,
                                 out=self.weight.g.view(self.cout, self.k, self.k, self.cin))
        elif self.mode == "lowered":
            col = self._col(x, reuse=True)
            if self.kpad != self.kdim:
                self.dw_pad.zero_()
                F.gemm(dy2, col, a_mn_major=True, b_mn_major=True, M=self.cout, N=self.kpad, K=m, out=self.dw_pad, accumulate=True)
                self.weight.g.view(self.cout, self.kdim).add_(self.dw_pad[:, : self.kdim])
            else:
                F.gemm(dy2, col, a_mn_major=True, b_mn_major=True, M=self.cout, N=self.kdim, K=m,
                       out=self.weight.g.view(self.cout, self.kdim), accumulate=True)
        else:
            col = self._col(x, reuse=True)
            F.gemm_simt(dy2, col, trans_a=True, M=self.cout, N=self.kdim, K=m, out=self.weight.g.view(self.cout, self.kdim), accumulate=True)
        if self.bias is not None:
            F.col_sum(dy2, out=self.bias.g, accumulate=True)

    # -- backward --------------------------------------------------------------------------------
    def backward(self, dy, dx_residual=None):
        """dy: [N,OH,OW,Cout]. Returns dx (or None for the first layer). ``dx_residual`` (implicit mode
        only) is added to dx inside the dgrad epilogue."""
        x = self.x
        net = self.net
        m = self.out_shape[0] * self.oh * self.ow
        dy2 = dy.view(m, self.cout)
        # ---- weight gradient first: the deepest parameters must leave for the master as early as possible. It only needs
        # x and dy, like the data gradient below, so it goes to the network's side stream and runs next to the data gradient
        # and the BatchNorm-backward passes that follow on the main stream (tensor-pipe bound next to HBM bound).
        if net.overlap_wgrad and self.mode != "simt" and not self.first:
            with torch.cuda.stream(net.side_stream()):
                self._wgrad(x, dy, dy2, m)
        else:
            self._wgrad(x, dy, dy2, m)
        if self.bias is not None:
            net.grad_ready(self.bias)
        net.grad_ready(self.weight)
        if self.first:
            return None
        # ---- data gradient
        if self.mode == "implicit":
            if self.halo_b:
                F.conv_halo_fprop(dy, self.w_t, out=self.dx, residual=dx_residual)
            else:
                F.conv_fprop(dy, self.w_t, pad=self.pad, out=self.dx, residual=dx_residual)
            return self.dx
        if self.mode == "strided" and self.parity_dgrad:
            if self.k == 1 and dx_residual is not None:
                F.conv_dgrad_s2(dy, self.w_t, R=1, S=1, pad=0, out=self.dx)
                F.add(self.dx, dx_residual, out=self.dx)
            else:
                F.conv_dgrad_s2(dy, self.w_t, R=self.k, S=self.k, pad=self.pad, out=self.dx, residual=dx_residual)
            return self.dx
        if self.mode in ("lowered", "strided"):
            wmat = self.w_pad if self.kpad != self.kdim else self.weight.w.view(self.cout, self.kdim)
            if not self.lowered_copy:
                F.gemm(dy2, wmat, b_mn_major=True, M=m, N=self.cin, K=self.cout, out=self.dx.view(-1, self.cin),
                       residual=dx_residual.view(-1, self.cin) if dx_residual is not None else None)
                return self.dx
            dcol = net.scratch(m * self.kpad).view(m, self.kpad)
            F.gemm(dy2, wmat, b_mn_major=True, M=m, N=self.kpad, K=self.cout, out=dcol)
            F.col2im(dcol, self.in_shape, self.k, self.k, self.stride, self.pad, out=self.dx)
        else:
            dcol = net.scratch(m * self.kpad).view(m, self.kpad)
            F.gemm_simt(dy2, self.weight.w.view(self.cout, self.kdim), M=m, N=self.kdim, K=self.cout, out=dcol)
            F.col2im(dcol, self.in_shape, self.k, self.k, self.stride, self.pad, out=self.dx)
        if dx_residual is not None:
            F.add(self.dx, dx_residual, out=self.dx)
        return self.dx


class BatchNorm(Layer):
    """Training-mode BatchNorm2d fused with the optional residual add and ReLU that follow it."""

    def __init__(self, net, name, c, in_shape, relu=False, eps=1e-5, momentum=0.1):
        self.net, self.name, self.c, self.relu = net, name, c, relu
        self.eps, self.momentum = eps, momentum
        dev = net.device
        self.weight = Param(name + ".weight", (c,), "vec")
        self.bias = Param(name + ".bias", (c,), "vec")
        self.running_mean = torch.zeros(c, device=dev, dtype=F32)
        self.running_var = torch.ones(c, device=dev, dtype=F32)
        self.num_batches_tracked = torch.zeros((), device=dev, dtype=torch.long)
        self.stats = torch.zeros(2, c, device=dev, dtype=F32)
        self.sums = torch.zeros(2, c, device=dev, dtype=F32)
        self.mean_invstd = torch.empty(2, c, device=dev, dtype=F32)
        self.y = torch.empty(in_shape, device=dev, dtype=net.dtype)
        self.dx = torch.empty(in_shape, device=dev, dtype=net.dtype)
        self.dres = None
        self.x = None
        self.out_shape = in_shape
        self.stats_ready = False

    def params(self):
        return [self.weight, self.bias]

    def forward(self, x, residual=None):
        self.x = x
        self._had_residual = residual is not None
        F.bn_forward(x, self.weight.wf, self.bias.wf, relu=self.relu, residual=residual, eps=self.eps,
                     momentum=self.momentum, running_mean=self.running_mean, running_var=self.running_var,
                     stats=self.stats, mean_invstd=self.mean_invstd, out=self.y, stats_ready=self.stats_ready,
                     prezeroed=self.net.pool_fresh)
        self.stats_ready = False
        return self.y

    def backward(self, dy):
        return self.backward_res(dy, False)[0]

    def backward_res(self, dy, want_dresidual=False):
        """Returns (dx, dresidual): dresidual is the gradient w.r.t. the tensor added before the ReLU."""
        if want_dresidual and self.dres is None:
            self.dres = torch.empty_like(self.dx)
        # without a residual the ReLU mask is rebuilt from x inside the kernels: y is not read again
        y = self.y if (self._had_residual or not self.relu or not self.net.bn_mask_from_x) else None
        F.bn_backward(self.x, y, dy, self.weight.wf, self.mean_invstd, relu=self.relu,
                      want_dresidual=want_dresidual, sums=self.sums, dx=self.dx, dresidual=self.dres,
                      dgamma=self.weight.g, dbeta=self.bias.g, beta=self.bias.wf, prezeroed=self.net.pool_fresh)
        self.net.grad_ready(self.bias)
        self.net.grad_ready(self.weight)
        return self.dx, (self.dres if want_dresidual else None)


class ReLU(Layer):
    def __init__(self, net, name, in_shape):
        self.net, self.name = net, name
        self.y = torch.empty(in_shape, device=net.device, dtype=net.dtype)
        self.dx = torch.empty(in_shape, device=net.device, dtype=net.dtype)
        self.out_shape = in_shape

    def forward(self, x):
        return F.relu(x, out=self.y)

    def backward(self, dy):
        return F.relu_backward(self.y, dy, out=self.dx)


class Sigmoid(Layer):
    def __init__(self, net, name, in_shape):
        self.net, self.name = net, name
        self.y = torch.empty(in_shape, device=net.device, dtype=net.dtype)
        self.dx = torch.empty(in_shape, device=net.device, dtype=net.dtype)
        self.out_shape = in_shape

    def forward(self, x):
        return F.sigmoid(x, out=self.y)

    def backward(self, dy):
        return F.sigmoid_backward(self.y, dy, out=self.dx)


class MaxPool2(Layer):
    def __init__(self, net, name, in_shape):
        n, h, w, c = in_shape
        self.net, self.name = net, name
        self.out_shape = (n, h // 2, w // 2, c)
        self.y = torch.empty(self.out_shape, device=net.device, dtype=net.dtype)
        self.dx = torch.empty(in_shape, device=net.device, dtype=net.dtype)
        self.x = None

    def forward(self, x):
        self.x = x
        return F.maxpool2(x, out=self.y)

    def backward(self, dy):
        return F.maxpool2_backward(self.x, self.y, dy.view(self.out_shape), out=self.dx)


class GlobalAvgPool(Layer):
    """avg_pool2d over the whole remaining map followed by flatten (ResNet head)."""

    def __init__(self, net, name, in_shape):
        n, h, w, c = in_shape
        self.net, self.name, self.hw = net, name, h * w
        self.in_shape = in_shape
        self.out_shape = (n, c)
        self.y = torch.empty(n, c, device=net.device, dtype=net.dtype)
        self.dx = torch.empty(in_shape, device=net.device, dtype=net.dtype)

    def forward(self, x):
        return F.avgpool(x, out=self.y)

    def backward(self, dy):
        F.avgpool_backward(dy, self.hw, out=self.dx.view(self.in_shape[0], self.hw, self.in_shape[3]))
        return self.dx


class Flatten(Layer):
    """NHWC flatten. The following Linear is declared ``hwc_from=(C,H,W)`` so that state_dict I/O
    permutes its columns to torch's CHW flatten order."""

    def __init__(self, net, name, in_shape):
        n, h, w, c = in_shape
        self.in_shape = in_shape
        self.out_shape = (n, h * w * c)
        self.name = name

    def forward(self, x):
        return x.view(self.out_shape)

    def backward(self, dy):
        return dy.view(self.in_shape)


class Linear(Layer):
    def __init__(self, net, name, fin, fout, in_shape, bias=True, hwc_from=None, first=False):
        self.net, self.name, self.fin, self.fout, self.first = net, name, fin, fout, first
        n = in_shape[0]
        assert in_shape[1] >= fin
        self.n = n
        dt, dev = net.dtype, net.device
        self.weight = Param(name + ".weight", (fout, fin), "linear_hwc" if hwc_from else "linear")
        if hwc_from:
            self.weight.extra["chw"] = hwc_from
        self.bias = Param(name + ".bias", (fout,), "vec") if bias else None
        self.ld_out = _round_up(fout, 8) if dt == BF16 else fout
        self.out_shape = (n, self.ld_out)
        self.y = torch.zeros(n, self.ld_out, device=dev, dtype=dt)
        self.dx = None if first else torch.zeros(in_shape, device=dev, dtype=dt)
        self.tc = dt == BF16 and fin % 8 == 0
        self.x = None

    def params(self):
        return [self.weight] + ([self.bias] if self.bias is not None else [])

    def forward(self, x):
        self.x = x
        w = self.weight.w.view(self.fout, self.fin)
        b = self.bias.wf if self.bias is not None else None
        if self.tc:
            F.gemm(x, w, M=self.n, N=self.fout, K=self.fin, lda=x.stride(0), out=self.y, bias=b)
        else:
            F.gemm_simt(x, w, trans_b=True, M=self.n, N=self.fout, K=self.fin, lda=x.stride(0), out=self.y, bias=b)
        return self.y

    def backward(self, dy):
        x = self.x
        w = self.weight.w.view(self.fout, self.fin)
        gw = self.weight.g.view(self.fout, self.fin)
        if self.bias is not None:
            F.col_sum(dy, M=self.n, Ccols=self.fout, ld=dy.stride(0), out=self.bias.g, accumulate=True)
            self.net.grad_ready(self.bias)
        if self.tc:
            F.gemm(dy, x, a_mn_major=True, b_mn_major=True, M=self.fout, N=self.fin, K=self.n, lda=dy.stride(0),
                   ldb=x.stride(0), out=gw, accumulate=True)
        else:
            F.gemm_simt(dy, x, trans_a=True, M=self.fout, N=self.fin, K=self.n, lda=dy.stride(0), ldb=x.stride(0),
                        out=gw, accumulate=True)
        self.net.grad_ready(self.weight)
        if self.first:
            return None
        if self.tc:
            F.gemm(dy, w, b_mn_major=True, M=self.n, N=self.fin, K=self.fout, lda=dy.stride(0), ldb=self.fin, out=self.dx,
                   ldo=self.dx.stride(0))
        else:
            F.gemm_simt(dy, w, M=self.n, N=self.fin, K=self.fout, lda=dy.stride(0), ldb=self.fin, out=self.dx, ldc=self.dx.stride(0))
        return self.dx


class BasicBlock(Layer):
    """ResNet basic block (reference: pytorch_code/model_ops/resnet.py:14-36, resnet_split.py:17-87)."""
    expansion = 1

    def __init__(self, net, name, cin, planes, stride, in_shape):
        self.net, self.name = net, name
        self.conv1 = Conv2d(net, f"{name}.conv1", cin, planes, 3, stride, 1, in_shape=in_shape)
        self.bn1 = BatchNorm(net, f"{name}.bn1", planes, self.conv1.out_shape, relu=True)
        self.conv2 = Conv2d(net, f"{name}.conv2", planes, planes, 3, 1, 1, in_shape=self.conv1.out_shape)
        self.bn2 = BatchNorm(net, f"{name}.bn2", planes, self.conv2.out_shape, relu=True)
        self.has_sc = stride != 1 or cin != planes
        if self.has_sc:
            self.sc_conv = Conv2d(net, f"{name}.shortcut.0", cin, planes, 1, stride, 0, in_shape=in_shape)
            self.sc_bn = BatchNorm(net, f"{name}.shortcut.1", planes, self.sc_conv.out_shape, relu=False)
        self.out_shape = self.conv2.out_shape

    def sublayers(self):
        ls = [self.conv1, self.bn1, self.conv2, self.bn2]
        if self.has_sc:
            ls += [self.sc_conv, self.sc_bn]
        return ls

    def params(self):
        return [p for l in self.sublayers() for p in l.params()]

    def prepare(self, net):
        for l in self.sublayers():
            l.prepare(net)

    def scratch_elems(self):
        return max(l.scratch_elems() for l in self.sublayers() if isinstance(l, Conv2d))

    def forward(self, x):
        a1 = self.bn1.forward(self.conv1.forward(x, bn=self.bn1))
        o2 = self.conv2.forward(a1, bn=self.bn2)
        s = self.sc_bn.forward(self.sc_conv.forward(x, bn=self.sc_bn)) if self.has_sc else x
        return self.bn2.forward(o2, residual=s)

    def backward(self, dy):
        do2, dres = self.bn2.backward_res(dy, True)
        if self.has_sc:
            # shortcut parameters sit after bn2 in parameters() order: their gradients leave first
            dso = self.sc_bn.backward(dres)
            dxs = self.sc_conv.backward(dso)
        else:
            dxs = dres
        da1 = self.conv2.backward(do2)
        do1 = self.bn1.backward(da1)
        return self.conv1.backward(do1, dx_residual=dxs)


class Bottleneck(Layer):
    """ResNet bottleneck block (reference: pytorch_code/model_ops/resnet.py:39-64). The reference's split
    variant leaves this block unsplit (resnet_split.py:90-116); here it streams gradients like any other."""
    expansion = 4

    def __init__(self, net, name, cin, planes, stride, in_shape):
        self.net, self.name = net, name
        out = planes * 4
        self.conv1 = Conv2d(net, f"{name}.conv1", cin, planes, 1, 1, 0, in_shape=in_shape)
        self.bn1 = BatchNorm(net, f"{name}.bn1", planes, self.conv1.out_shape, relu=True)
        self.conv2 = Conv2d(net, f"{name}.conv2", planes, planes, 3, stride, 1, in_shape=self.conv1.out_shape)
        self.bn2 = BatchNorm(net, f"{name}.bn2", planes, self.conv2.out_shape, relu=True)
        self.conv3 = Conv2d(net, f"{name}.conv3", planes, out, 1, 1, 0, in_shape=self.conv2.out_shape)
        self.bn3 = BatchNorm(net, f"{name}.bn3", out, self.conv3.out_shape, relu=True)
        self.has_sc = stride != 1 or cin != out
        if self.has_sc:
            self.sc_conv = Conv2d(net, f"{name}.shortcut.0", cin, out, 1, stride, 0, in_shape=in_shape)
            self.sc_bn = BatchNorm(net, f"{name}.shortcut.1", out, self.sc_conv.out_shape, relu=False)
        self.out_shape = self.conv3.out_shape

    def sublayers(self):
        ls = [self.conv1, self.bn1, self.conv2, self.bn2, self.conv3, self.bn3]
        if self.has_sc:
            ls += [self.sc_conv, self.sc_bn]
        return ls

    def params(self):
        return [p for l in self.sublayers() for p in l.params()]

    def prepare(self, net):
        for l in self.sublayers():
            l.prepare(net)

    def scratch_elems(self):
        return max(l.scratch_elems() for l in self.sublayers() if isinstance(l, Conv2d))

    def forward(self, x):
        a1 = self.bn1.forward(self.conv1.forward(x, bn=self.bn1))
        a2 = self.bn2.forward(self.conv2.forward(a1, bn=self.bn2))
        o3 = self.conv3.forward(a2, bn=self.bn3)
        s = self.sc_bn.forward(self.sc_conv.forward(x, bn=self.sc_bn)) if self.has_sc else x
        return self.bn3.forward(o3, residual=s)

    def backward(self, dy):
        do3, dres = self.bn3.backward_res(dy, True)
        if self.has_sc:
            dso = self.sc_bn.backward(dres)
            dxs = self.sc_conv.backward(dso)
        else:
            dxs = dres
        da2 = self.conv3.backward(do3)
        do2 = self.bn2.backward(da2)
        da1 = self.conv2.backward(do2)
        do1 = self.bn1.backward(da1)
        return self.conv1.backward(do1, dx_residual=dxs)
