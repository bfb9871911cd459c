"""The B200 engine's network: flat parameter / gradient buffers, bucketed gradient emission, state_dict I/O.

Parameters live in ONE flat buffer in ``parameters()`` order (same order as the torch reference models, so
parameter index == the reference's tag offset ``88 + l``, pytorch_code/sync_replicas_master_nn.py:278-281).
Three aligned views of that order exist:

    w32   fp32 master copy                (what the master updates and checkpoints)
    w     compute copy (bf16 or fp32)     (what the kernels read; the master's broadcast writes it)
    g     fp32 gradients                  (what wgrad kernels red.add into; what the push kernel sends)

Gradient buckets are contiguous ranges of the flat order, closed from the *end* (deepest layers first),
mirroring the reference's reverse-parameter-order streaming. ``on_bucket`` is invoked
as soon as every gradient of a bucket has been produced, while backward is still running.
"""
from __future__ import annotations

import os
from collections import OrderedDict
from typing import Callable, List, Optional

import torch

from ..ops import functional as F
from . import layers as L

BF16 = torch.bfloat16
F32 = torch.float32
ALIGN = 64  # elements; keeps every parameter 256-byte (fp32) / 128-byte (bf16) aligned for 16 B vectors and TMA


class Bucket:
    def __init__(self, index, start, end, param_ids):
        self.index, self.start, self.end, self.param_ids = index, start, end, param_ids
        self.pending = 0


class Network:
    def __init__(self, name: str, batch: int, in_chw, num_classes: int, dtype=BF16, device="cuda"):
        self.name = name
        self.batch = batch
        self.in_chw = tuple(in_chw)
        self.num_classes = num_classes
        self.dtype = dtype
        self.device = torch.device(device)
        self.layers: List[L.Layer] = []
        self.param_list: List[L.Param] = []
        self.total = 0
        self.w32 = self.w = self.g = None
        self.buckets: List[Bucket] = []
        self.on_bucket: Optional[Callable[[Bucket], None]] = None
        self.layer_hook: Optional[Callable[[int], None]] = None   # called before every backward layer (straggler injection)
        self._scratch = None
        self._param_bucket = {}
        self.loss = torch.zeros(1, device=self.device, dtype=F32)
        self.correct = torch.zeros(1, device=self.device, dtype=torch.int32)
        self.flip_table = None     # finalize(): device table for the one-launch mirrored-filter pass
        self.zpool = None          # finalize(): one buffer for everything that must start a step at zero
        self.pool_fresh = False
        self.dlogits = None
        self.x_in = None
        self.keep_col = True   # keep im2col matrices of lowered convs for backward (memory for time)
        # the halo conv epilogue produces the following BatchNorm's statistics (per-thread column sums over all tiles of
        # a CTA, one cross-lane reduction at the end): ResNet-18 b1024 step 7.41 ms against 7.49 ms with the separate pass.
        # (A first version that reduced across lanes per tile was 2x slower than the separate pass.) PDNN_BN_FUSE=0 disables.
        self.fuse_bn_stats = os.environ.get("PDNN_BN_FUSE", "1") != "0"
        self.bn_mask_from_x = os.environ.get("PDNN_BN_MASK_FROM_X", "1") == "1"   # BN backward without re-reading y (no residual)
        self.use_strided = os.environ.get("PDNN_NO_STRIDED") != "1"   # stride-2 convs as implicit GEMM (TMA element strides)
        self.use_halo = os.environ.get("PDNN_NO_HALO") != "1"   # halo-reuse conv kernel for the C<=128 stages
        # weight-gradient kernels (tensor-pipe bound) on a side stream, next to the data-gradient + BatchNorm-backward chain
        # (HBM bound) of the same layer: both only need dy. The side stream is joined before a bucket is pushed.
        self.batches_tracked = 0        # training steps this network has run (BatchNorm num_batches_tracked of every layer)
        self.overlap_wgrad = os.environ.get("PDNN_WGRAD_OVERLAP", "1") == "1" and self.device.type == "cuda"
        self._side = None
        self._side_event = None
        self._side_busy = False

    # ---- construction ---------------------------------------------------------------------------
    def add(self, layer):
        self.layers.append(layer)
        return layer

    def finalize(self, n_buckets: int = 8, external_w=None, external_g=None):
        """Lay parameters out in the flat buffers, bind views, size scratch and form gradient buckets."""
        off = 0
        for layer in self.layers:
            for p in layer.params():
                p.index = len(self.param_list)
                p.offset = off
                off += (p.numel + ALIGN - 1) // ALIGN * ALIGN
                self.param_list.append(p)
        self.total = off
        dev = self.device
        # zero pool: [loss, correct, then (stats, sums) of every BatchNorm], 64-float aligned slots
        bns = [l for l in self._all_layers() if isinstance(l, L.BatchNorm)]
        zoff = 64
        for bn in bns:
            bn._zoff = zoff
            zoff += (4 * bn.c + 63) // 64 * 64
        if dev.type != "meta":
            self.zpool = torch.zeros(zoff, device=dev, dtype=F32)
            self.loss = self.zpool[0:1]
            self.correct = self.zpool[1:2].view(torch.int32)
            for bn in bns:
                bn.stats = self.zpool[bn._zoff: bn._zoff + 2 * bn.c].view(2, bn.c)
                bn.sums = self.zpool[bn._zoff + 2 * bn.c: bn._zoff + 4 * bn.c].view(2, bn.c)
        self.w32 = torch.zeros(off, device=dev, dtype=F32)
        self.w = external_w if external_w is not None else torch.zeros(off, device=dev, dtype=self.dtype)
        self.g = external_g if external_g is not None else torch.zeros(off, device=dev, dtype=F32)
        assert self.w.numel() >= off and self.g.numel() >= off
        # fp32 shadow of the compute weights: BN scale/shift and biases are consumed in fp32 by the kernels
        self.wf = self.w if self.w.dtype == F32 else torch.zeros(off, device=dev, dtype=F32)
        for p in self.param_list:
            p.w = self.w[p.offset: p.offset + p.numel]
            p.g = self.g[p.offset: p.offset + p.numel]
            p.w32 = self.w32[p.offset: p.offset + p.numel]
            p.wf = self.wf[p.offset: p.offset + p.numel]
        if dev.type == "cuda" and self.w.dtype == BF16:
            rows = []
            for l in self._all_layers():
                if isinstance(l, L.Conv2d) and getattr(l, "w_t", None) is not None and l.mode in ("implicit", "strided") and not l.first:
                    rows.append([l.weight.offset, l.w_t.data_ptr(), l.cout, l.k, l.k, l.cin])
            if rows:
                self._n_flip = len(rows)
                self.flip_table = torch.tensor(rows, dtype=torch.long).to(dev)
        scratch = 0
        for layer in self.layers:
            if hasattr(layer, "scratch_elems"):
                scratch = max(scratch, layer.scratch_elems())
        if scratch:
            self._scratch = torch.empty(scratch, device=dev, dtype=self.dtype)
        self._make_buckets(n_buckets)
        last = self.layers[-1]
        self.dlogits = torch.zeros(last.out_shape, device=dev, dtype=self.dtype)
        return self

    def _make_buckets(self, n_buckets):
        n_buckets = max(1, min(n_buckets, len(self.param_list)))
        target = self.total / n_buckets
        self.buckets = []
        ids = []
        end = self.total
        acc = 0
        for p in reversed(self.param_list):   # closed from the end: deepest layers first
            ids.append(p.index)
            acc += (p.numel + ALIGN - 1) // ALIGN * ALIGN
            if acc >= target and len(self.buckets) < n_buckets - 1:
                self.buckets.append(Bucket(len(self.buckets), p.offset, end, ids))
                end = p.offset
                ids, acc = [], 0
        if ids:
            self.buckets.append(Bucket(len(self.buckets), 0, end, ids))
        for b in self.buckets:
            for i in b.param_ids:
                self._param_bucket[i] = b

    def scratch(self, elems):
        return self._scratch[:elems]

    # ---- parameters -----------------------------------------------------------------------------
    def init_parameters(self, seed: int = 0):
        """Random init with the statistics of torch's default initialisers (kaiming-uniform a=sqrt(5))."""
        gen = torch.Generator(device="cpu").manual_seed(seed)
        host = torch.zeros(self.total, dtype=F32)
        for p in self.param_list:
            if p.kind in ("conv", "linear", "linear_hwc"):
                fan_in = int(p.numel // p.torch_shape[0])
                bound = 1.0 / fan_in ** 0.5
                v = (torch.rand(p.numel, generator=gen) * 2 - 1) * bound
            elif p.name.endswith(".bias") and not self._is_bn(p):
                fan_in = p.extra.get("fan_in", p.numel)
                v = (torch.rand(p.numel, generator=gen) * 2 - 1) / fan_in ** 0.5
            elif p.name.endswith(".weight"):
                v = torch.ones(p.numel)     # BN gamma
            else:
                v = torch.zeros(p.numel)    # BN beta
            host[p.offset: p.offset + p.numel] = v
        self.w32.copy_(host)
        self.sync_compute_weights()

    def _is_bn(self, p):
        return p.kind == "vec" and (".bn" in p.name or "shortcut.1" in p.name or p.extra.get("bn", False))

    def sync_compute_weights(self):
        """Master copy -> compute copy (what the PS broadcast does across GPUs), then derived layouts."""
        if self.w.dtype == BF16:
            F.cast_to_bf16(self.w32, out=self.w[: self.total])
        elif self.w is not self.w32:
            self.w[: self.total].copy_(self.w32)
        self.prepare()

    def prepare(self):
        """Run once per step after new compute weights landed: fp32 shadow + per-layer derived layouts."""
        if self.wf is not self.w:
            F.cast_to_f32(self.w[: self.total], out=self.wf)
        if self.flip_table is not None:
            F.wgt_flip_transpose_multi(self.w, self.flip_table, self._n_flip)
        for layer in self.layers:
            layer.prepare(self)

    def state_dict(self) -> "OrderedDict[str, torch.Tensor]":
        """torch-layout (OIHW / CHW-flatten) fp32 state_dict built from the fp32 master copy."""
        sd = OrderedDict()
        bn_by_prefix = {}
        for layer in self._all_layers():
            if isinstance(layer, L.BatchNorm):
                bn_by_prefix[layer.name] = layer
        for p in self.param_list:
            sd[p.name] = self._export(p)
            if p.name.endswith(".bias"):
                bn = bn_by_prefix.get(p.name[: -len(".bias")])
                if bn is not None:
                    sd[bn.name + ".running_mean"] = bn.running_mean.clone()
                    sd[bn.name + ".running_var"] = bn.running_var.clone()
                    sd[bn.name + ".num_batches_tracked"] = torch.tensor(int(self.batches_tracked), dtype=torch.long)
        return sd

    def load_state_dict(self, sd, strict: bool = True):
        for p in self.param_list:
            if p.name not in sd:
                if strict:
                    raise KeyError(p.name)
                continue
            self._import(p, sd[p.name])
        for layer in self._all_layers():
            if isinstance(layer, L.BatchNorm):
                for k, buf in (("running_mean", layer.running_mean), ("running_var", layer.running_var)):
                    if f"{layer.name}.{k}" in sd:
                        buf.copy_(sd[f"{layer.name}.{k}"])
        self.sync_compute_weights()

    def bn_running_stats(self):
        """All BatchNorm running means / variances / batch counters as ONE flat fp32 CPU tensor (and back): what a training
        rank hands to a master-only rank before a checkpoint, since running statistics only exist where forward runs."""
        parts = []
        for layer in self._all_layers():
            if isinstance(layer, L.BatchNorm):
                parts += [layer.running_mean.float().reshape(-1).cpu(), layer.running_var.float().reshape(-1).cpu()]
        return torch.cat(parts + [torch.tensor([float(self.batches_tracked)])]) if parts else torch.zeros(0)

    def load_bn_running_stats(self, flat):
        off = 0
        for layer in self._all_layers():
            if isinstance(layer, L.BatchNorm):
                c = layer.c
                layer.running_mean.copy_(flat[off: off + c])
                layer.running_var.copy_(flat[off + c: off + 2 * c])
                off += 2 * c
        self.batches_tracked = int(flat[off])

    def _all_layers(self):
        for layer in self.layers:
            if hasattr(layer, "sublayers"):
                yield from layer.sublayers()
            else:
                yield layer

    def _export(self, p, src=None):
        flat = (self.w32 if src is None else src)[p.offset: p.offset + p.numel]
        if p.kind == "conv":
            o, i, r, s = p.torch_shape
            return flat.view(o, r, s, i).permute(0, 3, 1, 2).contiguous()
        if p.kind == "linear_hwc":
            c, h, w = p.extra["chw"]
            return flat.view(p.torch_shape[0], h, w, c).permute(0, 3, 1, 2).reshape(p.torch_shape).contiguous()
        return flat.view(p.torch_shape).clone()

    def _import(self, p, t):
        t = t.to(device=self.device, dtype=F32)
        dst = self.w32[p.offset: p.offset + p.numel]
        if p.kind == "conv":
            dst.copy_(t.permute(0, 2, 3, 1).reshape(-1))
        elif p.kind == "linear_hwc":
            c, h, w = p.extra["chw"]
            dst.copy_(t.view(p.torch_shape[0], c, h, w).permute(0, 2, 3, 1).reshape(-1))
        else:
            dst.copy_(t.reshape(-1))

    def export_grads(self) -> "OrderedDict[str, torch.Tensor]":
        return OrderedDict((p.name, self._export(p, self.g)) for p in self.param_list)

    # ---- execution ------------------------------------------------------------------------------
    def zero_grad(self):
        self.g[: self.total].zero_()
        for b in self.buckets:
            b.pending = len(b.param_ids)

    def side_stream(self):
        """Fork: returns the side stream, ordered behind everything enqueued on the current stream so far."""
        if self._side is None:
            self._side = torch.cuda.Stream(device=self.device)
            self._side_event = torch.cuda.Event()
        cur = torch.cuda.current_stream(self.device)
        self._side_event.record(cur)
        self._side.wait_event(self._side_event)
        self._side_busy = True
        return self._side

    def join_side(self):
        if self._side_busy:
            torch.cuda.current_stream(self.device).wait_stream(self._side)
            self._side_busy = False

    def grad_ready(self, p):
        b = self._param_bucket[p.index]
        b.pending -= 1
        if b.pending == 0 and self.on_bucket is not None:
            self.join_side()            # weight gradients of this bucket may still be running on the side stream
            self.on_bucket(b)

    def stage_input(self, x_nchw_f32):
        """NCHW fp32 images (device) -> the engine's NHWC compute layout."""
        if x_nchw_f32.dim() == 2:
            return x_nchw_f32
        n, c, h, w = x_nchw_f32.shape
        if self.x_in is None:
            self.x_in = torch.empty(n, h, w, c, device=self.device, dtype=self.dtype)
        return F.nchw_to_nhwc(x_nchw_f32, out=self.x_in)

    def forward(self, x_nchw_f32):
        x = self.stage_input(x_nchw_f32)
        layers = self.layers
        if self.zpool is not None:     # BatchNorm sums of both passes, loss and hit counter: one fill per step
            self.zpool.zero_()
            self.pool_fresh = True
        for i, layer in enumerate(layers):
            if isinstance(layer, L.Conv2d) and i + 1 < len(layers) and isinstance(layers[i + 1], L.BatchNorm):
                x = layer.forward(x, bn=layers[i + 1])     # the conv epilogue may produce that BatchNorm's statistics
            else:
                x = layer.forward(x)
        self.logits = x
        return x

    def loss_and_grad(self, labels):
        if not self.pool_fresh:
            self.loss.zero_()
            self.correct.zero_()
        F.cross_entropy(self.logits, labels, self.num_classes, loss=self.loss, dlogits=self.dlogits, correct=self.correct)
        return self.loss

    def backward(self):
        dy = self.dlogits
        for idx, layer in enumerate(reversed(self.layers)):
            if self.layer_hook is not None:     # the reference sleeps before every backward layer (pure_py_code/distributed_worker.py:130-132)
                self.layer_hook(idx)
            dy = layer.backward(dy)
        self.join_side()
        self.pool_fresh = False

    def train_step(self, x_nchw_f32, labels):
        """forward + loss + backward; gradients end up in ``self.g`` (and buckets fire ``on_bucket``)."""
        self.batches_tracked += 1
        self.zero_grad()
        self.forward(x_nchw_f32)
        self.loss_and_grad(labels)
        self.backward()
        return self.loss


class FlatParams:
    """A parameter vector without layers: the object the parameter-server kernels are benchmarked with
    (push / aggregate / broadcast sweeps, BASELINE config 5). ``train_step`` just declares every bucket ready."""

    def __init__(self, n_params: int, dtype=BF16, device="cuda", n_buckets: int = 1, external_w=None, external_g=None):
        self.name = f"Flat{n_params}"
        self.total = (n_params + ALIGN - 1) // ALIGN * ALIGN
        self.dtype = dtype
        self.device = torch.device(device)
        self.in_chw = (1, 1, 1)
        self.num_classes = 1
        dev = self.device
        self.w32 = torch.zeros(self.total, device=dev, dtype=F32)
        self.w = external_w if external_w is not None else torch.zeros(self.total, device=dev, dtype=dtype)
        self.g = external_g if external_g is not None else torch.zeros(self.total, device=dev, dtype=F32)
        self.loss = torch.zeros(1, device=dev, dtype=F32)
        self.on_bucket = None
        self.layer_hook = None
        n_buckets = max(1, n_buckets)
        step = (self.total // n_buckets + ALIGN - 1) // ALIGN * ALIGN
        self.buckets, end = [], self.total
        while end > 0:
            start = max(0, end - step)
            self.buckets.append(Bucket(len(self.buckets), start, end, []))
            end = start
        self.param_list = []

    def init_parameters(self, seed=0):
        if self.device.type != "meta":
            g = torch.Generator(device="cpu").manual_seed(seed)
            self.w32.copy_(torch.randn(self.total, generator=g))
            if self.w.dtype == BF16:
                self.w[: self.total].copy_(self.w32)
            elif self.w is not self.w32:
                self.w[: self.total].copy_(self.w32)

    def prepare(self):
        pass

    def load_state_dict(self, sd, strict=True):
        self.w32.copy_(sd["flat"])

    def state_dict(self):
        return OrderedDict(flat=self.w32.clone())

    def train_step(self, x=None, labels=None):
        for b in self.buckets:
            if self.on_bucket is not None:
                self.on_bucket(b)
        return self.loss


# -------------------------------------------------------------------------------------------------
# builders (architectures: reference pytorch_code/model_ops/lenet.py:12-33, resnet.py:67-113; VGG is the
# north-star addition, kuangliu CIFAR variant with BatchNorm)
# -------------------------------------------------------------------------------------------------
def build_lenet(batch, device="cuda", n_buckets=2, external_w=None, external_g=None) -> Network:
    net = Network("LeNet", batch, (1, 28, 28), 10, dtype=F32, device=device)
    shp = (batch, 28, 28, 1)
    c1 = net.add(L.Conv2d(net, "conv1", 1, 20, 5, 1, 0, bias=True, in_shape=shp, first=True))
    p1 = net.add(L.MaxPool2(net, "pool1", c1.out_shape))
    r1 = net.add(L.ReLU(net, "relu1", p1.out_shape))
    c2 = net.add(L.Conv2d(net, "conv2", 20, 50, 5, 1, 0, bias=True, in_shape=r1.out_shape))
    p2 = net.add(L.MaxPool2(net, "pool2", c2.out_shape))
    r2 = net.add(L.ReLU(net, "relu2", p2.out_shape))
    fl = net.add(L.Flatten(net, "flatten", r2.out_shape))
    f1 = net.add(L.Linear(net, "fc1", 800, 500, fl.out_shape, hwc_from=(50, 4, 4)))
    net.add(L.Linear(net, "fc2", 500, 10, f1.out_shape))
    net.finalize(n_buckets, external_w=external_w, external_g=external_g)
    _set_bias_fan_in(net)
    return net


def _set_bias_fan_in(net):
    by_name = {p.name: p for p in net.param_list}
    for p in net.param_list:
        if p.name.endswith(".bias"):
            w = by_name.get(p.name[:-5] + ".weight")
            if w is not None and w.kind != "vec":
                p.extra["fan_in"] = w.numel // w.torch_shape[0]
            elif w is not None:
                p.extra["bn"] = True


_RESNET_CFG = {
    "ResNet18": (L.BasicBlock, [2, 2, 2, 2]), "ResNet34": (L.BasicBlock, [3, 4, 6, 3]),
    "ResNet50": (L.Bottleneck, [3, 4, 6, 3]), "ResNet101": (L.Bottleneck, [3, 4, 23, 3]),
    "ResNet152": (L.Bottleneck, [3, 8, 36, 3]),
}


def build_resnet(name, batch, device="cuda", n_buckets=8, num_classes=10, external_w=None, external_g=None) -> Network:
    block, nblocks = _RESNET_CFG[name]
    net = Network(name, batch, (3, 32, 32), num_classes, dtype=BF16, device=device)
    shp = (batch, 32, 32, 3)
    c = net.add(L.Conv2d(net, "conv1", 3, 64, 3, 1, 1, in_shape=shp, first=True))
    b = net.add(L.BatchNorm(net, "bn1", 64, c.out_shape, relu=True))
    shp = b.out_shape
    cin = 64
    for stage, (planes, stride0, nb) in enumerate(zip([64, 128, 256, 512], [1, 2, 2, 2], nblocks), start=1):
        for i in range(nb):
            blk = net.add(block(net, f"layer{stage}.{i}", cin, planes, stride0 if i == 0 else 1, shp))
            shp = blk.out_shape
            cin = planes * block.expansion
    ap = net.add(L.GlobalAvgPool(net, "avgpool", shp))
    net.add(L.Linear(net, "linear", cin, num_classes, ap.out_shape))
    net.finalize(n_buckets, external_w=external_w, external_g=external_g)
    _set_bias_fan_in(net)
    return net


_VGG_CFG = {
    "VGG11": [64, "M", 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "VGG13": [64, 64, "M", 128, 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "VGG16": [64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"],
    "VGG19": [64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M", 512, 512, 512, 512, "M"],
}


def build_vgg(name, batch, device="cuda", n_buckets=8, num_classes=10, external_w=None, external_g=None) -> Network:
    net = Network(name, batch, (3, 32, 32), num_classes, dtype=BF16, device=device)
    shp = (batch, 32, 32, 3)
    cin = 3
    idx = 0
    first = True
    for v in _VGG_CFG[name]:
        if v == "M":
            mp = net.add(L.MaxPool2(net, f"features.{idx}", shp))
            shp = mp.out_shape
            idx += 1
        else:
            c = net.add(L.Conv2d(net, f"features.{idx}", cin, v, 3, 1, 1, bias=True, in_shape=shp, first=first))
            b = net.add(L.BatchNorm(net, f"features.{idx + 1}", v, c.out_shape, relu=True))
            b.weight.extra["bn"] = b.bias.extra["bn"] = True
            shp = b.out_shape
            cin = v
            idx += 3    # conv, bn, relu occupy three slots of the torch nn.Sequential
            first = False
    fl = net.add(L.Flatten(net, "flatten", shp))
    net.add(L.Linear(net, "classifier", 512, num_classes, fl.out_shape))
    net.finalize(n_buckets, external_w=external_w, external_g=external_g)
    _set_bias_fan_in(net)
    for layer in net.layers:
        if isinstance(layer, L.BatchNorm):
            layer.weight.extra["bn"] = layer.bias.extra["bn"] = True
    return net


def build_mlp(dims, batch, device="cuda", n_buckets=4, activation="sigmoid", external_w=None, external_g=None) -> Network:
    """Fully connected sigmoid MLP (reference: MPI_code/src/distributed_nn.cpp:35-47 784-500-500-800-800-200-100-100-10;
    pure_py_code/distributed_nn.py:41-96)."""
    net = Network("MLP" + "-".join(map(str, dims)), batch, (1, 1, dims[0]), dims[-1], dtype=F32, device=device)
    shp = (batch, dims[0])
    for i in range(len(dims) - 1):
        fc = net.add(L.Linear(net, f"layers.{i}", dims[i], dims[i + 1], shp, first=(i == 0)))
        shp = fc.out_shape
        if i < len(dims) - 2:
            act = net.add((L.Sigmoid if activation == "sigmoid" else L.ReLU)(net, f"act{i}", shp))
            shp = act.out_shape
    net.finalize(n_buckets, external_w=external_w, external_g=external_g)
    _set_bias_fan_in(net)
    return net


def build_network(name: str, batch: int, device="cuda", n_buckets: int = 8, external_w=None, external_g=None) -> Network:
    key = name.replace("Split", "")
    if key.startswith("Flat"):
        return FlatParams(int(key[4:]), BF16, device, n_buckets, external_w=external_w, external_g=external_g)
    if key in ("LeNet",):
        return build_lenet(batch, device, min(n_buckets, 2), external_w=external_w, external_g=external_g)
    if key in ("ResNet", "Resnet"):
        key = "ResNet18"
    if key in _RESNET_CFG:
        return build_resnet(key, batch, device, n_buckets, external_w=external_w, external_g=external_g)
    if key in _VGG_CFG:
        return build_vgg(key, batch, device, n_buckets, external_w=external_w, external_g=external_g)
    if key in ("FC", "MLP"):
        return build_mlp([784, 500, 500, 800, 800, 200, 100, 100, 10], batch, device, min(n_buckets, 4), external_w=external_w, external_g=external_g)
    if key == "FCDeep":
        from ..models.mlp import DIMS_DEEP
        return build_mlp(list(DIMS_DEEP), batch, device, min(n_buckets, 4), external_w=external_w, external_g=external_g)
    raise ValueError(f"unknown network {name!r}")
