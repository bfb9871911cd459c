"""All-reduce data parallelism, for comparison with the parameter server
(reference pytorch_code/data_parallel_dist/data_parallel_dist.py: a vendored copy of torch-0.2's
``DistributedDataParallel`` kept as study material -- parameter broadcast at construction, ~1 MB gradient buckets,
one all-reduce per bucket as soon as its gradients exist, buffer broadcast every forward).

This is the same algorithm written against today's ``torch.distributed`` (any backend): it is the ``ddp`` comparator
of BASELINE.md, not a product path -- the product is the parameter server.
"""
from __future__ import annotations

import torch
import torch.distributed as dist
import torch.nn as nn


class DistributedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, bucket_bytes: int = 1 << 20, group=None):
        super().__init__()
        self.module = module
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        # parameters and buffers start identical on every rank (reference :45-46)
        if self.world > 1:
            for t in list(module.state_dict().values()):
                if t.is_floating_point() or t.dtype in (torch.int64, torch.int32):
                    dist.broadcast(t, src=0, group=group)
        # buckets in reverse parameter order: the last layers' gradients are ready first (reference :70-83)
        self.buckets, cur, size = [], [], 0
        for p in reversed([p for p in module.parameters() if p.requires_grad]):
            cur.append(p)
            size += p.numel() * p.element_size()
            if size >= bucket_bytes:
                self.buckets.append(cur)
                cur, size = [], 0
        if cur:
            self.buckets.append(cur)
        self._pending = {}
        self._handles = []
        for bi, bucket in enumerate(self.buckets):
            for p in bucket:
                p.register_post_accumulate_grad_hook(self._make_hook(bi))

    def _make_hook(self, bi):
        def hook(param):
            if self.world == 1:
                return
            left = self._pending.get(bi, len(self.buckets[bi])) - 1
            self._pending[bi] = left
            if left == 0:                                      # bucket complete: reduce it while backward continues
                flat = torch.cat([p.grad.reshape(-1) for p in self.buckets[bi]])
                h = dist.all_reduce(flat, group=self.group, async_op=True)
                self._handles.append((h, flat, bi))
        return hook

    def forward(self, *args, **kwargs):
        self._pending = {}
        if self.world > 1:                                     # buffers (BN statistics) follow rank 0 (reference :126-144)
            for b in self.module.buffers():
                if b.is_floating_point():
                    dist.broadcast(b, src=0, group=self.group)
        return self.module(*args, **kwargs)

    def finish_backward(self):
        """Wait for the bucket reductions and write the averaged gradients back (call after ``loss.backward()``)."""
        for h, flat, bi in self._handles:
            h.wait()
            flat.div_(self.world)
            off = 0
            for p in self.buckets[bi]:
                n = p.numel()
                p.grad.copy_(flat[off:off + n].view_as(p.grad))
                off += n
        self._handles = []
