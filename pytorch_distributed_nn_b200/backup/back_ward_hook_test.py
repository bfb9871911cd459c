"""Autograd hook walkthrough: where a parameter server can tap gradients while backward is still running.

Reference: pytorch_code/backup/back_ward_hook_test.py:1-70 — an old-style ``Function`` (a hand-written Linear) plus
``register_backward_hook`` on the module and ``register_hook`` on a tensor, printing when each fires. Same experiment on
today's API: a ``torch.autograd.Function`` with static forward / backward, a full backward hook on the module and a
tensor hook that rescales the gradient (x 0.1, as in the reference). ``run()`` returns the order of events so that the
test can assert it; the split-tape models (models/split.py) are the production form of the same idea.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn


class LinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, inp, weight, bias=None):
        ctx.save_for_backward(inp, weight, bias)
        out = inp.mm(weight.t())
        if bias is not None:
            out = out + bias.unsqueeze(0)
        return out

    @staticmethod
    def backward(ctx, grad_output):
        inp, weight, bias = ctx.saved_tensors
        EVENTS.append("function backward")
        grad_input = grad_output.mm(weight) if ctx.needs_input_grad[0] else None
        grad_weight = grad_output.t().mm(inp) if ctx.needs_input_grad[1] else None
        grad_bias = grad_output.sum(0) if bias is not None and ctx.needs_input_grad[2] else None
        return grad_input, grad_weight, grad_bias


EVENTS = []


class Linear(nn.Module):
    def __init__(self, in_features, out_features, bias=True):
        super().__init__()
        stdv = 1.0 / math.sqrt(in_features)
        self.weight = nn.Parameter(torch.empty(out_features, in_features).uniform_(-stdv, stdv))
        self.bias = nn.Parameter(torch.empty(out_features).uniform_(-stdv, stdv)) if bias else None

    def forward(self, x):
        return LinearFn.apply(x, self.weight, self.bias)


def run(seed=0, log=print):
    """One forward / backward through two custom Linear layers with both kinds of hook installed."""
    torch.manual_seed(seed)
    EVENTS.clear()
    l1, l2 = Linear(8, 6), Linear(6, 4)
    l2.register_full_backward_hook(lambda m, gi, go: EVENTS.append("module hook") or None)
    x = torch.randn(5, 8)
    h = l1(x)
    h.register_hook(lambda g: (EVENTS.append("tensor hook"), g * 0.1)[1])     # the layer below sees a 10x smaller gradient
    out = l2(h)
    out.sum().backward()
    # reference gradient of l1.weight without the hook, for comparison
    expect = (torch.ones(5, 4) @ l2.weight.detach()).t() @ x
    ratio = float((l1.weight.grad / expect).mean())
    for e in EVENTS:
        log(e)
    log(f"l1.weight.grad / unhooked gradient = {ratio:.3f}")
    return list(EVENTS), ratio


if __name__ == "__main__":
    run()
