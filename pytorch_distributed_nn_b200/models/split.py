"""Layer-split autograd: run backward one op at a time so that gradients can stream out mid-backward.

The reference implements this separately (and partially) for LeNet and ResNet
(pytorch_code/model_ops/lenet.py:56-221, resnet_split.py:161-440): ``forward`` detaches before every op and
stores ``input[i]``/``output[i]``; ``backward(g, communicator, req_send_check)`` walks the ops last-to-first,
calls ``output[i].backward(grad)``, and ``Isend``s each parameter gradient when its ``.grad`` appears;
``backward_signal_kill`` additionally polls for the master's kill message before every op.

Here one generic mechanism serves every model: a network is an ordered list of :class:`Step` s over named
tape values. A value consumed by several steps (a residual branch point) is ONE detached leaf, so its
``.grad`` accumulates the contributions of all consumers before the producing step is differentiated --
which is what the reference achieves with its in-place ``out += shortcut(x)``.

Gradients are handed to ``communicator.send_grad(param_index, tensor)`` (our Communicator interface) or, for
API fidelity with the reference, to ``communicator.Isend([buf, dtype], dest=0, tag=88+idx)`` when the object
passed is MPI-like. Emission order is strictly reverse ``parameters()`` order per step (bias before weight).
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence

import torch.nn as nn

GRAD_TAG_BASE = 88   # reference: tag = 88 + parameter index (sync_replicas_master_nn.py:278-281)


class Step:
    """One differentiable op of the tape: ``outputs[out_key] = fn(*[leaf[k] for k in in_keys])``."""

    def __init__(self, name: str, fn: Callable, in_keys: Sequence[str], out_key: str, modules: Sequence[nn.Module] = ()):
        self.name, self.fn, self.in_keys, self.out_key = name, fn, list(in_keys), out_key
        self.modules = list(modules)

    def parameters(self):
        for m in self.modules:
            yield from m.parameters()


class SplitNet(nn.Module):
    def __init__(self):
        super().__init__()
        self._steps: List[Step] = []
        self._leaf = {}
        self._out = {}
        self._param_index = {}

    # ---- construction ---------------------------------------------------------------------------
    def set_steps(self, steps: List[Step]):
        self._steps = steps
        self._param_index = {id(p): i for i, p in enumerate(self.parameters())}

    @property
    def full_modules(self):
        """Modules that own parameters, in ``parameters()`` order (reference attribute name)."""
        return [m for s in self._steps for m in s.modules]

    @property
    def fetch_init_channel_index(self):
        """Reference name for '#parameter tensors' (lenet.py:102-104, resnet_split.py:226-233)."""
        return len(self._param_index)

    # ---- forward / backward ---------------------------------------------------------------------
    def forward(self, x):
        self._leaf = {"x": x.detach()}
        self._out = {}
        for st in self._steps:
            ins = [self._leaf[k] for k in st.in_keys]
            out = st.fn(*ins)
            self._out[st.out_key] = out
            self._leaf[st.out_key] = out.detach().requires_grad_(True)
        return self._leaf[self._steps[-1].out_key]

    def _emit(self, communicator, st: Step, req_send_check: list):
        params = list(st.parameters())
        for p in reversed(params):          # bias first, then weight: reverse parameters() order
            if p.grad is None:
                continue
            idx = self._param_index[id(p)]
            if hasattr(communicator, "send_grad"):
                req = communicator.send_grad(idx, p.grad)
            else:                           # MPI-like object (mpi4py API shape)
                buf = p.grad.detach().cpu().numpy().astype("float64")
                req = communicator.Isend([buf, None], dest=0, tag=GRAD_TAG_BASE + idx)
            if req is not None:
                req_send_check.append(req)

    def backward(self, g, communicator=None, req_send_check: Optional[list] = None):
        """Chain rule op by op starting from ``g = dL/d(output)``; returns ``req_send_check``."""
        reqs, _ = self._run_backward(g, communicator, req_send_check, kill_probe=None)
        return reqs

    def backward_signal_kill(self, g, communicator=None, req_send_check: Optional[list] = None, cur_step: int = 0):
        """Like :meth:`backward` but asks ``communicator.should_abort(cur_step)`` before every op and abandons
        the rest of backward when the master closed the step. Returns ``(req_send_check, killed)``."""
        probe = None
        if communicator is not None and hasattr(communicator, "should_abort"):
            probe = lambda: communicator.should_abort(cur_step)  # noqa: E731
        return self._run_backward(g, communicator, req_send_check, kill_probe=probe)

    def backward_timeout_kill(self, g, communicator=None, req_send_check=None, cur_step=0, deadline_s=None):
        """Deadline variant (the reference only stubs it, lenet.py:223-225): abandon once ``deadline_s`` passed."""
        import time
        t_end = None if deadline_s is None else time.monotonic() + deadline_s
        probe = (lambda: time.monotonic() > t_end) if t_end is not None else None
        return self._run_backward(g, communicator, req_send_check, kill_probe=probe)

    def _run_backward(self, g, communicator, req_send_check, kill_probe):
        req_send_check = [] if req_send_check is None else req_send_check
        last_key = self._steps[-1].out_key
        grads = {last_key: g}
        killed = False
        for st in reversed(self._steps):
            if kill_probe is not None and kill_probe():
                killed = True
                break
            out = self._out[st.out_key]
            gout = grads.pop(st.out_key, None)
            if gout is None:
                gout = self._leaf[st.out_key].grad
            if gout is None:
                continue
            out.backward(gout.reshape(out.shape))
            if communicator is not None:
                self._emit(communicator, st, req_send_check)
        return req_send_check, killed
