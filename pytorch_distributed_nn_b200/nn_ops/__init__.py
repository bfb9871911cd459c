"""``accuracy`` and the single-machine ``NN_Trainer`` (reference pytorch_code/nn_ops/__init__.py:1-85).

``NN_Trainer`` keeps the reference contract -- ``NN_Trainer(**kwargs).build_model(); .train(train_loader)`` with the
kwargs ``batch_size, learning_rate, max_epochs, momentum, network`` -- and runs either the torch modules on the CPU
(``device='cpu'``) or the B200 engine + fused momentum-SGD kernel on a GPU. Both paths are the same maths:
forward, mean softmax cross-entropy, backward, torch-semantics momentum SGD.
"""
from __future__ import annotations

import time

import torch
import torch.nn as nn


def accuracy(output, target, topk=(1,)):
    """prec@k in percent for each k (reference nn_ops/__init__.py:13-26). On a GPU this is one kernel."""
    batch = target.size(0)
    if output.is_cuda:
        from ..ops import functional as F
        correct = F.topk_correct(output.float().contiguous(), target.contiguous(), tuple(topk), classes=output.shape[1])
        return [c.float().mul_(100.0 / batch).reshape(1) for c in correct]
    maxk = max(topk)
    _, pred = output.topk(maxk, 1, True, True)
    hit = pred.t().eq(target.view(1, -1).expand_as(pred.t()))
    return [hit[:k].reshape(-1).float().sum(0, keepdim=True).mul_(100.0 / batch) for k in topk]


class NN_Trainer(object):
    def __init__(self, **kwargs):
        self.batch_size = kwargs["batch_size"]
        self.lr = kwargs["learning_rate"]
        self.max_epochs = kwargs["max_epochs"]
        self.momentum = kwargs["momentum"]
        self.network_config = kwargs["network"]
        self.device = kwargs.get("device", "cuda" if torch.cuda.is_available() else "cpu")
        self.max_steps = kwargs.get("max_steps", 0)
        self.log_interval = kwargs.get("log_interval", 10)
        self.history = []

    def build_model(self):
        from ..models import build_model
        if str(self.device).startswith("cuda"):
            from ..engine.network import build_network
            self.engine = build_network(self.network_config, self.batch_size, device=self.device)
            self.engine.init_parameters(seed=1)
            self.momentum_buf = torch.zeros(self.engine.total, device=self.device)
            self.network = None
        else:
            self.engine = None
            self.network = build_model(self.network_config)
            self.optimizer = torch.optim.SGD(self.network.parameters(), lr=self.lr, momentum=self.momentum)
            self.criterion = nn.CrossEntropyLoss()
        return self

    def state_dict(self):
        return self.engine.state_dict() if self.engine is not None else self.network.state_dict()

    def train_step(self, x, y):
        """One optimisation step; returns (loss, prec1, prec5) as floats-on-host only when asked for logging."""
        if self.engine is not None:
            from ..ops import functional as F
            eng = self.engine
            x = x.to(self.device, non_blocking=True).float()
            y = y.to(self.device, non_blocking=True)
            loss = eng.train_step(x.flatten(1) if eng.name.startswith("MLP") else x, y)
            F.sgd_momentum(eng.w32, eng.g, self.momentum_buf, lr=self.lr, mu=self.momentum,
                           w_bf16=eng.w if eng.w.dtype == torch.bfloat16 else None)
            if eng.w.dtype != torch.bfloat16 and eng.w is not eng.w32:
                eng.w[: eng.total].copy_(eng.w32)
            eng.prepare()
            return loss, eng.logits[:, : eng.num_classes]
        self.network.train()
        self.optimizer.zero_grad()
        out = self.network(x.float())
        loss = self.criterion(out, y)
        loss.backward()
        self.optimizer.step()
        return loss.detach(), out.detach()

    def train(self, train_loader):
        step = 0
        for epoch in range(self.max_epochs):
            n_batches = max(1, len(train_loader)) if hasattr(train_loader, "__len__") else 1
            for batch_idx in range(n_batches):
                t0 = time.time()
                x, y = train_loader.next_batch() if hasattr(train_loader, "next_batch") and not hasattr(train_loader, "dataset_next") \
                    else next(iter(train_loader))
                loss, out = self.train_step(x, y)
                step += 1
                if step % self.log_interval == 0 or step == 1:
                    yy = y.to(out.device)
                    p1, p5 = accuracy(out.float(), yy, topk=(1, 5))
                    rec = {"epoch": epoch, "step": step, "loss": float(loss), "prec1": float(p1), "prec5": float(p5),
                           "iter_s": time.time() - t0}
                    self.history.append(rec)
                    print("Train Epoch: {} Step: {} Loss: {:.6f} Prec@1: {:.2f} Prec@5: {:.2f} Time Cost: {:.4f}".format(
                        epoch, step, rec["loss"], rec["prec1"], rec["prec5"], rec["iter_s"]))
                if self.max_steps and step >= self.max_steps:
                    return self.history
        return self.history
