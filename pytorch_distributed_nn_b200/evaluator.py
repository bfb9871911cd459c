"""Evaluator role (reference: MPI_code/src/distributed/evaluator_nn.h, distributed_TF/src/nn_eval.py).

Two forms, as in the reference's two stacks:
* in-job rank (C++ form): receives step + weights like a worker, never computes gradients, and appends
  ``step time_ms loss err`` rows to ``outfiles/time_loss_out_<scheme>`` (evaluator_nn.h:29-60);
* stand-alone process (TF form): polls ``train_dir`` for the newest ``model_step_<n>`` checkpoint, restores it and
  prints ``Num examples ... Precision @ 1 ... Loss ...`` (nn_eval.py:49-111).
"""
from __future__ import annotations

import os
import time

import torch
import torch.nn.functional as F

from .parallel.comm import TERMINATE_STEP
from .parallel.protocol import TAG_WEIGHT_BASE
from .utils import checkpoint as ckpt


def evaluate(network, dataset, batch_size=1000, max_batches=None, device="cpu"):
    """Mean loss and error rate over ``dataset`` (a utils.data.DataSet)."""
    network.eval()
    n = dataset.num_examples
    tot_loss, wrong, seen = 0.0, 0, 0
    with torch.no_grad():
        for b, start in enumerate(range(0, n, batch_size)):
            if max_batches is not None and b >= max_batches:
                break
            x = torch.from_numpy(dataset.images[start:start + batch_size]).float().to(device)
            y = torch.from_numpy(dataset.labels[start:start + batch_size]).to(device)
            out = network(x)
            tot_loss += float(F.cross_entropy(out, y, reduction="sum"))
            wrong += int((out.argmax(1) != y).sum())
            seen += x.shape[0]
    network.train()
    return tot_loss / max(1, seen), wrong / max(1, seen), seen


class EvaluatorNN(object):
    """In-job evaluator rank for the message backends."""

    def __init__(self, comm, dataset, **kwargs):
        self.comm = comm
        self.dataset = dataset
        self.network_config = kwargs["network"]
        self.comm_type = kwargs.get("comm_method", "Bcast")
        self.out_dir = kwargs.get("train_dir") or "outfiles"
        self.scheme = kwargs.get("scheme_name", "SyncReplicas")
        self.max_batches = kwargs.get("eval_batches", 2)
        self.rows = []

    def build_model(self):
        from .models import build_model
        self.network = build_model(self.network_config, split=False)
        self.bufs = [torch.zeros(p.numel(), dtype=self.comm.wire_dtype) for p in self.network.parameters()]
        return self

    def train(self):
        os.makedirs(self.out_dir, exist_ok=True)
        path = os.path.join(self.out_dir, f"time_loss_out_{self.scheme}")
        t0 = time.monotonic()
        while True:
            step = self.comm.recv_step(0)
            if step == TERMINATE_STEP:
                break
            if self.comm_type == "Bcast":
                for b in self.bufs:
                    self.comm.bcast(b, root=0)
            else:
                reqs = [self.comm.irecv(b, 0, TAG_WEIGHT_BASE + l) for l, b in enumerate(self.bufs)]
                self.comm.wait_all(reqs)
            with torch.no_grad():
                for p, b in zip(self.network.parameters(), self.bufs):
                    p.copy_(b.view_as(p).to(p.dtype))
            loss, err, _ = evaluate(self.network, self.dataset, max_batches=self.max_batches)
            row = (step, (time.monotonic() - t0) * 1e3, loss, err)
            self.rows.append(row)
            with open(path, "a") as f:
                f.write("%d %.3f %.6f %.6f\n" % row)
        return self.rows


def evaluate_checkpoints(train_dir, network_name, dataset, eval_interval_secs=5.0, run_once=False, max_polls=None):
    """Stand-alone evaluator process: restore the newest checkpoint whenever a new one appears."""
    from .models import build_model
    net = build_model(network_name)
    last = -1
    polls = 0
    results = []
    while True:
        found = ckpt.latest_checkpoint(train_dir)
        if found is not None and found[1] != last:
            sd, step, _ = ckpt.load_checkpoint(found[0])
            net.load_state_dict(sd)
            t0 = time.time()
            loss, err, seen = evaluate(net, dataset)
            last = step
            results.append((step, loss, 1.0 - err))
            print("Num examples: %d Precision @ 1: %.4f Loss: %.4f Time: %.3f (step %d)" % (seen, 1.0 - err, loss, time.time() - t0, step), flush=True)
        polls += 1
        if run_once or (max_polls is not None and polls >= max_polls):
            return results
        time.sleep(eval_interval_secs)
