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


class B200EvaluatorNN(object):
    """In-job evaluator rank on the device backend (reference: MPI_code/src/distributed/evaluator_nn.h:29-60).

    The rank is a passive *receiver* of the parameter server: the master's broadcast (one multicast store, or one more
    unicast store) lands every step's weights in this GPU's landing buffers and raises its ``ready`` word. Every
    ``eval_freq`` steps (and at the end) the evaluator snapshots the landing buffer of that step -- re-reading the
    ready word afterwards, seqlock style, because the master reuses the buffer two steps later -- runs the engine's
    forward + loss kernels on held-out batches and appends ``step time_ms loss err`` to ``time_loss_out_<scheme>``.
    BatchNorm layers normalise with the statistics of the evaluation batch (the engine keeps running statistics only on
    ranks that train), which the output file's header line states."""

    def __init__(self, comm, dataset, **kwargs):
        self.comm, self.dataset, self.kwargs = comm, dataset, kwargs
        self.network_config = kwargs["network"]
        self.batch = int(kwargs.get("test_batch_size", 1000) or 1000)
        self.eval_freq = int(kwargs.get("eval_freq", 0) or 0) or 10
        self.max_steps = int(kwargs.get("max_steps", 0) or 0) or 20000
        self.out_dir = kwargs.get("train_dir") or "outfiles"
        self.max_batches = int(kwargs.get("eval_batches", 2) or 2)
        self.rows = []

    def build_model(self):
        from .parallel.b200_ps import B200ParameterServer
        kw = self.kwargs
        world = self.comm.Get_size()
        colocate = bool(kw.get("colocate_worker", False))
        policy = kw.get("straggler_policy", "kofn")
        n_workers = (world if colocate else world - 1) - 1
        k = kw.get("kill_threshold")
        if policy == "all" or not k:
            k = n_workers
        k = max(1, min(int(k), n_workers))
        interval_ms = float(kw.get("interval_ms", 0.0) or 0.0)
        self.batch = min(self.batch, self.dataset.num_examples) // 32 * 32
        self.ps = B200ParameterServer(self.network_config, self.batch, self.comm.Get_rank(), world, lr=kw["learning_rate"],
                                      momentum=kw["momentum"], num_aggregate=k, avg_mode=kw.get("avg_mode", "k"),
                                      grad_dtype="bf16" if kw.get("grad_dtype", "bf16") == "bf16" else "fp32", colocate=colocate,
                                      n_buckets=int(kw.get("n_buckets", 8) or 8), seed=int(kw.get("seed", 1) or 1),
                                      deadline_ms=interval_ms if policy == "interval" else 0.0, group=self.comm.group,
                                      device=self.comm.device, n_listeners=1)
        self.scheme = "SyncReplicasWithBackup%d_%d" % (k, n_workers)      # reference scheme_full_name(), distributed_defines.h:30-41
        self.network = self.ps.net
        return self

    def _snapshot(self):
        """Copy the newest complete weights into the network; returns the step they belong to."""
        ps = self.ps
        while True:
            r = int(ps.ready_word)
            ps.step = r
            ps.fetch_weights()                      # landing buffer of parity r & 1 -> compute copy (+ derived layouts)
            torch.cuda.current_stream(ps.device).synchronize()
            if int(ps.ready_word) == r:             # the master has not started to refill that buffer
                return r

    def train(self):
        ps, net = self.ps, self.ps.net
        os.makedirs(self.out_dir, exist_ok=True)
        path = os.path.join(self.out_dir, f"time_loss_out_{self.scheme}")
        with open(path, "w") as f:
            f.write("# step time_ms loss err   (BatchNorm: statistics of the evaluation batch)\n")
        t0 = time.monotonic()
        dev = ps.device
        last_eval = -1
        while True:
            r = int(ps.ready_word)                  # weights of step r are complete (r - 1 updates applied)
            updates = r - 1
            due = updates >= self.max_steps or last_eval < 0 or updates // self.eval_freq > last_eval // self.eval_freq
            if r < 1 or not due:
                time.sleep(0.0005)
                continue
            r = self._snapshot()
            updates = r - 1
            tot_loss, wrong, seen = 0.0, 0, 0
            n = self.dataset.num_examples
            for b, start in enumerate(range(0, n - self.batch + 1, self.batch)):
                if b >= self.max_batches:
                    break
                x = torch.from_numpy(self.dataset.images[start:start + self.batch]).float().to(dev)
                y = torch.from_numpy(self.dataset.labels[start:start + self.batch]).long().to(dev)
                net.forward(x)
                net.loss_and_grad(y)
                torch.cuda.current_stream(dev).synchronize()
                tot_loss += float(net.loss) * self.batch
                wrong += self.batch - int(net.correct)
                seen += self.batch
            row = (updates, (time.monotonic() - t0) * 1e3, tot_loss / max(1, seen), wrong / max(1, seen))
            self.rows.append(row)
            with open(path, "a") as f:
                f.write("%d %.3f %.6f %.6f\n" % row)
            last_eval = updates
            if updates >= self.max_steps:
                break
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


def main(argv=None):
    """``python -m pytorch_distributed_nn_b200.evaluator --checkpoint-dir DIR --network N --dataset D``: the stand-alone
    evaluator process (flags of distributed_TF/src/nn_eval.py:36-45: checkpoint_dir, eval_interval_secs, run_once; eval_dir
    becomes ``--out``, a JSON-lines file of (step, loss, precision@1) rows)."""
    import argparse
    import json
    from .utils.data import read_cifar10, read_mnist
    ap = argparse.ArgumentParser(description="poll a training directory and evaluate every new checkpoint")
    ap.add_argument("--checkpoint-dir", "--train-dir", dest="checkpoint_dir", required=True)
    ap.add_argument("--network", default="LeNet")
    ap.add_argument("--dataset", default="MNIST", choices=["MNIST", "Cifar10"])
    ap.add_argument("--data-dir", default=None)
    ap.add_argument("--synthetic", action="store_true", help="random data of the dataset's shape (no files needed)")
    ap.add_argument("--eval-interval-secs", type=float, default=1.0)
    ap.add_argument("--run-once", action="store_true")
    ap.add_argument("--max-polls", type=int, default=None)
    ap.add_argument("--out", default="", help="append one JSON line per evaluated checkpoint")
    a = ap.parse_args(argv)
    if a.dataset == "MNIST":
        data = read_mnist(a.data_dir or "MNIST_data", synthetic=a.synthetic).validation
    else:
        data = read_cifar10(a.data_dir or "cifar10_data", synthetic=a.synthetic).validation
    rows = evaluate_checkpoints(a.checkpoint_dir, a.network, data, a.eval_interval_secs, a.run_once, a.max_polls)
    if a.out:
        with open(a.out, "a") as f:
            for step, loss, prec in rows:
                f.write(json.dumps({"step": step, "loss": loss, "precision_at_1": prec}) + "\n")
    return rows


if __name__ == "__main__":
    main()
