"""Worker role: fetch step + weights, forward / backward on a local batch, stream gradients to the master while
backward is still running, abandon the step when the master says so.

Public names follow the reference (pytorch_code/distributed_worker.py): ``ModelBuffer``, ``DistributedWorker`` with
``build_model`` / ``train`` / ``init_recv_buf`` / ``sync_fetch_step`` / ``async_fetch_step`` / ``update_step`` /
``async_fetch_weights_bcast`` / ``async_fetch_weights_async`` / ``model_update``.

Message backends run the reference's loop on ``torch.distributed`` (gradients leave through
``network.backward_signal_kill``, one message per parameter tensor, deepest first, kill probe between ops).
On the b200 backend the worker enqueues kernels only (see parallel/b200_ps.py).
"""
from __future__ import annotations

import time

import torch
import torch.nn as nn

from .nn_ops import NN_Trainer
from .parallel.comm import TERMINATE_STEP
from .parallel.protocol import STEP_START_, TAG_WEIGHT_BASE
from .utils.metrics import JsonlLogger

__all__ = ["ModelBuffer", "DistributedWorker"]


class ModelBuffer(object):
    """Receive buffers for the weights, one per parameter tensor, plus the step each was last refreshed at
    (reference distributed_worker.py:34-53). BN running statistics are local state and never received."""

    def __init__(self, network, wire_dtype=torch.float32):
        self.recv_buf = []
        self.layer_cur_step = []
        for p in network.parameters():
            self.recv_buf.append(torch.zeros(p.numel(), dtype=wire_dtype))
            self.layer_cur_step.append(0)


class DistributedWorker(NN_Trainer):
    _kill_aware = True

    def __init__(self, comm, **kwargs):
        self.comm = comm
        self.world_size = comm.Get_size()
        self.rank = comm.Get_rank()
        self.cur_step = 0
        self.next_step = 0
        self.batch_size = kwargs["batch_size"]
        self.max_epochs = kwargs["max_epochs"]
        self.momentum = kwargs["momentum"]
        self.lr = kwargs["learning_rate"]
        self.network_config = kwargs["network"]
        self.comm_type = kwargs.get("comm_method", "Bcast")
        self.max_steps = int(kwargs.get("max_steps", 0) or 0)
        self.inject_delay = kwargs.get("inject_delay", "")
        self.log_interval = int(kwargs.get("log_interval", 10) or 10)
        self.grad_dtype = kwargs.get("grad_dtype", "fp32")
        self.colocate = bool(kwargs.get("colocate_worker", False))
        self.n_buckets = int(kwargs.get("n_buckets", 8) or 8)
        self.train_dir = kwargs.get("train_dir", "")
        self.has_evaluator = bool(kwargs.get("evaluator", False))
        self.kwargs = kwargs
        self.logger = JsonlLogger(self.train_dir or None, self.rank, "worker")
        self.history = []
        self.killed_steps = []
        self.ps = None

    def build_model(self):
        if self.comm.backend == "b200":
            from .parallel.b200_ps import B200ParameterServer
            from .parallel.delays import DeviceDelay
            kw = self.kwargs
            device_delay = DeviceDelay(self.inject_delay, seed=int(kw.get("seed", 1) or 1)) if self.inject_delay else None
            policy = kw.get("straggler_policy", "kofn" if self._kill_aware else "all")
            n_workers = (self.world_size if self.colocate else self.world_size - 1) - (1 if self.has_evaluator else 0)
            k = kw.get("kill_threshold")
            if policy == "all" or not k:
                k = n_workers
            interval_ms = float(kw.get("interval_ms", 0.0) or 0.0)
            self.ps = B200ParameterServer(self.network_config, self.batch_size, self.rank, self.world_size, lr=self.lr,
                                          momentum=self.momentum, num_aggregate=max(1, min(int(k), n_workers)), avg_mode=kw.get("avg_mode", "k"),
                                          grad_dtype="bf16" if self.grad_dtype == "bf16" else "fp32", colocate=self.colocate,
                                          n_buckets=self.n_buckets, seed=int(kw.get("seed", 1) or 1),
                                          deadline_ms=interval_ms if policy == "interval" else 0.0, device_delay=device_delay,
                                          group=self.comm.group, device=self.comm.device, n_listeners=1 if self.has_evaluator else 0)
            self.network = self.ps.net
            return self
        from .models import build_model
        from .parallel.delays import parse_delay_spec
        self.network = build_model(self.network_config, split=True)
        self.optimizer = torch.optim.SGD(self.network.parameters(), lr=self.lr, momentum=self.momentum)  # zero_grad only
        self.criterion = nn.CrossEntropyLoss()
        self.init_recv_buf()
        self._param_idx = self.network.fetch_init_channel_index - 1          # index of the last parameter tensor
        self.worker_id = self.rank - 1
        self.delay_fn = parse_delay_spec(self.inject_delay) if self.inject_delay else None
        return self

    def init_recv_buf(self):
        self.model_recv_buf = ModelBuffer(self.network, self.comm.wire_dtype)

    # ---------------------------------------------------------------------------------------------------
    def train(self, train_loader):
        if self.ps is not None:
            return self._train_b200(train_loader)
        comm = self.comm
        comm.start_kill_listener()
        self.sync_fetch_step()
        assert self.update_step(), "the first step message must start a new step"
        assert self.cur_step == STEP_START_ or self.cur_step > 0
        first = True
        n_steps = 0
        while True:
            if not first:
                self.async_fetch_step()
                if self.next_step == TERMINATE_STEP:
                    break
                if not self.update_step():
                    continue
            first = False
            comm.cur_step = self.cur_step
            t0 = time.monotonic()
            if self.comm_type == "Bcast":
                self.async_fetch_weights_bcast()
            else:
                self.async_fetch_weights_async()
            t_fetch = time.monotonic()
            x, y = train_loader.next_batch(self.batch_size)
            self.network.train()
            self.optimizer.zero_grad()
            logits = self.network(x.float())                    # a detached leaf: dL/dlogits seeds the manual chain rule
            loss = self.criterion(logits, y)
            loss.backward()
            t_fwd = time.monotonic()
            if self.delay_fn is not None:
                ns = self.delay_fn(self.worker_id, self.cur_step)
                if ns > 0:
                    self._sleep_or_abort(ns / 1e9)
            reqs = []
            if self._kill_aware:
                reqs, killed = self.network.backward_signal_kill(logits.grad, comm, reqs, self.cur_step)
            else:
                reqs, killed = self.network.backward(logits.grad, comm, reqs), False
            comm.wait_all(reqs)
            t_bwd = time.monotonic()
            if killed:
                self.killed_steps.append(self.cur_step)
            rec = {"step": self.cur_step, "loss": float(loss), "killed": bool(killed), "total_s": t_bwd - t0,
                   "fetch_s": t_fetch - t0, "forward_s": t_fwd - t_fetch, "backward_s": t_bwd - t_fwd}
            self.history.append(rec)
            self.logger.log(**rec)
            n_steps += 1
            if self.cur_step % self.log_interval == 0 or n_steps == 1:
                print("Worker: {}, Cur Step: {}, Train Loss: {:.4f}, Killed: {}, Time Cost: {:.4f}, FetchWeight: {:.4f}, "
                      "Forward: {:.4f}, Backward: {:.4f}".format(self.rank, self.cur_step, rec["loss"], killed, rec["total_s"],
                                                                  rec["fetch_s"], rec["forward_s"], rec["backward_s"]), flush=True)
        self._send_tombstones()
        self.logger.close()
        return self.history

    def _sleep_or_abort(self, seconds):
        """Injected straggler delay that still honours the kill signal (the reference sleeps blindly, §5.3)."""
        t_end = time.monotonic() + seconds
        while time.monotonic() < t_end:
            if self._kill_aware and self.comm.should_abort(self.cur_step):
                return
            time.sleep(min(0.001, max(0.0, t_end - time.monotonic())))

    def _send_tombstones(self):
        """End-of-stream marker so that the master's receiver thread for this worker exits."""
        self.comm.wait_all([self.comm.send_grad_eos()])

    def _train_b200(self, train_loader):
        """Same loop as the master's colocated worker: ``B200ParameterServer.fit`` (pinned staging + CUDA-graph replay)."""
        ps = self.ps
        if train_loader is not None and hasattr(train_loader, "pin_memory"):
            train_loader.pin_memory()           # batches are then DMA'd from where they are
        steps = self.max_steps or 20000

        eval_freq = int(self.kwargs.get("eval_freq", 0) or 0)
        # a master-only rank 0 never runs forward, so its BatchNorm running statistics would stay at their initial values in
        # every checkpoint: the first worker hands its own over whenever the master is about to write one
        ships_bn = (not self.colocate) and self.rank == 1 and bool(self.train_dir) and eval_freq > 0 and hasattr(ps.net, "bn_running_stats")

        def on_log(i, loss):
            self.history.append({"step": i, "loss": loss})
            self.logger.log(step=i, loss=loss)              # worker_rank<r>.jsonl, read by tools/plot_study.py / chrome_trace.py
            print("Worker: {}, Cur Step: {}, Train Loss: {:.4f}".format(self.rank, i, loss), flush=True)
            if ships_bn and i % eval_freq == 0:
                import torch.distributed as dist
                flat = ps.net.bn_running_stats()
                if flat.numel():
                    dist.send(flat, dst=0)

        def should_log(i):     # the same predicate as the master's: every rank synchronises with its device at the same steps
            return i % self.log_interval == 0 or (eval_freq and i % eval_freq == 0)

        ps.fit(train_loader, steps, log_interval=self.log_interval, on_log=on_log, should_log=should_log,
               cuda_graph=bool(self.kwargs.get("cuda_graph", True)))
        return self.history

    # ---- message-backend pieces (reference method names) -------------------------------------------------------
    def sync_fetch_step(self):
        self.next_step = self.comm.recv_step(0)

    def async_fetch_step(self):
        self.next_step = self.comm.recv_step(0)

    def update_step(self):
        """Adopt ``next_step`` if it is new (reference :245-249)."""
        changed = self.next_step != self.cur_step and self.next_step != TERMINATE_STEP
        if changed:
            self.cur_step = self.next_step
        return changed

    def async_fetch_weights_bcast(self):
        """Take part in the master's per-tensor broadcasts for every stale layer (reference :229-243)."""
        layers = []
        for l, buf in enumerate(self.model_recv_buf.recv_buf):
            if self.model_recv_buf.layer_cur_step[l] < self.cur_step:
                self.comm.bcast(buf, root=0)
                self.model_recv_buf.layer_cur_step[l] = self.cur_step
                layers.append(buf)
        self.model_update(self.model_recv_buf.recv_buf)

    def async_fetch_weights_async(self):
        """Irecv every stale layer (tag 11+l), then wait (reference :210-227)."""
        reqs = []
        for l, buf in enumerate(self.model_recv_buf.recv_buf):
            if self.model_recv_buf.layer_cur_step[l] < self.cur_step:
                reqs.append(self.comm.irecv(buf, 0, TAG_WEIGHT_BASE + l))
                self.model_recv_buf.layer_cur_step[l] = self.cur_step
        self.comm.wait_all(reqs)
        self.model_update(self.model_recv_buf.recv_buf)

    def model_update(self, weights_to_update):
        """Load the received tensors into the parameters (reference :251-264); buffers other than parameters
        (BN running statistics, num_batches_tracked) are left alone."""
        with torch.no_grad():
            for p, w in zip(self.network.parameters(), weights_to_update):
                assert p.numel() == w.numel()
                p.copy_(w.view_as(p).to(p.dtype))
