"""Master role: synchronous SGD with backup workers -- aggregate the first k of n gradient sets, kill the stragglers.

Public names follow the reference (pytorch_code/sync_replicas_master_nn.py): ``STEP_START_``, ``MAX_NUM_ITERATIONS``,
``update_params_dist_version``, ``GradientAccumulator``, ``SyncReplicasMaster_NN`` with ``build_model`` / ``train`` /
``async_bcast_step`` / ``async_bcast_layer_weights_bcast`` / ``async_bcast_layer_weights_async`` /
``async_fetch_gradient_start`` / ``aggregate_gradient`` / ``model_update`` / ``meset_grad_buffer`` /
``send_kill_signal``.

Two execution paths behind the same class:

* message backends (``comm.backend`` = gloo / nccl-ref): the reference's loop restated on ``torch.distributed`` --
  step message, per-tensor weight Bcast / Isend, pre-armed per-(worker, tensor) receives polled in arrival order,
  kill signal to the stragglers, update, next step -- with the defects of SURVEY §2.11 fixed (right buffer per
  arrival, complete sets only, step-stamped messages, 1/k scaling by default, real momentum, clean termination).
* b200 backend: the loop body is ONE kernel (`pdnn_ps_master_step`, see parallel/b200_ps.py); this class only
  enqueues it, checkpoints and logs.
"""
from __future__ import annotations

import queue
import time
from typing import List

import torch

from .nn_ops import NN_Trainer
from .parallel.comm import TERMINATE_STEP, TorchDistCommunicator
from .parallel.protocol import MAX_NUM_ITERATIONS, STEP_START_, TAG_WEIGHT_BASE
from .utils import checkpoint as ckpt
from .utils.metrics import JsonlLogger, StepTimer, Timeline

__all__ = ["STEP_START_", "MAX_NUM_ITERATIONS", "update_params_dist_version", "GradientAccumulator", "SyncReplicasMaster_NN"]


def update_params_dist_version(param, avg_grad, learning_rate):
    """``param -= lr * avg_grad`` in place (reference sync_replicas_master_nn.py:22-28); returns param."""
    param.add_(avg_grad.to(param.dtype).view_as(param), alpha=-learning_rate)
    return param


class GradientAccumulator(object):
    """Receive buffers and counters of the master (reference :30-79).

    ``gradient_aggregator[l][w]`` is the wire buffer of parameter tensor ``l`` from worker ``w``; ``gradient_aggregate_counter[l]`` counts valid arrivals of tensor ``l`` in the
    current step; ``model_index_range`` lists the tensor indices. Buffers are allocated once and reused
    (the reference re-allocates them with ``np.zeros`` every step, defect §2.11-11).
    """

    def __init__(self, module, num_worker, wire_dtype=torch.float32):
        self.gradient_aggregate_counter: List[int] = []
        self.model_index_range: List[int] = []
        self.gradient_aggregator: List[List[torch.Tensor]] = []
        self.shapes = []
        for idx, p in enumerate(module.parameters()):
            self.shapes.append(tuple(p.shape))
            self.gradient_aggregator.append([torch.zeros(p.numel(), dtype=wire_dtype) for _ in range(num_worker)])
            self.gradient_aggregate_counter.append(0)
            self.model_index_range.append(idx)
        self.num_worker = num_worker
        self.received = [[False] * len(self.shapes) for _ in range(num_worker)]

    def meset_everything(self):
        """(sic, reference spelling) reset the per-step bookkeeping; buffers are overwritten by the next receives."""
        self.gradient_aggregate_counter = [0] * len(self.shapes)
        self.received = [[False] * len(self.shapes) for _ in range(self.num_worker)]


class SyncReplicasMaster_NN(NN_Trainer):
    _policy_default = "kofn"

    def __init__(self, comm, **kwargs):
        self.comm = comm
        self.world_size = comm.Get_size()
        self.cur_step = STEP_START_
        self.lr = kwargs["learning_rate"]
        self.momentum = kwargs["momentum"]
        self.network_config = kwargs["network"]
        self.comm_type = kwargs.get("comm_method", "Bcast")
        self.batch_size = kwargs.get("batch_size", 128)
        self.max_epochs = kwargs.get("max_epochs", 100)
        self.has_evaluator = bool(kwargs.get("evaluator", False))
        self.colocate = bool(kwargs.get("colocate_worker", False)) and comm.backend == "b200"
        n_workers = self.world_size - 1 - (1 if self.has_evaluator else 0)
        if self.colocate or self.world_size == 1:
            n_workers += 1
        self._num_grad_to_collect = n_workers                       # reference name: n = #workers
        self.policy = kwargs.get("straggler_policy", self._policy_default)
        k = kwargs.get("kill_threshold")
        if self.policy == "all" or not k:
            k = n_workers
        self._should_kill_threshold = max(1, min(int(k), n_workers))  # reference name: k
        self.avg_mode = kwargs.get("avg_mode", "k")
        self.interval_ms = float(kwargs.get("interval_ms", 0.0) or 0.0)
        self.max_steps = int(kwargs.get("max_steps", 0) or 0) or MAX_NUM_ITERATIONS
        self.train_dir = kwargs.get("train_dir", "")
        self.eval_freq = int(kwargs.get("eval_freq", 0) or 0)
        self.resume = kwargs.get("resume", "")
        self.log_interval = int(kwargs.get("log_interval", 10) or 10)
        self.grad_dtype = kwargs.get("grad_dtype", "fp32")
        self.n_buckets = int(kwargs.get("n_buckets", 8) or 8)
        self.seed = int(kwargs.get("seed", 1) or 1)
        self.inject_delay = kwargs.get("inject_delay", "")
        self.cuda_graph = bool(kwargs.get("cuda_graph", True))
        self.logger = JsonlLogger(self.train_dir or None, comm.Get_rank(), "master")
        self.timeline = Timeline((self.train_dir + "/timeline_master") if self.train_dir else None)
        self.step_log = []          # (step, selected workers, killed workers, gather seconds)
        self.ps = None
        self._pending_ckpt = None

    # ---------------------------------------------------------------------------------------------------
    def build_model(self):
        if self.comm.backend == "b200":
            return self._build_b200()
        from .models import build_model
        torch.manual_seed(self.seed)
        self.network = build_model(self.network_config, split=True)
        self.params = [p for p in self.network.parameters()]
        self.momentum_buffers = [torch.zeros_like(p.data) for p in self.params]
        self.optimizer = None       # the update is applied by model_update (fused momentum SGD), not by an optimizer object
        self.grad_accumulator = GradientAccumulator(self.network, self._num_grad_to_collect, self.comm.wire_dtype)
        self.init_model_shapes()
        self._worker_ranks = [r for r in range(1, self.world_size - (1 if self.has_evaluator else 0))]
        self._evaluator_rank = self.world_size - 1 if self.has_evaluator else None
        if self.resume:
            sd, step, mom = ckpt.load_checkpoint(self.resume)
            self.network.load_state_dict(sd)
            if mom is not None:                 # optimizer state continues too (flat, parameters() order)
                off = 0
                for m in self.momentum_buffers:
                    m.copy_(mom[off: off + m.numel()].view_as(m))
                    off += m.numel()
            self.cur_step = step + 1
        return self

    def _build_b200(self):
        from .parallel.b200_ps import B200ParameterServer
        from .parallel.delays import DeviceDelay
        device_delay = DeviceDelay(self.inject_delay, seed=self.seed) if self.inject_delay else None
        init_sd, init_mom, self.step_offset = None, None, 0
        if self.resume:
            # the device protocol always counts from 1; the global step (checkpoint names) continues from the file
            init_sd, self.step_offset, init_mom = ckpt.load_checkpoint(self.resume)
        self.ps = B200ParameterServer(self.network_config, self.batch_size, self.comm.Get_rank(), self.world_size, lr=self.lr,
                                      momentum=self.momentum, num_aggregate=self._should_kill_threshold, avg_mode=self.avg_mode,
                                      grad_dtype="bf16" if self.grad_dtype == "bf16" else "fp32", colocate=self.colocate or self.world_size == 1,
                                      n_buckets=self.n_buckets, seed=self.seed,
                                      deadline_ms=self.interval_ms if self.policy == "interval" else 0.0, device_delay=device_delay,
                                      group=self.comm.group, device=self.comm.device, init_state_dict=init_sd, init_momentum=init_mom,
                                      n_listeners=1 if self.has_evaluator else 0)
        self.network = self.ps.net
        return self

    def init_model_shapes(self):
        self._model_shapes = [tuple(p.shape) for p in self.network.parameters()]

    # ---------------------------------------------------------------------------------------------------
    def train(self, train_loader=None):
        if self.ps is not None:
            return self._train_b200(train_loader)
        comm: TorchDistCommunicator = self.comm
        n, k = self._num_grad_to_collect, self._should_kill_threshold
        timer = StepTimer(self.batch_size * k)
        self.async_bcast_step()                                            # reference :121
        first = self.cur_step
        for i in range(first, self.max_steps + 1):
            t_step = time.monotonic()
            self.timeline.mark(self.cur_step, "step_start")
            if self.comm_type == "Bcast":
                self.async_bcast_layer_weights_bcast()
            else:
                self.async_bcast_layer_weights_async()
            self.async_fetch_gradient_start()
            complete_order: List[int] = []
            t_first = None
            deadline = None if self.policy != "interval" or self.interval_ms <= 0 else time.monotonic() + self.interval_ms / 1e3
            while len(complete_order) < k:                                 # reference :148 "Waitany" loop
                try:
                    item = self._rx_queue.get(timeout=0.002 if deadline is not None else 1.0)
                except queue.Empty:
                    if deadline is not None and time.monotonic() > deadline and len(complete_order) >= 1:
                        break
                    continue
                if deadline is not None and time.monotonic() > deadline and len(complete_order) >= 1:
                    self._rx_queue.put(item)    # belongs to whoever closes the next step (or is dropped there as stale)
                    break
                w, l, stamp = item[0], item[1], item[2]
                if l < 0:
                    continue
                if stamp != self.cur_step or self.grad_accumulator.received[w][l]:
                    continue            # stale: a killed worker's late message (C++ rule tag != cur_step) -> dropped
                if t_first is None:
                    t_first = time.monotonic()
                self._stash(w, l, item[3])
                self.grad_accumulator.received[w][l] = True
                self.grad_accumulator.gradient_aggregate_counter[l] += 1
                if all(self.grad_accumulator.received[w]):
                    complete_order.append(w)
                    self.timeline.mark(self.cur_step, "worker_complete", w)
            selected = sorted(complete_order[:k]) if deadline is None else sorted(complete_order)
            killed = [w for w in range(n) if w not in selected]
            self.send_kill_signal(killed)                                  # reference :172-187 (no wait on the critical path)
            gather_s = time.monotonic() - (t_first or t_step)
            # ---- aggregate + update (reference :212-223) ----
            self._apply_update(selected)
            self.meset_grad_buffer()
            self.step_log.append((self.cur_step, selected, killed, gather_s))
            step_ms = (time.monotonic() - t_step) * 1e3
            timer.add(step_ms)
            if self.cur_step % self.log_interval == 0 or self.cur_step == first:
                print("Master Step: {} selected: {} killed: {} Gradient Gather Time: {:.4f} Step Time: {:.4f}".format(
                    self.cur_step, selected, killed, gather_s, step_ms / 1e3), flush=True)
            self.logger.log(step=self.cur_step, selected=selected, killed=killed, gather_s=gather_s, step_ms=step_ms)
            if self.eval_freq and self.train_dir and self.cur_step % self.eval_freq == 0:
                self.save_checkpoint()
            self.cur_step += 1
            if i < self.max_steps:
                self.async_bcast_step()
        self._shutdown()
        self.summary = timer.summary()
        self.timeline.flush()
        self.logger.log(event="summary", **self.summary)
        self.logger.close()
        return self.summary

    # ---- b200 path ---------------------------------------------------------------------------------------
    def _train_b200(self, train_loader):
        """The device backend's loop is ``B200ParameterServer.fit`` (pinned double-buffered input staging, the whole step
        replayed from a CUDA graph after two eager steps); this role logs, checkpoints and keeps the step records."""
        ps = self.ps
        if train_loader is not None and hasattr(train_loader, "pin_memory"):
            train_loader.pin_memory()           # batches are then DMA'd from where they are
        timer = StepTimer(self.batch_size * ps.k)
        mark = {"t": time.perf_counter(), "step": ps.step - 1}
        first_step = ps.step

        def on_log(i, loss):
            sel, kill, nrecv, st = ps.last_result()
            now = time.perf_counter()
            ms = (now - mark["t"]) * 1e3 / max(1, i - mark["step"])        # mean step time since the previous log line
            mark["t"], mark["step"] = now, i
            if i != first_step or self.max_steps == first_step:      # the first step carries one-time set-up (lazy kernel loads)
                timer.add(ms)
            selected = [w for w in range(ps.n_workers) if sel >> w & 1]
            killed = [w for w in range(ps.n_workers) if kill >> w & 1]
            self.step_log.append((i, selected, killed, ms / 1e3))
            print("Master Step: {} selected: {} killed: {} Step Time: {:.6f}".format(i, selected, killed, ms / 1e3), flush=True)
            self.logger.log(step=i, selected=selected, killed=killed, step_ms=ms, loss=loss, device_timeline_us=ps.last_timing_us)
            if self.eval_freq and self.train_dir and i % self.eval_freq == 0:
                if not ps.is_worker and hasattr(ps.net, "bn_running_stats"):
                    # master-only rank: BatchNorm running statistics come from the first worker (distributed_worker.py)
                    import torch.distributed as dist
                    flat = ps.net.bn_running_stats()
                    if flat.numel():
                        dist.recv(flat, src=1)
                        ps.net.load_bn_running_stats(flat.to(ps.device))
                self.cur_step = i + getattr(self, "step_offset", 0)
                self.save_checkpoint()
                mark["t"] = time.perf_counter()

        def should_log(i):
            return i % self.log_interval == 0 or (self.eval_freq and i % self.eval_freq == 0)

        ps.fit(train_loader if ps.is_worker else None, self.max_steps, log_interval=self.log_interval, on_log=on_log,
               should_log=should_log, cuda_graph=self.cuda_graph)
        if self._pending_ckpt is not None:
            self._pending_ckpt.join()
        self.cur_step = ps.step
        self.summary = timer.summary()
        self.logger.log(event="summary", **self.summary)
        self.logger.close()
        return self.summary

    # ---- message-backend pieces (reference method names) -------------------------------------------------------
    def async_bcast_step(self):
        """Send the current step to every worker (and the evaluator): tag 10 (reference :235-241)."""
        reqs = [self.comm.send_step(self.cur_step, r) for r in self._worker_ranks]
        if self._evaluator_rank is not None:
            reqs.append(self.comm.send_step(self.cur_step, self._evaluator_rank))
        return reqs

    def async_bcast_layer_weights_bcast(self):
        """One broadcast per parameter tensor, in ``parameters()`` order (reference :259-272)."""
        for p in self.params:
            self.comm.bcast(self.comm.to_wire(p.data), root=0)

    def async_bcast_layer_weights_async(self):
        """One Isend per (worker, parameter tensor), tag 11+l (reference :243-257)."""
        reqs = []
        dests = list(self._worker_ranks) + ([self._evaluator_rank] if self._evaluator_rank is not None else [])
        for l, p in enumerate(self.params):
            wire = self.comm.to_wire(p.data)
            for r in dests:
                reqs.append(self.comm.isend(wire, r, TAG_WEIGHT_BASE + l))
        self.comm.wait_all(reqs)

    def async_fetch_gradient_start(self):
        """Arm the receive side for tag 88+l from every worker (reference :275-284). With gloo this is one receiver
        thread per worker, started once; messages of abandoned steps are recognised by their step stamp."""
        if not getattr(self, "_rx_started", False):
            self._rx_queue = queue.Queue()
            self.comm.start_grad_receivers(self._worker_ranks, self.grad_accumulator.gradient_aggregator, self._rx_queue)
            self._rx_started = True
        return self._rx_queue

    def _stash(self, w, l, buf):
        # keep the arrived gradient until the contributing set is known (complete sets only, §2.11-3)
        if not hasattr(self, "_stashed"):
            self._stashed = {}
        self._stashed[(w, l)] = buf

    def aggregate_gradient(self, gradient, layer_idx):
        """Add one gradient into the per-tensor sum (reference :286-288)."""
        if self._sum[layer_idx] is None:
            self._sum[layer_idx] = gradient.to(torch.float32).clone()
        else:
            self._sum[layer_idx] += gradient.to(torch.float32)

    def _apply_update(self, selected):
        n, k = self._num_grad_to_collect, max(1, len(selected))
        scale = 1.0 / n if self.avg_mode == "n" else 1.0 / k
        self._sum = [None] * len(self.params)
        for w in sorted(selected):                                     # rank order: deterministic sum
            for l in self.grad_accumulator.model_index_range:
                self.aggregate_gradient(self._stashed[(w, l)], l)
        tmp_module = []
        for l, p in enumerate(self.params):
            g = self._sum[l] * scale
            m = self.momentum_buffers[l]
            m.mul_(self.momentum).add_(g.view_as(m))
            tmp_module.append(update_params_dist_version(p.data.clone(), m, self.lr))
        self.model_update(tmp_module)
        self._stashed = {}

    def model_update(self, tmp_module):
        """Install the updated tensors (reference :290-303). Only parameters move: BN running statistics of the
        master keep their initial values, as in the reference (§2.11-7)."""
        with torch.no_grad():
            for p, new in zip(self.params, tmp_module):
                assert p.shape == new.shape
                p.copy_(new)

    def meset_grad_buffer(self):
        self.grad_accumulator.meset_everything()

    def send_kill_signal(self, should_kill_list):
        """tag 77 to every straggler, carrying the step being closed (reference :309-314)."""
        return [self.comm.send_kill(self._worker_ranks[w], self.cur_step) for w in should_kill_list]

    def save_checkpoint(self):
        sd = self.network.state_dict()
        if self._pending_ckpt is not None:
            self._pending_ckpt.join()
        mom = self.ps.momentum if self.ps is not None else torch.cat([m.reshape(-1) for m in self.momentum_buffers])
        path, th = ckpt.save_checkpoint(self.train_dir, self.cur_step, sd, mom, background=True)
        self._pending_ckpt = th
        return path

    def _shutdown(self):
        """Tell everyone to stop and drain the armed receives (workers answer with one tombstone per tensor)."""
        comm = self.comm
        reqs = [comm.send_step(TERMINATE_STEP, r) for r in self._worker_ranks]
        reqs += [comm.send_kill(r, TERMINATE_STEP) for r in self._worker_ranks]
        if self._evaluator_rank is not None:
            reqs.append(comm.send_step(TERMINATE_STEP, self._evaluator_rank))
        comm.join_grad_receivers()
        comm.wait_all(reqs)
        if self._pending_ckpt is not None:
            self._pending_ckpt.join()
