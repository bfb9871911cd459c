"""Device-resident synchronous-SGD parameter server (the B200 backend).

Role model (reference: rank 0 = master holding the weights, ranks >= 1 = workers, pytorch_code/distributed_nn.py:91-102):

* ``colocate=False``  rank 0 is master-only, ranks 1..G-1 are the n = G-1 workers (reference semantics).
* ``colocate=True``   GPU 0 additionally hosts a worker, so n = G (and G = 1 is a complete system: one
                      master + one worker sharing a stream). This is what the throughput bench uses.

Per step s (1-based, as ``STEP_START_`` in the reference):

  worker  wait(ready >= s) -> prepare derived weight layouts -> [injected delay] -> forward -> loss ->
          backward; each finished gradient bucket is pushed (bf16 or fp32) into the master's slot for this
          worker by ``pdnn_ps_push``; the last bucket's push publishes ``arrival[w] = s``.
  master  ``pdnn_ps_master_step``: one kernel polls the arrival flags, takes the first k complete sets, stores
          ``closed = s`` into every late worker's memory (their remaining kernels become no-ops and their push
          is dropped), reduces the selected slots, applies 1/k + momentum SGD on the fp32 master weights,
          casts and stores the new weights into every worker's weight buffer and finally ``ready = s+1``.

Nothing in the step touches the host: each rank simply enqueues its kernels; streams block on device flags.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Callable, Optional

import torch

from ..engine.network import Network, build_network
from ..ops import _lib
from ..ops import functional as F
from .peer import make_arena
from .protocol import AVG_MODES, STEP_START_

U32 = 4
FLAG_WORDS = 256          # per-rank flag block (uint32 words)
W_READY, W_CLOSED, W_STEP = 0, 1, 2         # worker-side words
M_ARRIVAL, M_CONTROL, M_RESULT, M_STEPW = 32, 64, 128, 160  # master-side words (arrival[n], control[64], result[4], step)


class B200ParameterServer:
    """One instance per rank. Builds the engine network, the peer arena and drives the step loop."""

    def __init__(self, network: str, batch_size: int, rank: int = 0, world_size: int = 1, *, lr: float = 0.01,
                 momentum: float = 0.0, num_aggregate: Optional[int] = None, avg_mode: str = "k",
                 grad_dtype: str = "bf16", colocate: bool = True, n_buckets: int = 8, seed: int = 0,
                 deadline_ms: float = 0.0, delay_fn: Optional[Callable[[int, int], int]] = None, group=None,
                 device=None, init_state_dict=None, init_momentum=None, poison_slots: Optional[bool] = None):
        assert torch.cuda.is_available(), "the b200 backend needs a GPU (use --backend gloo on CPU)"
        # race-detection mode: the master kernel overwrites every slot it has summed with NaN before releasing
        # the workers; a gradient that is ever consumed twice (stale step) shows up as NaN weights. Env: PDNN_PS_POISON=1.
        self.poison_slots = (os.environ.get("PDNN_PS_POISON") == "1") if poison_slots is None else bool(poison_slots)
        self.rank, self.world = rank, world_size
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.colocate = colocate or world_size == 1
        self.n_workers = world_size if self.colocate else world_size - 1
        assert 1 <= self.n_workers <= _lib.MAX_WORKERS
        self.is_master = rank == 0
        self.worker_id = rank if self.colocate else rank - 1          # -1 on a master-only rank
        self.is_worker = self.worker_id >= 0
        self.k = self.n_workers if not num_aggregate else min(int(num_aggregate), self.n_workers)
        self.lr, self.mu = float(lr), float(momentum)
        self.avg_mode = AVG_MODES[avg_mode]
        self.grad_bf16 = grad_dtype == "bf16"
        self.deadline_ns = int(deadline_ms * 1e6)
        self.delay_fn = delay_fn
        self.step = STEP_START_
        self.batch_size = batch_size
        self._graph_mode = False
        self._graph = None
        self._check_agreement(network, group)

        # ---- network (every rank builds it: the master needs the layout, workers need the layers)
        self.net: Network = self._build(network, batch_size, n_buckets, group)
        net = self.net
        if self.is_master:
            net.init_parameters(seed)          # fills w32 and (through sync_compute_weights) the local w
            if init_state_dict is not None:    # resume: the checkpoint replaces the random init before the first broadcast
                net.load_state_dict(init_state_dict)
            self.momentum = torch.zeros(net.total, device=self.device, dtype=torch.float32)
            if init_momentum is not None:      # resume: the optimizer state continues as well (TF Supervisor restores everything)
                assert init_momentum.numel() == net.total, "momentum sidecar does not match this network"
                self.momentum.copy_(init_momentum.to(self.device, torch.float32).view(-1))
        self._install_hooks()
        self._initial_broadcast(group)

    def _check_agreement(self, network, group):
        """Every rank derives its data path from (k, n, deadline, wire dtype): a worker that believes k = n writes its
        gradients into its LOCAL slot for the in-switch sum while a master with k < n reads pushed slots nobody fills.
        Arrival flags would still fire, so nothing hangs and training silently stops learning -- hence a hard check."""
        if self.world == 1:
            return
        import torch.distributed as dist
        mine = (str(network), self.n_workers, self.k, self.deadline_ns, self.grad_bf16, bool(self.colocate), self.avg_mode)
        everyone = [None] * self.world
        dist.all_gather_object(everyone, mine, group=group)
        if any(tuple(o) != mine for o in everyone):
            raise RuntimeError("B200ParameterServer: ranks disagree on (network, n_workers, k, deadline_ns, grad_bf16, colocate, avg_mode): "
                               + "; ".join(f"rank {r}: {o}" for r, o in enumerate(everyone)))

    # ---- construction ---------------------------------------------------------------------------
    def _build(self, network, batch, n_buckets, group) -> Network:
        # a meta-device pass gives the flat layout (size, dtype) without allocating anything
        proto = build_network(network, 128, device="meta", n_buckets=n_buckets)
        total, wdtype = proto.total, proto.dtype
        del proto
        self.wbytes = total * (2 if wdtype == torch.bfloat16 else 4)
        self.slot_elem = 2 if self.grad_bf16 else 4
        slot_bytes = total * self.slot_elem
        # symmetric size keeps the carve offsets identical on every rank (only rank 0's master-side slots are used)
        arena_bytes = 4096 + FLAG_WORDS * U32 + self.wbytes + 512 + (self.n_workers + 1) * (slot_bytes + 256)
        self.arena = make_arena(arena_bytes, self.rank, self.world, self.device, group,
                                prefer_symmetric=os.environ.get("PDNN_NO_SYMM") != "1")
        self.off_flags = self.arena.carve(FLAG_WORDS * U32)
        self.off_w = self.arena.carve(self.wbytes)
        self.off_gslot = self.arena.carve(slot_bytes)            # this rank's LOCAL gradient slot (NVLS gather mode)
        self.off_slots = [self.arena.carve(slot_bytes) for _ in range(self.n_workers)]
        # NVLS (in-switch) paths: weights go out with one multimem.st; when every worker is aggregated the gradients
        # come in with one multimem.ld_reduce over the workers' local slots instead of n pushed copies
        self.nvls = bool(getattr(self.arena, "mc_ptr", 0)) and os.environ.get("PDNN_NO_NVLS") != "1"
        self.nvls_gather = self.nvls and self.k == self.n_workers and self.deadline_ns == 0
        self.flags = self.arena.tensor(self.off_flags, FLAG_WORDS * U32, torch.int32)
        w_ext = self.arena.tensor(self.off_w, self.wbytes, wdtype)
        real_batch = batch if self.is_worker else self._min_batch(network)
        return build_network(network, real_batch, device=self.device, n_buckets=n_buckets, external_w=w_ext)

    @staticmethod
    def _min_batch(network):
        return 32

    def _flag_ptr(self, rank, word):
        return self.arena.ptr(rank, self.off_flags + word * U32)

    def _worker_rank(self, w):
        return w if self.colocate else w + 1

    def _install_hooks(self):
        net = self.net
        self.closed_word = self.flags[W_CLOSED: W_CLOSED + 1]
        self.step_word = self.flags[W_STEP: W_STEP + 1]
        if self.is_worker:
            w = self.worker_id
            if self.nvls_gather:
                self._slot_ptr = self.arena.ptr(self.rank, self.off_gslot)     # local slot, summed in the switch by the master
            else:
                self._slot_ptr = self.arena.ptr(0, self.off_slots[w])          # my slot lives on the master (rank 0)
            self._arrival_ptr = self._flag_ptr(0, M_ARRIVAL + w)
            last_bucket = net.buckets[-1].index

            def on_bucket(b):
                flag = self._arrival_ptr if b.index == last_bucket else None
                if self._graph_mode:      # step tag read from the device word: the launch is replayable
                    _lib.call("pdnn_ps_push_dev", net.g.data_ptr() + b.start * 4, self._slot_ptr + b.start * self.slot_elem,
                              b.end - b.start, int(self.grad_bf16), flag, self.step_word.data_ptr(),
                              self.closed_word.data_ptr(), self.step_word.data_ptr(), 0, _lib.stream_ptr())
                else:
                    _lib.call("pdnn_ps_push", net.g.data_ptr() + b.start * 4, self._slot_ptr + b.start * self.slot_elem,
                              b.end - b.start, int(self.grad_bf16), flag, self.step,
                              self.closed_word.data_ptr(), self.step_word.data_ptr(), 0, _lib.stream_ptr())

            net.on_bucket = on_bucket
        if self.is_master:
            m = _lib.PSMaster()
            for w in range(self.n_workers):
                r = self._worker_rank(w)
                m.slots[w] = self.arena.ptr(r, self.off_gslot) if self.nvls_gather else self.arena.ptr(0, self.off_slots[w])
                m.worker_weights[w] = self.arena.ptr(r, self.off_w)
                m.worker_ready[w] = self._flag_ptr(r, W_READY)
                m.worker_closed[w] = self._flag_ptr(r, W_CLOSED)
            m.arrival_flags = self._flag_ptr(0, M_ARRIVAL)
            m.multicast_weights = self.arena.mc(self.off_w) if self.nvls else None
            m.multicast_slots = self.arena.mc(self.off_gslot) if self.nvls_gather else None
            m.master_w = net.w32.data_ptr()
            m.momentum = self.momentum.data_ptr()
            m.control = self._flag_ptr(0, M_CONTROL)
            m.n = net.total
            m.n_workers = self.n_workers
            m.k = self.k
            m.slot_is_bf16 = int(self.grad_bf16)
            m.weights_are_bf16 = int(net.w.dtype == torch.bfloat16)
            m.avg_mode = self.avg_mode
            m.lr, m.mu = self.lr, self.mu
            m.deadline_ns = self.deadline_ns
            m.max_ctas = 0
            m.poison_slots = int(self.poison_slots)
            self._margs = m
            self.result = self.flags[M_RESULT: M_RESULT + 8]
            self.history = []

    def _initial_broadcast(self, group):
        """Step-1 weights: the master stores its initial weights into every worker buffer and releases them
        (reference: first async_bcast_step + weight Bcast, sync_replicas_master_nn.py:121-140)."""
        net = self.net
        if self.is_master:
            src = net.w32
            for w in range(self.n_workers):
                r = self._worker_rank(w)
                dst = self.arena.ptr(r, self.off_w)
                if r == 0:
                    continue                      # init_parameters already filled the local compute copy
                if net.w.dtype == torch.bfloat16:
                    _lib.call("pdnn_cast_f32_to_bf16", src, dst, net.total, _lib.stream_ptr())
                else:
                    _lib.call("pdnn_p2p_copy", src, dst, net.total * 4, 0, _lib.stream_ptr())
            for w in range(self.n_workers):
                _lib.call("pdnn_set_word", self._flag_ptr(self._worker_rank(w), W_READY), STEP_START_, _lib.stream_ptr())
        torch.cuda.synchronize()
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier(group=group)

    # ---- step loop ------------------------------------------------------------------------------
    def worker_step(self, x_nchw_f32, labels):
        """Enqueue one worker step (no host synchronisation). Returns the device loss tensor."""
        net = self.net
        st = _lib.stream_ptr()
        if self._graph_mode:
            _lib.call("pdnn_ps_wait_next", self.flags.data_ptr() + W_READY * U32, self.step_word.data_ptr(), 0, st)
        else:
            _lib.call("pdnn_ps_wait_flag", self.flags.data_ptr() + W_READY * U32, self.step, self.step_word.data_ptr(), 0, st)
        F.set_abort_guard(self.closed_word, self.step_word)
        try:
            net.prepare()
            if self.delay_fn is not None:
                ns = int(self.delay_fn(self.worker_id, self.step))
                if ns > 0:
                    _lib.call("pdnn_delay", ns, self.closed_word.data_ptr(), self.step_word.data_ptr(), st)
            loss = net.train_step(x_nchw_f32, labels)
        finally:
            F.set_abort_guard(None, None)
        return loss

    def master_step(self):
        """Enqueue the master's aggregate/update/broadcast kernel for the current step."""
        m = self._margs
        m.step = self.step
        m.step_ptr = self._flag_ptr(0, M_STEPW) if self._graph_mode else None
        _lib.call("pdnn_ps_master_step", C.byref(m), self.result.data_ptr(), _lib.stream_ptr())

    def enable_cuda_graph(self, x_example=None, y_example=None):
        """Capture one complete step of this rank (wait -> layouts -> forward -> loss -> backward with bucket pushes ->
        [master kernel]) into a CUDA graph; later ``train_step`` calls copy the batch into static buffers and replay.
        The step number moves to device words (`pdnn_ps_wait_next`, `pdnn_ps_push_dev`, master ``step_ptr``) so nothing in
        the captured launches depends on it. Call on EVERY rank at the same step (one real step runs as warm-up).
        Host-computed per-step delays (``delay_fn``) cannot be replayed and are rejected."""
        if self.delay_fn is not None:
            raise ValueError("CUDA-graph replay is incompatible with host-scheduled delay injection")
        torch.cuda.synchronize()
        dev = self.device
        if self.is_worker:
            self.flags[W_STEP] = self.step - 1         # last completed step; wait_next advances it
        if self.is_master:
            self.flags[M_STEPW] = self.step
        self._graph_mode = True
        if self.is_worker:
            self._gx = x_example.clone()
            self._gy = y_example.clone()
        else:
            self._gx = self._gy = None
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):                  # one eager step in graph mode (also the warm-up torch asks for)
            self._step_body(self._gx, self._gy)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize()
        self.step += 1
        self._graph = torch.cuda.CUDAGraph()
        n0 = _lib.launch_count()
        with torch.cuda.graph(self._graph):
            self._graph_loss = self._step_body(self._gx, self._gy)
        self.graph_kernel_launches = _lib.launch_count() - n0      # library kernels inside one replay
        return self

    def local_slot(self, w: int) -> torch.Tensor:
        """Rank 0, unicast-slot mode: the master-side gradient slot of worker ``w`` as a tensor (tests, debugging)."""
        assert self.is_master and not self.nvls_gather, "slots live on the workers in NVLS gather mode"
        dt = torch.bfloat16 if self.grad_bf16 else torch.float32
        return self.arena.tensor(self.off_slots[w], self.net.total * self.slot_elem, dt)

    def _step_body(self, x, y):
        loss = None
        if self.is_worker:
            loss = self.worker_step(x, y)
        if self.is_master:
            self.master_step()
        return loss

    def train_step(self, x_nchw_f32=None, labels=None):
        """One synchronous-SGD step of this rank's role(s); advances the step counter."""
        if self._graph is not None:
            if self.is_worker:
                self._gx.copy_(x_nchw_f32, non_blocking=True)
                self._gy.copy_(labels, non_blocking=True)
            self._graph.replay()
            _lib.add_launch_count(self.graph_kernel_launches)
            self.step += 1
            return self._graph_loss
        loss = None
        if self.is_worker:
            loss = self.worker_step(x_nchw_f32, labels)
        if self.is_master:
            self.master_step()
        self.step += 1
        return loss

    def last_result(self):
        """(selected_mask, kill_mask, n_received, step) of the most recent master step (synchronises)."""
        r = [int(v) & 0xFFFFFFFF for v in self.result.cpu().tolist()]
        if r[3] == 0xDEAD:
            raise RuntimeError("parameter-server master timed out: fewer than k gradient sets arrived (a worker died or hung); "
                               f"arrived mask={r[0]:b}")
        self.last_timing_us = {"wait_for_kth_arrival": r[4], "reduce_update_broadcast": r[5]}
        return tuple(r[:4])

    def master_state_dict(self):
        return self.net.state_dict()

    def close(self):
        torch.cuda.synchronize()
        self.arena.close()
