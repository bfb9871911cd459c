"""Device-resident synchronous-SGD parameter server (the B200 backend).

Role model (reference: rank 0 = master holding the weights, ranks >= 1 = workers, pytorch_code/distributed_nn.py:91-102):

* ``colocate=False``  rank 0 is master-only, ranks 1.. are the workers (reference semantics).
* ``colocate=True``   GPU 0 additionally hosts a worker, so n = G (and G = 1 is a complete system: one
                      master + one worker on one GPU). This is what the throughput bench uses.
* ``n_listeners``     the last ranks take no part in training but receive every step's weights (the evaluator
                      role of MPI_code/src/distributed/evaluator_nn.h).

Per step s (1-based, as ``STEP_START_`` in the reference):

  worker  wait(ready >= s) -> copy this step's landing buffer (parity s & 1) into the private compute weights ->
          derived weight layouts -> [injected delay] -> forward -> loss -> backward; every finished gradient bucket
          (deepest layers first) is pushed (bf16 or fp32) into the master's slot for this worker by ``pdnn_ps_push``
          and published with its own step-tagged arrival flag ``arrival[w][b] = s``.
  master  ``pdnn_ps_master_step`` on the master's OWN stream: for every bucket in arrival order one warp waits for
          the first k flags of that bucket, then a grid reduces the selected slots (or lets the NVSwitch sum them),
          applies 1/k + momentum SGD on the fp32 master weights, casts and stores the new weights into every
          receiver's landing buffer of parity (s+1) & 1 -- while the workers are still in the backward pass of the
          shallower layers. The last bucket decides the kill set (workers not among its first k: ``closed = s`` in
          their memory, their remaining kernels become no-ops and their pushes are dropped) and publishes
          ``ready = s+1``.

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
FLAG_WORDS = 1024         # per-rank flag block (uint32 words)
W_READY, W_CLOSED, W_STEP = 0, 1, 2         # worker-side words
M_CONTROL, M_RESULT, M_STEPW, M_ARRIVAL = 64, 320, 400, 448  # master-side: control[256], result[80], step word, arrival[n][stride]
ARRIVAL_STRIDE = _lib.MAX_BUCKETS


class B200ParameterServer:
    """One instance per rank. Builds the engine network, the peer arena and drives the step loop."""

    def __init__(self, network: str, batch_size: int, rank: int = 0, world_size: int = 1, *, lr: float = 0.01,
                 momentum: float = 0.0, num_aggregate: Optional[int] = None, avg_mode: str = "k",
                 grad_dtype: str = "bf16", colocate: bool = True, n_buckets: int = 8, seed: int = 0,
                 deadline_ms: float = 0.0, delay_fn: Optional[Callable[[int, int], int]] = None, group=None,
                 device=None, init_state_dict=None, init_momentum=None, poison_slots: Optional[bool] = None,
                 n_listeners: int = 0, device_delay=None):
        assert torch.cuda.is_available(), "the b200 backend needs a GPU (use --backend gloo on CPU)"
        # race-detection mode: the master kernel overwrites every slot it has summed with NaN before releasing
        # the workers; a gradient that is ever consumed twice (stale step) shows up as NaN weights. Env: PDNN_PS_POISON=1.
        self.poison_slots = (os.environ.get("PDNN_PS_POISON") == "1") if poison_slots is None else bool(poison_slots)
        self.rank, self.world = rank, world_size
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.colocate = colocate or world_size == 1
        self.n_listeners = int(n_listeners)
        self.n_workers = world_size - self.n_listeners - (0 if self.colocate else 1)
        assert 1 <= self.n_workers <= _lib.MAX_WORKERS, f"{self.n_workers} workers (world {world_size}, listeners {n_listeners})"
        self.is_master = rank == 0
        self.is_listener = rank >= world_size - self.n_listeners
        self.worker_id = -1 if self.is_listener else (rank if self.colocate else rank - 1)   # -1: no worker on this rank
        self.is_worker = self.worker_id >= 0
        self.k = self.n_workers if not num_aggregate else min(int(num_aggregate), self.n_workers)
        self.lr, self.mu = float(lr), float(momentum)
        self.avg_mode = AVG_MODES[avg_mode]
        self.grad_bf16 = grad_dtype == "bf16"
        self.deadline_ns = int(deadline_ms * 1e6)
        self.delay_fn = delay_fn
        self.device_delay = device_delay          # parallel.delays.DeviceDelay: sampled on the GPU, replayable in a CUDA graph
        self.step = STEP_START_
        self.batch_size = batch_size
        self._graph_mode = False
        self._graph = self._mgraph = None
        self._graph_ready = False
        self.last_timing_us = {}
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
            self.master_stream = torch.cuda.Stream(device=self.device, priority=int(os.environ.get("PDNN_MASTER_PRIO", "-1")))
        # How the master's kernels meet the colocated worker's on GPU 0:
        #   gated   (all-n, no deadline, rank 0 trains too): bucket b's {wait, reduce} pair goes to the master's stream behind
        #           an event recorded after this rank's own push of b -- without that push the bucket cannot be complete
        #           anyway, and the polling warp is then resident for the skew between workers only;
        #   free    (k < n, interval policy, master-only rank 0): the whole step's pairs run free on the master's stream, so
        #           a slow worker 0 cannot hold the master back;
        #   inline  (A/B switch): the pairs sit in the worker's own stream.
        default_mode = "gated" if (self.is_master and self.is_worker and self.k == self.n_workers and self.deadline_ns == 0) else "free"
        self.master_mode = os.environ.get("PDNN_MASTER_MODE", default_mode) if (self.is_master and self.is_worker) else "free"
        self._bucket_events = {}
        self._install_hooks()
        self._initial_broadcast(group)

    def _check_agreement(self, network, group):
        """Every rank derives its data path from (k, n, deadline, wire dtype): a worker that believes k = n writes its
        gradients into its LOCAL slot for the in-switch sum while a master with k < n reads pushed slots nobody fills.
        Arrival flags would still fire, so nothing hangs and training silently stops learning -- hence a hard check."""
        if self.world == 1:
            return
        import torch.distributed as dist
        mine = (str(network), self.n_workers, self.k, self.deadline_ns, self.grad_bf16, bool(self.colocate), self.avg_mode,
                self.n_listeners)
        everyone = [None] * self.world
        dist.all_gather_object(everyone, mine, group=group)
        if any(tuple(o) != mine for o in everyone):
            raise RuntimeError("B200ParameterServer: ranks disagree on (network, n_workers, k, deadline_ns, grad_bf16, colocate, "
                               "avg_mode, n_listeners): " + "; ".join(f"rank {r}: {o}" for r, o in enumerate(everyone)))

    # ---- construction ---------------------------------------------------------------------------
    def _build(self, network, batch, n_buckets, group) -> Network:
        # a meta-device pass gives the flat layout (size, dtype) without allocating anything
        n_buckets = max(1, min(int(n_buckets), _lib.MAX_BUCKETS))
        proto = build_network(network, 128, device="meta", n_buckets=n_buckets)
        total, wdtype = proto.total, proto.dtype
        assert len(proto.buckets) <= ARRIVAL_STRIDE
        del proto
        self.wbytes = total * (2 if wdtype == torch.bfloat16 else 4)
        self.slot_elem = 2 if self.grad_bf16 else 4
        slot_bytes = total * self.slot_elem
        # symmetric size keeps the carve offsets identical on every rank (only rank 0's master-side slots are used)
        arena_bytes = 8192 + FLAG_WORDS * U32 + 2 * (self.wbytes + 256) + (self.n_workers + 1) * (slot_bytes + 256)
        self.arena = make_arena(arena_bytes, self.rank, self.world, self.device, group,
                                prefer_symmetric=os.environ.get("PDNN_NO_SYMM") != "1")
        self.off_flags = self.arena.carve(FLAG_WORDS * U32)
        self.off_land = [self.arena.carve(self.wbytes), self.arena.carve(self.wbytes)]   # weight landing buffers by step parity
        self.off_gslot = self.arena.carve(slot_bytes)            # this rank's LOCAL gradient slot (NVLS gather mode)
        self.off_slots = [self.arena.carve(slot_bytes) for _ in range(self.n_workers)]
        # NVLS (in-switch) paths: weights go out with one multimem.st; when every worker is aggregated the gradients
        # come in with one multimem.ld_reduce over the workers' local slots instead of n pushed copies
        self.nvls = bool(getattr(self.arena, "mc_ptr", 0)) and os.environ.get("PDNN_NO_NVLS") != "1"
        self.nvls_gather = self.nvls and self.k == self.n_workers and self.deadline_ns == 0
        self.flags = self.arena.tensor(self.off_flags, FLAG_WORDS * U32, torch.int32)
        self.land = [self.arena.tensor(o, self.wbytes, wdtype) for o in self.off_land]
        self.gslot = self.arena.tensor(self.off_gslot, slot_bytes, torch.bfloat16 if self.grad_bf16 else torch.float32)
        real_batch = batch if (self.is_worker or self.is_listener) else self._min_batch(network)
        # zero-copy push (fp32 wire + in-switch sum): the gradient buffer the weight-gradient kernels accumulate into IS the
        # slot the master's multimem.ld_reduce reads, so a finished bucket only needs its flag published -- no push kernel
        self.zero_copy = self.nvls_gather and not self.grad_bf16 and os.environ.get("PDNN_PS_ZERO_COPY", "1") != "0"
        return build_network(network, real_batch, device=self.device, n_buckets=n_buckets,
                             external_g=self.gslot if self.zero_copy else None)

    @staticmethod
    def _min_batch(network):
        return 32

    def _flag_ptr(self, rank, word):
        return self.arena.ptr(rank, self.off_flags + word * U32)

    def _worker_rank(self, w):
        return w if self.colocate else w + 1

    def _receiver_ranks(self):
        """Ranks that get every step's weights and the ready flag: the workers, then the listeners."""
        return [self._worker_rank(w) for w in range(self.n_workers)] + list(range(self.world - self.n_listeners, self.world))

    def _install_hooks(self):
        net = self.net
        self.closed_word = self.flags[W_CLOSED: W_CLOSED + 1]
        self.step_word = self.flags[W_STEP: W_STEP + 1]
        self.ready_word = self.flags[W_READY: W_READY + 1]
        if self.is_worker:
            w = self.worker_id
            if self.nvls_gather:
                self._slot_ptr = self.arena.ptr(self.rank, self.off_gslot)     # local slot, summed in the switch by the master
            else:
                self._slot_ptr = self.arena.ptr(0, self.off_slots[w])          # my slot lives on the master (rank 0)
            self._arrival_ptr = self._flag_ptr(0, M_ARRIVAL + w * ARRIVAL_STRIDE)

            def on_bucket(b):
                # every bucket carries its own step-tagged arrival flag: the master starts on bucket b while backward continues
                flag = self._arrival_ptr + b.index * U32
                if self.zero_copy:
                    _lib.call("pdnn_ps_publish", flag, self.step, self.step_word.data_ptr() if self._graph_mode else None,
                              self.closed_word.data_ptr(), self.step_word.data_ptr(), _lib.stream_ptr())
                elif self._graph_mode:      # step tag read from the device word: the launch is replayable
                    _lib.call("pdnn_ps_push_dev", net.g.data_ptr() + b.start * 4, self._slot_ptr + b.start * self.slot_elem,
                              b.end - b.start, int(self.grad_bf16), flag, self.step_word.data_ptr(),
                              self.closed_word.data_ptr(), self.step_word.data_ptr(), 0, _lib.stream_ptr())
                else:
                    _lib.call("pdnn_ps_push", net.g.data_ptr() + b.start * 4, self._slot_ptr + b.start * self.slot_elem,
                              b.end - b.start, int(self.grad_bf16), flag, self.step,
                              self.closed_word.data_ptr(), self.step_word.data_ptr(), 0, _lib.stream_ptr())

            if self.is_master and self.master_mode in ("gated", "inline"):
                push = on_bucket

                def on_bucket(b, _push=push):     # noqa: F811 -- the master's kernels for bucket b follow this rank's own push of b
                    _push(b)
                    self._master_bucket(b.index)

            net.on_bucket = on_bucket
        if self.is_master:
            m = _lib.PSMaster()
            for w in range(self.n_workers):
                r = self._worker_rank(w)
                m.slots[w] = self.arena.ptr(r, self.off_gslot) if self.nvls_gather else self.arena.ptr(0, self.off_slots[w])
                m.worker_closed[w] = self._flag_ptr(r, W_CLOSED)
            recv = self._receiver_ranks()
            assert len(recv) <= _lib.MAX_RECEIVERS
            for i, r in enumerate(recv):
                for par in range(2):
                    m.recv_weights[par][i] = self.arena.ptr(r, self.off_land[par])
                m.recv_ready[i] = self._flag_ptr(r, W_READY)
            m.n_receivers = len(recv)
            m.arrival_flags = self._flag_ptr(0, M_ARRIVAL)
            m.arrival_stride = ARRIVAL_STRIDE
            for par in range(2):
                m.multicast_weights[par] = self.arena.mc(self.off_land[par]) if self.nvls else None
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
            m.hard_timeout_ns = int(float(os.environ.get("PDNN_PS_TIMEOUT_S", "0")) * 1e9)
            # a free-running master next to a colocated worker polls every 2 us: measured on ResNet-18 b1024, a 400 ns poll
            # loop resident for the whole backward pass stretches the worker's step from 6.6 to 8.4 ms, 5 us to 7.4 ms
            m.poll_sleep_ns = int(os.environ.get("PDNN_PS_POLL_NS", "2000" if self.is_worker else "0"))
            m.n_buckets = len(net.buckets)
            for b in net.buckets:             # bucket.index is the arrival order (deepest layers first)
                m.bucket_lo[b.index], m.bucket_hi[b.index] = b.start, b.end
            self._margs = m
            self.result = self.flags[M_RESULT: M_RESULT + 80]
            self.history = []

    def _initial_broadcast(self, group):
        """Step-1 weights: the master stores its initial weights into every receiver's landing buffer of parity 1 and
        releases them (reference: first async_bcast_step + weight Bcast, sync_replicas_master_nn.py:121-140)."""
        net = self.net
        if self.is_master:
            src = net.w32
            par = STEP_START_ & 1
            for r in self._receiver_ranks():
                dst = self.arena.ptr(r, self.off_land[par])
                if net.w.dtype == torch.bfloat16:
                    _lib.call("pdnn_cast_f32_to_bf16", src, dst, net.total, _lib.stream_ptr())
                else:
                    _lib.call("pdnn_p2p_copy", src, dst, net.total * 4, 0, _lib.stream_ptr())
            for r in self._receiver_ranks():
                _lib.call("pdnn_set_word", self._flag_ptr(r, W_READY), STEP_START_, _lib.stream_ptr())
        torch.cuda.synchronize()
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier(group=group)

    # ---- step loop ------------------------------------------------------------------------------
    def fetch_weights(self):
        """After the ready flag: this step's landing buffer -> the private compute copy (+ derived layouts)."""
        net = self.net
        _lib.call("pdnn_ps_fetch_weights", self.land[0].data_ptr(), self.land[1].data_ptr(), net.w.data_ptr(), self.wbytes,
                  self.step_word.data_ptr() if self._graph_mode else None, self.step, _lib.stream_ptr())
        net.prepare()

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
            self.fetch_weights()
            if self.delay_fn is not None:
                ns = int(self.delay_fn(self.worker_id, self.step))
                if ns > 0:
                    _lib.call("pdnn_delay", ns, self.closed_word.data_ptr(), self.step_word.data_ptr(), st)
            if self.device_delay is not None:
                dd, wid = self.device_delay, self.worker_id
                net.layer_hook = lambda idx: dd.launch(wid, idx, self.step_word, self.closed_word, st)
            loss = net.train_step(x_nchw_f32, labels)
        finally:
            net.layer_hook = None
            F.set_abort_guard(None, None)
        return loss

    def master_step(self):
        """Enqueue the master's per-bucket wait / reduce / update / broadcast kernels for the current step on the master's
        OWN stream. No stream dependency ties them to the worker's stream: arrival flags order the master behind the
        pushes, the ready flag orders the next worker step behind the master (both on the device)."""
        m = self._margs
        m.step = self.step
        m.step_ptr = self._flag_ptr(0, M_STEPW) if self._graph_mode else None
        ms = self.master_stream
        _lib.call("pdnn_ps_master_step", C.byref(m), self.result.data_ptr(), ms.cuda_stream)
        if self.poison_slots and self.nvls_gather and not self.is_worker:
            with torch.cuda.stream(ms):       # a master-only rank never rewrites its own (all-zero) slot: undo the multicast poison
                self.gslot.zero_()

    def _master_bucket(self, b):
        m = self._margs
        m.step = self.step
        m.step_ptr = self._flag_ptr(0, M_STEPW) if self._graph_mode else None
        cur = torch.cuda.current_stream(self.device)
        if self.master_mode == "inline":
            stream = cur.cuda_stream
        else:
            ev = self._bucket_events.get(b)
            if ev is None:
                ev = self._bucket_events[b] = torch.cuda.Event()
            ev.record(cur)
            self.master_stream.wait_event(ev)
            stream = self.master_stream.cuda_stream
        _lib.call("pdnn_ps_master_bucket", C.byref(m), self.result.data_ptr(), b, stream)

    def join(self):
        """Make the current stream wait for everything the master's stream has been given so far (timing brackets)."""
        if self.is_master:
            torch.cuda.current_stream(self.device).wait_stream(self.master_stream)

    def enable_cuda_graph(self, x_example=None, y_example=None):
        """Capture one complete step of this rank into CUDA graphs: the worker's step (wait -> fetch weights -> forward ->
        loss -> backward with bucket pushes) as one graph replayed on the caller's stream. On rank 0 the master's kernels
        are part of that graph in *gated* mode (each bucket's {wait, reduce} pair forks to the master's stream behind
        this rank's own push of the bucket and joins at the end of the step) and a second graph replayed on the master's
        stream in *free* mode (k < n, interval policy, master-only rank 0: the master must not wait for worker 0).
        Later ``train_step`` calls copy the batch into static buffers and replay. The
        step number moves to device words (`pdnn_ps_wait_next`, `pdnn_ps_push_dev`, master ``step_ptr``) so nothing in the
        captured launches depends on it. Call on EVERY rank at the same step (one real step runs as warm-up).
        Host-computed per-step delays (``delay_fn``) cannot be replayed and are rejected; device-sampled delays
        (``device_delay``) are part of the graph."""
        if self.delay_fn is not None:
            raise ValueError("CUDA-graph replay is incompatible with host-scheduled delay injection (use a device delay spec)")
        torch.cuda.synchronize()
        dev = self.device
        if self.is_worker:
            self.flags[W_STEP] = self.step - 1         # last completed step; wait_next advances it
        if self.is_master:
            self.flags[M_STEPW] = self.step
        self._graph_mode = True
        if self.is_worker and x_example is not None:
            self._gx = x_example.clone()
            self._gy = y_example.clone()
        else:
            self._gx = self._gy = None          # no batch: master-only ranks, and the flat parameter vector of the comm sweep
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):                  # one eager step in graph mode (also the warm-up torch asks for)
            self._first_graph_loss = self._step_body(self._gx, self._gy)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize()
        self.step += 1
        n0 = _lib.launch_count()
        self._graph = self._mgraph = None
        self._graph_loss = None
        if self.is_worker:
            self._graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._graph):
                self._graph_loss = self._worker_part(self._gx, self._gy)
            if hasattr(self.net, "batches_tracked"):
                self.net.batches_tracked -= 1          # the capture enqueued nothing: not a training step
        if self.is_master and self.master_mode == "free":
            self._mgraph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._mgraph, stream=self.master_stream):
                self.master_step()
        self.graph_kernel_launches = _lib.launch_count() - n0      # library kernels inside one replay
        self._graph_ready = True
        return self

    def local_slot(self, w: int) -> torch.Tensor:
        """Rank 0, unicast-slot mode: the master-side gradient slot of worker ``w`` as a tensor (tests, debugging)."""
        assert self.is_master and not self.nvls_gather, "slots live on the workers in NVLS gather mode"
        dt = torch.bfloat16 if self.grad_bf16 else torch.float32
        return self.arena.tensor(self.off_slots[w], self.net.total * self.slot_elem, dt)

    def _step_body(self, x, y):
        """Worker kernels first, master kernels second in HOST order (on the device they run side by side): the first
        launch of a kernel loads its module lazily and that load can wait for the context to drain, which never happens
        while a master kernel that spins on a flag is already resident ("Lazy kernel loading")."""
        loss = None
        if self.is_worker:
            loss = self._worker_part(x, y)
        if self.is_master and self.master_mode == "free":
            self.master_step()
        return loss

    def _worker_part(self, x, y):
        loss = self.worker_step(x, y)
        if self.is_master and self.master_mode == "gated":
            torch.cuda.current_stream(self.device).wait_stream(self.master_stream)     # joins the per-bucket forks (graph capture needs it)
        return loss

    def train_step(self, x_nchw_f32=None, labels=None):
        """One synchronous-SGD step of this rank's role(s); advances the step counter."""
        if self._graph_ready:
            if self.is_worker:
                if self._gx is not None:
                    self._gx.copy_(x_nchw_f32, non_blocking=True)
                    self._gy.copy_(labels, non_blocking=True)
                self._graph.replay()
                self.net.batches_tracked = getattr(self.net, "batches_tracked", 0) + 1      # a replay does not pass Network.train_step
            if self._mgraph is not None:
                with torch.cuda.stream(self.master_stream):
                    self._mgraph.replay()
            _lib.add_launch_count(self.graph_kernel_launches)
            self.step += 1
            return self._graph_loss
        loss = self._step_body(x_nchw_f32, labels)
        self.step += 1
        return loss

    # ---- the loop the roles (and the bench's end-to-end arm) run -------------------------------------------------
    def fit(self, train_loader, last_step: int, *, log_interval: int = 10, on_log=None, should_log=None, cuda_graph: bool = True,
            graph_after: int = 2, read_loss_every_step: bool = False):
        """Train from ``self.step`` through ``last_step`` (inclusive). Batches come from ``train_loader.next_batch(B)``
        (host tensors); they are staged through pinned memory on a side stream while the previous step computes
        (``PinnedBatchStager``), and after ``graph_after`` eager steps the whole step becomes one CUDA-graph launch.
        ``on_log(step, loss_value_or_None)`` runs every ``log_interval`` steps (or where ``should_log(step)`` says so) after
        a host synchronisation (that is what bounds how far the host runs ahead); ``read_loss_every_step`` reads the loss back every step instead.
        Master-only and listener ranks have no batches and pass ``train_loader=None``. Every rank must call this with the
        same ``last_step`` / ``graph_after`` (graph capture happens at the same step everywhere)."""
        from ..utils.data import PinnedBatchStager
        dev = self.device
        first = self.step
        use_graph = cuda_graph and self.delay_fn is None
        slot = None
        if self.is_worker:
            bx, by = train_loader.next_batch(self.batch_size)
            if getattr(self, "_stager", None) is None:
                self._stager = PinnedBatchStager(tuple(bx.shape), dev)
            stager = self._stager
            slot = stager.stage(bx, by)
        losses = []
        pending = None
        loss_bufs = getattr(self, "_loss_bufs", None)
        if loss_bufs is None and self.is_worker:
            loss_bufs = self._loss_bufs = [torch.zeros(1).pin_memory(), torch.zeros(1).pin_memory()]

        def collect(p):
            step_i, event, host_buf = p
            event.synchronize()
            losses.append((step_i, float(host_buf)))

        while self.step <= last_step:
            i = self.step
            x = y = None
            if self.is_worker:
                x, y = stager.acquire(slot)
            if use_graph and not self._graph_ready and i == first + graph_after:
                self.enable_cuda_graph(x, y)           # runs step i on this batch eagerly, then captures
                loss = self._first_graph_loss
            else:
                loss = self.train_step(x, y)
            if self.is_worker:
                stager.release(slot)
                if self.step <= last_step:
                    slot = stager.stage(*train_loader.next_batch(self.batch_size))      # overlaps with the step just enqueued
            want_log = (should_log(i) if should_log is not None else i % log_interval == 0) or i == first or i == last_step
            if read_loss_every_step or want_log:
                # Every step's loss is copied to pinned host memory right behind the step. It is READ one step late: the
                # host first hands step i+1 to the device and then waits for step i's copy, so the device never idles
                # for the launch latency of the next step. A log step (callback, checkpoint) is read at once.
                ev = None
                if loss is not None:
                    buf = loss_bufs[i & 1]
                    buf.copy_(loss, non_blocking=True)
                    ev = torch.cuda.Event()
                    ev.record(torch.cuda.current_stream(dev))
                if pending is not None:
                    collect(pending)
                    pending = None
                if want_log:
                    torch.cuda.current_stream(dev).synchronize()
                    val = None
                    if loss is not None:
                        val = float(buf)
                        losses.append((i, val))
                    if on_log is not None:
                        on_log(i, val)
                elif ev is not None:
                    pending = (i, ev, buf)
        if pending is not None:
            collect(pending)
        torch.cuda.synchronize()               # both streams: the master's last release is part of the run
        return losses

    def last_result(self):
        """(selected_mask, kill_mask, n_received, step) of the most recent master step (synchronises)."""
        self.master_stream.synchronize()
        r = [int(v) & 0xFFFFFFFF for v in self.result.cpu().tolist()]
        if r[3] == 0xDEAD:
            raise RuntimeError("parameter-server master timed out: fewer than k gradient sets arrived (a worker died or hung); "
                               f"arrived mask={r[0]:b}")
        nb = r[7]
        self.last_timing_us = {"wait_for_kth_arrival": r[4], "reduce_update_broadcast": r[5],
                               "bucket_ready_us": r[40: 40 + nb]}
        self.last_bucket_masks = r[8: 8 + nb]
        self.total_sets_applied = r[6]
        self.selected_counts = r[56: 56 + min(16, self.n_workers)]      # per worker: steps in which it was among the first k
        return tuple(r[:4])

    def master_state_dict(self):
        return self.net.state_dict()

    def close(self):
        torch.cuda.synchronize()
        self.arena.close()
