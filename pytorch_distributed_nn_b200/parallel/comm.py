"""Communicators: what the roles talk through.

The reference hands an ``MPI.COMM_WORLD`` object to its roles and lets models call ``communicator.Isend`` directly
(SURVEY §1 "layering violations"). Here roles receive a :class:`Communicator`:

* :class:`TorchDistCommunicator` -- message passing over ``torch.distributed`` (``gloo`` on CPU: the plumbing config
  of BASELINE.json; ``nccl`` only for the all-n baseline because NCCL has no tags and no any-order matching).
  It implements the reference's wire protocol (tags 10 / 11+l / 88+l / 77) with two fixes: every
  gradient and kill message carries the step it belongs to (stale messages are recognised and dropped, as the C++
  master does with ``tag == cur_step``), and workers can terminate.
* ``B200Communicator`` -- a handle for the device-resident backend (`parallel/b200_ps.py`): no messages exist
  there, the roles just enqueue kernels.

``Get_rank`` / ``Get_size`` keep the mpi4py spelling the reference uses.
"""
from __future__ import annotations

import os
from typing import List, Optional

import torch
import torch.distributed as dist

from .protocol import KILL_SIGNAL, TAG_GRAD_BASE, TAG_KILL, TAG_STEP

TERMINATE_STEP = -1


class Communicator:
    backend = "none"

    def Get_rank(self) -> int:
        raise NotImplementedError

    def Get_size(self) -> int:
        raise NotImplementedError


class B200Communicator(Communicator):
    """Rendezvous handle for the B200 backend: ranks, the (optional) process group, the local device."""
    backend = "b200"

    def __init__(self, rank=None, world_size=None, group=None, device=None):
        self.rank = int(os.environ.get("RANK", 0)) if rank is None else rank
        self.world = int(os.environ.get("WORLD_SIZE", 1)) if world_size is None else world_size
        self.group = group
        self.device = device

    def Get_rank(self):
        return self.rank

    def Get_size(self):
        return self.world


class TorchDistCommunicator(Communicator):
    """Point-to-point + broadcast message passing with the reference's tags."""

    def __init__(self, backend: str = "gloo", wire_dtype: torch.dtype = torch.float32, group=None, device="cpu"):
        assert dist.is_initialized(), "init_process_group first (see parallel.launch.init_distributed)"
        self.backend = backend
        self.group = group
        self.wire_dtype = wire_dtype
        self.device = torch.device(device)
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self._keep: List = []            # in-flight (request, tensor) pairs: buffers must outlive their isend
        self._kill_req = None
        self._kill_buf = None
        self.cur_step = 0                # worker side: step the rank is currently computing
        self.bytes_sent = 0

    def Get_rank(self):
        return self.rank

    def Get_size(self):
        return self.world

    # ---- small helpers ----------------------------------------------------------------------------------
    def _gc(self):
        self._keep = [(r, t) for (r, t) in self._keep if not r.is_completed()]

    def isend(self, t: torch.Tensor, dest: int, tag: int):
        req = dist.isend(t, dst=dest, tag=tag, group=self.group)
        self._keep.append((req, t))
        self.bytes_sent += t.numel() * t.element_size()
        if len(self._keep) > 4096:
            self._gc()
        return req

    def irecv(self, buf: torch.Tensor, src: int, tag: int):
        return dist.irecv(buf, src=src, tag=tag, group=self.group)

    def barrier(self):
        dist.barrier(group=self.group)

    # ---- step messages (tag 10) -------------------------------------------------------------------------
    def send_step(self, step: int, dest: int):
        return self.isend(torch.tensor([step], dtype=torch.int64), dest, TAG_STEP)

    def recv_step(self, src: int = 0) -> int:
        buf = torch.zeros(1, dtype=torch.int64)
        dist.recv(buf, src=src, tag=TAG_STEP, group=self.group)
        return int(buf)

    # ---- weights -----------------------------------------------------------------------------------------
    def to_wire(self, t: torch.Tensor) -> torch.Tensor:
        return t.detach().to(self.wire_dtype).contiguous().view(-1)

    def bcast(self, buf: torch.Tensor, root: int = 0):
        dist.broadcast(buf, src=root, group=self.group)

    # ---- gradients: header (tag 87: tensor index, step) + payload (tag 88 + l: the flat gradient) ----------
    # gloo offers neither MPI_Waitany nor completion polling on point-to-point requests, so the master runs one
    # receiver thread per worker: header first (fixed size), then the tensor on the reference's tag 88+l. A worker
    # that abandons a step simply stops sending; nothing is left half-matched.
    def send_grad(self, idx: int, grad: torch.Tensor):
        """Worker -> master, issued from inside the layer-split backward (reference lenet.py:119, tag 88+idx)."""
        self.isend(torch.tensor([idx, self.cur_step], dtype=torch.int64), 0, TAG_GRAD_BASE - 1)
        return self.isend(self.to_wire(grad), 0, TAG_GRAD_BASE + idx)

    def send_grad_eos(self):
        """End of stream: lets the master's receiver thread for this worker exit."""
        return self.isend(torch.tensor([-1, TERMINATE_STEP], dtype=torch.int64), 0, TAG_GRAD_BASE - 1)

    def start_grad_receivers(self, worker_ranks, buffers, out_queue):
        """Master side: ``buffers[l][w]`` receives tensor l of worker w; ``(w, l, step)`` goes into ``out_queue`` in
        arrival order (the Waitany order of the reference's hot loop)."""
        import threading
        self._rx_threads = []

        def loop(w, rank):
            hdr = torch.zeros(2, dtype=torch.int64)
            while True:
                dist.recv(hdr, src=rank, tag=TAG_GRAD_BASE - 1, group=self.group)
                l, step = int(hdr[0]), int(hdr[1])
                if l < 0:
                    out_queue.put((w, -1, TERMINATE_STEP))
                    return
                dist.recv(buffers[l][w], src=rank, tag=TAG_GRAD_BASE + l, group=self.group)
                out_queue.put((w, l, step, buffers[l][w].clone()))

        for w, rank in enumerate(worker_ranks):
            t = threading.Thread(target=loop, args=(w, rank), daemon=True)
            t.start()
            self._rx_threads.append(t)

    def join_grad_receivers(self, timeout=60.0):
        for t in getattr(self, "_rx_threads", []):
            t.join(timeout)

    # ---- kill signal (tag 77): payload = step being closed -----------------------------------------------
    def send_kill(self, dest: int, step: int):
        return self.isend(torch.tensor([KILL_SIGNAL, step], dtype=torch.int64), dest, TAG_KILL)

    def start_kill_listener(self):
        """Worker side: a thread blocks in recv(tag 77) and records the newest closed step
        (the reference polls ``Iprobe(0, 77)`` up to 10 000 times before every op, lenet.py:167-179)."""
        import threading
        self.closed_step = 0
        self.terminated = False

        def loop():
            buf = torch.zeros(2, dtype=torch.int64)
            while True:
                dist.recv(buf, src=0, tag=TAG_KILL, group=self.group)
                closed = int(buf[1])
                if closed == TERMINATE_STEP:
                    self.terminated = True
                    return
                self.closed_step = max(self.closed_step, closed)

        self._kill_thread = threading.Thread(target=loop, daemon=True)
        self._kill_thread.start()

    def should_abort(self, step: Optional[int] = None) -> bool:
        """Has the master closed ``step`` (default: the step this worker is computing)?"""
        step = self.cur_step if step is None else step
        return getattr(self, "closed_step", 0) >= step or getattr(self, "terminated", False)

    def wait_all(self, reqs, timeout_s: float = 120.0):
        for r in reqs:
            r.wait()
        self._keep = []
